// Source of correlated signal rates (reference: src/source.h:15-34; the Json::Value
// constructor lives with the config reader, not on the hot path).
#ifndef SXMC_B200_SOURCE_H
#define SXMC_B200_SOURCE_H
#include <cstddef>
#include <string>

struct Source {
  Source() : index(0), mean(1), sigma(0), fixed(false) {}
  Source(const std::string& _name, size_t _index, float _mean, float _sigma, bool _fixed)
      : name(_name), index(_index), mean(_mean), sigma(_sigma), fixed(_fixed) {}
  std::string name;
  size_t index;   // index in the list of sources
  float mean;     // mean expectation (scaling, 1.0 is nominal)
  float sigma;    // Gaussian constraint (fractional), 0 = unconstrained
  bool fixed;
};
#endif
