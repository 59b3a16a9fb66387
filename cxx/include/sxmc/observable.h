// Observable of the fit (reference: src/observable.h:16-35).
#ifndef SXMC_B200_OBSERVABLE_H
#define SXMC_B200_OBSERVABLE_H
#include <cstddef>
#include <string>
#include <vector>

struct Observable {
  Observable() : field_index(0), bins(0), lower(0), upper(0), logscale(false) {}
  std::string name, title, field, units;
  std::vector<float> yrange;
  size_t field_index;   // column of the sample rows
  size_t bins;
  float lower, upper;
  bool logscale;
};
#endif
