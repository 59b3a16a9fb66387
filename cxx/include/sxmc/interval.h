// reference: src/interval.h:11-28
#ifndef SXMC_B200_INTERVAL_H
#define SXMC_B200_INTERVAL_H
#include <string>

struct Interval {
  Interval() : one_sided(false), point_estimate(-1), lower(-1), upper(-1), cl(-1), coverage(-1) {}
  virtual ~Interval() {}
  std::string str() const;
  bool one_sided;
  float point_estimate, lower, upper, cl, coverage;
};
#endif
