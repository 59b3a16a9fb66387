// LikelihoodSpace -- the sink of a chain (reference: src/likelihood.h:25-100).  The TNtuple
// of the reference is a row-major float table here (names + rows); the Contour estimator
// (src/error_estimators/contour.cpp:18-69) runs as device reductions (sxmc_rows_contour); the
// Projection estimator (projection.cpp:14-77) is host code, as in the reference, with ROOT's
// auto-binned TH1F and Gaussian fit restated (cxx/src/projection.cpp).
#ifndef SXMC_B200_LIKELIHOOD_H
#define SXMC_B200_LIKELIHOOD_H
#include <map>
#include <string>
#include <vector>

#include <sxmc/interval.h>

typedef enum { ERROR_PROJECTION, ERROR_CONTOUR } ErrorType;   // src/error_estimator.h:13

struct NtupleRows {                     // stands in for TNtuple
  std::vector<std::string> names;       // parameter names, then "likelihood"
  std::vector<float> rows;              // [nrows][names.size()]
  long nrows() const { return names.empty() ? 0 : (long)(rows.size() / names.size()); }
};

namespace sxmc {
namespace errors {
// Projection::get_interval on one column: values[i * stride], i < n
Interval projection_interval(const float* values, long n, long stride, float cl, int nbins = 100);
void root_axis_limits(double xmin, double xmax, int nbins, double& lo, double& hi, int& nb);
}  // namespace errors
}  // namespace sxmc

class LikelihoodSpace {
 public:
  LikelihoodSpace(NtupleRows* samples, float _cl = 0.683, ErrorType _error_type = ERROR_CONTOUR);
  virtual ~LikelihoodSpace();
  std::map<std::string, Interval> get_best_fit() const { return ml_params; }
  std::map<std::string, Interval> get_best_fit(float& _nll) const { _nll = nll; return ml_params; }
  void print_best_fit() const;
  std::map<std::string, Interval> extract_best_fit(float& ml, float cl, ErrorType error_type = ERROR_CONTOUR);
  const NtupleRows* get_samples() { return samples; }

 private:
  std::map<std::string, Interval> ml_params;
  ErrorType error_type;
  float nll;
  float cl;
  NtupleRows* samples;   // owned (src/likelihood.h:30)
};
#endif
