// Fit-level systematic parameter (reference: src/systematic.h:17-40).
#ifndef SXMC_B200_SYSTEMATIC_H
#define SXMC_B200_SYSTEMATIC_H
#include <cstddef>
#include <string>
#include <vector>

#include <sxmc/pdfz.h>

class Systematic {
 public:
  Systematic() : means(NULL), sigmas(NULL), observable_field_index(0), truth_field_index(0), npars(0),
                 type(pdfz::Systematic::SHIFT), fixed(false) {}
  virtual ~Systematic() {}
  std::string name, title, observable_field, truth_field;
  double* means;    // [npars] power-series coefficients' central values (borrowed)
  double* sigmas;   // [npars] their Gaussian widths, 0 = unconstrained (borrowed)
  size_t observable_field_index;
  size_t truth_field_index;
  size_t npars;
  std::vector<short> pidx;   // global coefficient indices into the pdfz parameter array (src/config.cpp:118-120)
  pdfz::Systematic::Type type;
  bool fixed;
};
#endif
