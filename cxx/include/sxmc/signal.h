// One signal of the fit: its PDF and expectation (reference: src/signal.h:24-112).  The
// file-reading constructor belongs to the config reader; build() is Signal::build_pdfz
// (src/signal.cpp:131-168) for callers that already hold the sample rows.
#ifndef SXMC_B200_SIGNAL_H
#define SXMC_B200_SIGNAL_H
#include <string>
#include <vector>

#include <sxmc/observable.h>
#include <sxmc/source.h>
#include <sxmc/systematic.h>

namespace pdfz { class Eval; }

class Signal {
 public:
  Signal() : dataset(0), nexpected(0), n_mc(0), histogram(NULL) {}
  // samples: rows of nfields floats.  Keeps the hemi::Array<short> index arrays alive in
  // `owned_pars` (the reference leaks them, src/signal.cpp:139).
  void build(const std::vector<float>& samples, int nfields, const std::vector<Observable>& observables,
             const std::vector<Systematic>& systematics);
  double get_efficiency(std::vector<Systematic>& systematics);        // src/signal.cpp:172-197
  std::string name, title, filename;
  unsigned dataset;
  Source source;
  double nexpected;
  size_t n_mc;
  pdfz::Eval* histogram;
  std::vector<std::string> systematic_names;
  std::vector<hemi::Array<short>*> owned_pars;
};
#endif
