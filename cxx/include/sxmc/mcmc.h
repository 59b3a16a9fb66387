// MCMC -- Markov-chain sampler of the likelihood space (reference: src/mcmc.h:36-138).  Same
// constructor and call operator; the chain runs device-resident inside libsxmc_b200
// (sxmc_chain_*), so the reference's protected nll() helper and launch-shape members are gone.
#ifndef SXMC_B200_MCMC_H
#define SXMC_B200_MCMC_H
#include <cstdint>
#include <string>
#include <vector>

#include <sxmc/likelihood.h>
#include <sxmc/observable.h>
#include <sxmc/signal.h>
#include <sxmc/source.h>
#include <sxmc/systematic.h>

struct sxmc_chain;

class MCMC {
 public:
  MCMC(const std::vector<Source>& sources, const std::vector<Signal>& signals,
       const std::vector<Systematic>& systematics, const std::vector<Observable>& observables);
  ~MCMC();

  // data: rows of nobservables+1 floats (observables, dataset id), src/mcmc.cpp:231
  LikelihoodSpace* operator()(std::vector<float>& data, unsigned nsteps, float burnin_fraction,
                              const bool debug_mode = false, unsigned sync_interval = 10000);

  // Philox seed and stream (the reference seeds from gRandom->GetSeed(), src/mcmc.cpp:124-126)
  void SetSeed(uint64_t seed, uint64_t stream_id = 0);
  long accepted() const;
  sxmc_chain* handle() const { return chain; }

 private:
  size_t nsources, nsignals, nsystematics, nparameters, nobservables, nfloat;
  bool systematics_fixed;
  std::vector<std::string> parameter_names;
  sxmc_chain* chain;
};
#endif
