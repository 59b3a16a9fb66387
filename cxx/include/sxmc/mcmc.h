// MCMC -- Markov-chain sampler of the likelihood space (reference: src/mcmc.h:36-138).  Same
// constructor and call operator; the chain runs device-resident inside libsxmc_b200
// (sxmc_chain_*), so the reference's protected nll() helper and launch-shape members are gone.
#ifndef SXMC_B200_MCMC_H
#define SXMC_B200_MCMC_H
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include <sxmc/likelihood.h>
#include <sxmc/observable.h>
#include <sxmc/signal.h>
#include <sxmc/source.h>
#include <sxmc/systematic.h>

struct sxmc_chain;
struct sxmc_batch;
struct sxmc_generator;

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
// The fits of an ensemble advanced together (no reference counterpart: src/sxmc.cpp:62-134 runs
// them one after the other): up to 148 chains over the same signals in ONE cooperative launch,
// each with its own data events and Philox stream.  Small fits only (the samples must fit in
// shared memory); Run() throws NotSmallFit otherwise and the caller falls back to MCMC.  Chain
// `slot` is the chain MCMC would produce for the same data and seed, row for row.
class MCMCBatch {
 public:
  struct NotSmallFit : public std::runtime_error {
    explicit NotSmallFit(const std::string& what) : std::runtime_error(what) {}
  };
  MCMCBatch(const std::vector<Source>& sources, const std::vector<Signal>& signals,
            const std::vector<Systematic>& systematics, const std::vector<Observable>& observables,
            int max_chains = 148);
  ~MCMCBatch();
  int max_chains() const { return nmax; }
  // make_fake_dataset(signals, systematics, observables, poisson) for this slot, drawn and kept on
  // the device (src/generator.cpp:10-48); observed[j] = events of signal j
  void GenerateData(int slot, bool poisson, uint64_t seed, uint64_t experiment, std::vector<unsigned>* observed = NULL);
  void SetData(int slot, const std::vector<float>& data);
  void SetSeed(int slot, uint64_t seed, uint64_t stream_id);
  void Run(int nchains, unsigned nsteps, float burnin_fraction, bool debug_mode = false);
  LikelihoodSpace* GetLikelihoodSpace(int slot);
  long accepted(int slot) const;

 private:
  int nmax;
  size_t nparameters, nsignals;
  std::vector<std::string> parameter_names;
  sxmc_batch* batch;
  sxmc_generator* gen;
};
#endif
