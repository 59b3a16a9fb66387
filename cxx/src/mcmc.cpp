// Signal / LikelihoodSpace / MCMC over the C ABI (reference bodies: src/signal.cpp:111-197,
// src/likelihood.cpp:19-137, src/error_estimators/contour.cpp:18-69, src/mcmc.cpp:28-387).
#include <sxmc/mcmc.h>

#include <cmath>
#include <cstdio>
#include <iostream>
#include <sstream>
#include <stdexcept>

#include "sxmc_b200.h"

static void check(int rc) {
  if (rc == SXMC_OK) return;
  if (rc == SXMC_EINVAL) throw pdfz::Error(sxmc_last_error());
  throw std::runtime_error(sxmc_last_error());
}

// ---- Interval (src/interval.cpp) -------------------------------------------------------

std::string Interval::str() const {
  std::ostringstream o;
  float lo = point_estimate - lower, hi = upper - point_estimate;
  if (one_sided) o << "<" << upper << " (" << 100 * cl << "% CL)";
  else o << point_estimate << " -" << lo << " +" << hi;
  return o.str();
}

// ---- Signal ----------------------------------------------------------------------------

void Signal::build(const std::vector<float>& samples, int nfields, const std::vector<Observable>& observables,
                   const std::vector<Systematic>& systematics) {       // src/signal.cpp:111-168
  std::vector<double> lower(observables.size()), upper(observables.size());
  std::vector<int> nbins(observables.size());
  for (size_t i = 0; i < observables.size(); ++i) {
    lower[i] = observables[i].lower;
    upper[i] = observables[i].upper;
    nbins[i] = (int)observables[i].bins;
  }
  n_mc = nfields > 0 ? samples.size() / nfields : 0;
  histogram = new pdfz::EvalHist(samples, nfields, (int)observables.size(), lower, upper, nbins, dataset);
  for (size_t i = 0; i < systematics.size(); ++i) {
    const Systematic& y = systematics[i];
    hemi::Array<short>* pars = new hemi::Array<short>(y.npars, true);
    owned_pars.push_back(pars);
    for (size_t k = 0; k < y.pidx.size() && k < y.npars; ++k) pars->writeOnlyHostPtr()[k] = y.pidx[k];
    const int o = (int)y.observable_field_index, t = (int)y.truth_field_index;
    switch (y.type) {
      case pdfz::Systematic::SHIFT: histogram->AddSystematic(pdfz::ShiftSystematic(o, pars)); break;
      case pdfz::Systematic::SCALE: histogram->AddSystematic(pdfz::ScaleSystematic(o, pars)); break;
      case pdfz::Systematic::CTSCALE: histogram->AddSystematic(pdfz::CosThetaScaleSystematic(o, pars)); break;
      case pdfz::Systematic::RESOLUTION_SCALE:
        histogram->AddSystematic(pdfz::ResolutionScaleSystematic(o, t, pars)); break;
      default: throw pdfz::Error("Signal::build_pdfz: Unknown systematic ID");
    }
    systematic_names.push_back(y.name);
  }
}

double Signal::get_efficiency(std::vector<Systematic>& systematics) {  // src/signal.cpp:172-197
  size_t npars = 0;
  for (size_t i = 0; i < systematics.size(); ++i) npars += systematics[i].npars;
  hemi::Array<double> param_buffer(npars ? npars : 1, true);
  double* p = param_buffer.writeOnlyHostPtr();
  size_t k = 0;
  for (size_t i = 0; i < systematics.size(); ++i)
    for (size_t j = 0; j < systematics[i].npars; ++j) p[k++] = systematics[i].means[j];
  hemi::Array<unsigned> norms_buffer(1, true);
  norms_buffer.writeOnlyHostPtr();
  histogram->SetNormalizationBuffer(&norms_buffer);
  histogram->SetParameterBuffer(&param_buffer);
  histogram->EvalAsync(false);
  histogram->EvalFinished();
  const unsigned norm = norms_buffer.readOnlyHostPtr()[0];
  // unlike the reference, do not leave the PDF pointing at the dead stack buffers
  histogram->SetNormalizationBuffer(NULL);
  histogram->SetParameterBuffer(NULL);
  return n_mc ? 1.0 * norm / n_mc : 0.0;
}

// ---- LikelihoodSpace -------------------------------------------------------------------

// TMath::ChisquareQuantile(cl, 1) = z^2 with erf(z / sqrt 2) = cl; Newton on erf in double.
static double chisquare_quantile_1dof(double cl) {
  if (!(cl > 0)) return 0;
  if (cl >= 1) return INFINITY;
  double z = 1.0;
  for (int it = 0; it < 100; ++it) {
    const double f = std::erf(z * M_SQRT1_2) - cl;
    const double d = std::sqrt(2.0 / M_PI) * std::exp(-0.5 * z * z);
    const double step = f / d;
    z -= step;
    if (z <= 0) z = 1e-12;
    if (std::fabs(step) < 1e-15 * z) break;
  }
  return z * z;
}

LikelihoodSpace::LikelihoodSpace(NtupleRows* _samples, float _cl, ErrorType _error_type)
    : error_type(_error_type), nll(0), cl(_cl), samples(_samples) {    // src/likelihood.cpp:19-24
  ml_params = extract_best_fit(nll, cl, error_type);
}

LikelihoodSpace::~LikelihoodSpace() { delete samples; }

std::map<std::string, Interval>
LikelihoodSpace::extract_best_fit(float& lm, float _cl, ErrorType _error_type) {   // src/likelihood.cpp:104-137
  if (_error_type != ERROR_CONTOUR && _error_type != ERROR_PROJECTION) {
    std::cerr << "LikelihoodSpace::extract_best_fit: Unknown error type" << std::endl;
    throw(5);
  }
  std::map<std::string, Interval> best_fit;
  const long nrows = samples->nrows();
  const int ncols = (int)samples->names.size();
  if (nrows <= 0 || ncols < 2) { lm = 0; return best_fit; }
  const int np = ncols - 1;
  if (_error_type == ERROR_PROJECTION) {          // host estimator on the rows (projection.cpp:14-77)
    for (int i = 0; i < np; ++i)
      best_fit[samples->names[i]] = sxmc::errors::projection_interval(samples->rows.data() + i, nrows, ncols, _cl);
    float lmin = samples->rows[np];
    for (long r = 1; r < nrows; ++r) lmin = std::min(lmin, samples->rows[(size_t)r * ncols + np]);
    lm = lmin;                                    // samples->GetMinimum("likelihood")
    return best_fit;
  }
  void* dev = NULL;
  check(sxmc_buf_alloc(sizeof(float) * samples->rows.size(), &dev));
  int rc = sxmc_buf_upload(dev, samples->rows.data(), sizeof(float) * samples->rows.size());
  std::vector<float> lo(np), hi(np), pt(np);
  float lmin = 0;
  long ncontour = 0;
  const float delta = (float)(0.5 * chisquare_quantile_1dof(_cl));       // contour.cpp:20
  if (rc == SXMC_OK)
    rc = sxmc_rows_contour((const float*)dev, nrows, np, delta, lo.data(), hi.data(), pt.data(), &lmin, &ncontour);
  sxmc_buf_free(dev);
  check(rc);
  for (int i = 0; i < np; ++i) {                                          // contour.cpp:40-69
    Interval iv;
    iv.point_estimate = pt[i];
    iv.lower = lo[i];
    iv.upper = hi[i];
    iv.cl = _cl;
    iv.one_sided = false;
    iv.coverage = -1;
    best_fit[samples->names[i]] = iv;
  }
  lm = lmin;
  return best_fit;
}

void LikelihoodSpace::print_best_fit() const {                            // src/likelihood.cpp:32-40
  std::cout << "-- Best fit --" << std::endl;
  for (std::map<std::string, Interval>::const_iterator it = ml_params.begin(); it != ml_params.end(); ++it)
    std::cout << " " << it->first << ": " << it->second.str() << std::endl;
}

// ---- MCMC ------------------------------------------------------------------------------

// The parameter tables of MCMC::MCMC (src/mcmc.cpp:28-130) in the C ABI's form.
namespace {
struct ChainTables {
  size_t nsources, nsignals, nsystematics, nparameters, nobservables, nfloat;
  bool systematics_fixed;
  std::vector<double> means, sigma, nexpected;
  std::vector<int> fixed;
  std::vector<sxmc_hist*> pdfs;
  std::vector<unsigned> n_mc;
  std::vector<short> source_id;
  std::vector<std::string> names;
  sxmc_chain_config cfg;

  ChainTables(const std::vector<Source>& sources, const std::vector<Signal>& signals,
              const std::vector<Systematic>& systematics, const std::vector<Observable>& observables)
      : nsources(sources.size()), nsignals(signals.size()), nsystematics(systematics.size()), nparameters(0),
        nobservables(observables.size()), nfloat(0), systematics_fixed(true) {
    size_t npars = 0;
    for (size_t i = 0; i < systematics.size(); ++i) npars += systematics[i].npars;
    nparameters = nsources + npars;
    means.resize(nparameters);
    sigma.resize(nparameters);
    fixed.resize(nparameters);
    for (size_t i = 0; i < nsources; ++i) {
      means[i] = sources[i].mean;
      sigma[i] = sources[i].sigma;
      fixed[i] = sources[i].fixed ? 1 : 0;
      nfloat += sources[i].fixed ? 0 : 1;
      names.push_back(sources[i].name);
    }
    size_t k = nsources;
    for (size_t i = 0; i < systematics.size(); ++i) {
      if (!systematics[i].fixed) systematics_fixed = false;
      for (size_t j = 0; j < systematics[i].npars; ++j, ++k) {
        means[k] = systematics[i].means[j];
        sigma[k] = systematics[i].sigmas[j];
        fixed[k] = systematics[i].fixed ? 1 : 0;
        nfloat += systematics[i].fixed ? 0 : 1;
        std::ostringstream oss;
        oss << systematics[i].name << "_" << j;
        names.push_back(oss.str());
      }
    }
    names.push_back("likelihood");
    pdfs.resize(nsignals);
    nexpected.resize(nsignals);
    n_mc.resize(nsignals);
    source_id.resize(nsignals);
    for (size_t i = 0; i < nsignals; ++i) {
      pdfs[i] = signals[i].histogram ? signals[i].histogram->handle() : NULL;
      nexpected[i] = signals[i].nexpected;
      n_mc[i] = (unsigned)signals[i].n_mc;
      source_id[i] = (short)signals[i].source.index;
    }
    cfg.nsources = (int)nsources;
    cfg.nsignals = (int)nsignals;
    cfg.nparameters = (int)nparameters;
    cfg.nobservables = (int)nobservables;
    cfg.parameter_means = means.data();
    cfg.parameter_sigma = sigma.data();
    cfg.parameter_fixed = fixed.data();
    cfg.systematics_fixed = systematics_fixed ? 1 : 0;
    cfg.nsystematics = (int)nsystematics;
    cfg.pdfs = pdfs.data();
    cfg.nexpected = nexpected.data();
    cfg.n_mc = n_mc.data();
    cfg.source_id = source_id.data();
  }
};
}  // namespace

MCMC::MCMC(const std::vector<Source>& sources, const std::vector<Signal>& signals,
           const std::vector<Systematic>& systematics, const std::vector<Observable>& observables)
    : nsources(sources.size()), nsignals(signals.size()), nsystematics(systematics.size()), nparameters(0),
      nobservables(observables.size()), nfloat(0), systematics_fixed(true), chain(NULL) {   // src/mcmc.cpp:28-130
  ChainTables t(sources, signals, systematics, observables);
  nparameters = t.nparameters;
  nfloat = t.nfloat;
  systematics_fixed = t.systematics_fixed;
  parameter_names = t.names;
  if (systematics.size() > 0 && systematics_fixed)
    std::cout << "MCMC::MCMC: All systematics are fixed, will not reevaluate PDFs." << std::endl;
  check(sxmc_chain_create(&t.cfg, &chain));
}

MCMC::~MCMC() {
  if (chain) sxmc_chain_destroy(chain);
}

void MCMC::SetSeed(uint64_t seed, uint64_t stream_id) { check(sxmc_chain_set_seed(chain, seed, stream_id)); }

long MCMC::accepted() const {
  long a = 0;
  check(sxmc_chain_accepted(chain, &a));
  return a;
}

LikelihoodSpace* MCMC::operator()(std::vector<float>& data, unsigned nsteps, float burnin_fraction,
                                  const bool debug_mode, unsigned sync_interval) {   // src/mcmc.cpp:143-387
  long nrows = 0;
  check(sxmc_chain_run(chain, data.data(), data.size(), nsteps, burnin_fraction, debug_mode ? 1 : 0,
                       sync_interval, &nrows));
  NtupleRows* nt = new NtupleRows;
  nt->names = parameter_names;
  nt->rows.resize((size_t)nrows * (nparameters + 1));
  if (nrows > 0) check(sxmc_chain_read(chain, nt->rows.data(), nrows));
  long acc = 0;
  sxmc_chain_accepted(chain, &acc);
  float ms = 0;
  sxmc_chain_loop_ms(chain, &ms);
  std::cout << "MCMC: Elapsed time: " << ms * 1e-3f << "s, accepted " << acc << "/" << nsteps << " ("
            << (nsteps ? (float)acc / nsteps : 0.0f) << ")" << std::endl;   // src/mcmc.cpp:379-385
  return new LikelihoodSpace(nt);
}

// ---- MCMCBatch -------------------------------------------------------------------------

MCMCBatch::MCMCBatch(const std::vector<Source>& sources, const std::vector<Signal>& signals,
                     const std::vector<Systematic>& systematics, const std::vector<Observable>& observables,
                     int max_chains)
    : nmax(max_chains), nparameters(0), nsignals(signals.size()), batch(NULL), gen(NULL) {
  ChainTables t(sources, signals, systematics, observables);
  nparameters = t.nparameters;
  parameter_names = t.names;
  int rc = sxmc_batch_create(&t.cfg, max_chains, &batch);
  if (rc == SXMC_ENOTSMALL) throw NotSmallFit(sxmc_last_error());
  check(rc);
  // the fake-data generator of the ensemble: efficiencies and sampling histograms at the
  // systematics' means, observables' limits for the rejection test (src/generator.cpp:17-29)
  std::vector<double> syst_vals;
  for (size_t i = 0; i < systematics.size(); ++i)
    for (size_t j = 0; j < systematics[i].npars; ++j) syst_vals.push_back(systematics[i].means[j]);
  std::vector<float> upper, lower;
  for (size_t i = 0; i < observables.size(); ++i) { upper.push_back(observables[i].upper); lower.push_back(observables[i].lower); }
  rc = sxmc_generator_create(t.pdfs.data(), (int)t.pdfs.size(), t.nexpected.data(), t.n_mc.data(),
                             syst_vals.empty() ? NULL : syst_vals.data(), (int)syst_vals.size(), upper.data(),
                             lower.data(), (int)observables.size(), &gen);
  if (rc != SXMC_OK) { sxmc_batch_destroy(batch); batch = NULL; check(rc); }
}

MCMCBatch::~MCMCBatch() {
  if (gen) sxmc_generator_destroy(gen);
  if (batch) sxmc_batch_destroy(batch);
}

void MCMCBatch::GenerateData(int slot, bool poisson, uint64_t seed, uint64_t experiment, std::vector<unsigned>* observed) {
  std::vector<unsigned> obs(nsignals, 0);
  check(sxmc_batch_generate_data(batch, slot, gen, poisson ? 1 : 0, seed, experiment, obs.data()));
  if (observed) *observed = obs;
}

void MCMCBatch::SetData(int slot, const std::vector<float>& data) {
  check(sxmc_batch_set_data(batch, slot, data.data(), data.size()));
}

void MCMCBatch::SetSeed(int slot, uint64_t seed, uint64_t stream_id) {
  check(sxmc_batch_set_seed(batch, slot, seed, stream_id));
}

void MCMCBatch::Run(int nchains, unsigned nsteps, float burnin_fraction, bool debug_mode) {
  const int rc = sxmc_batch_run(batch, nchains, nsteps, burnin_fraction, debug_mode ? 1 : 0);
  if (rc == SXMC_ENOTSMALL) throw NotSmallFit(sxmc_last_error());
  check(rc);
}

long MCMCBatch::accepted(int slot) const {
  long n = 0, a = 0;
  check(sxmc_batch_nrows(batch, slot, &n, &a));
  return a;
}

LikelihoodSpace* MCMCBatch::GetLikelihoodSpace(int slot) {
  long nrows = 0, acc = 0;
  check(sxmc_batch_nrows(batch, slot, &nrows, &acc));
  NtupleRows* nt = new NtupleRows;
  nt->names = parameter_names;
  nt->rows.resize((size_t)nrows * (nparameters + 1));
  if (nrows > 0) check(sxmc_batch_read(batch, slot, nt->rows.data(), nrows));
  return new LikelihoodSpace(nt);
}
