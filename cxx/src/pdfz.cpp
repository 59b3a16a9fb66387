// pdfz::Eval / pdfz::EvalHist over the C ABI (reference bodies: src/pdfz.cpp:57-236, 441-620,
// 817-940).  Nothing is computed here: every method validates nothing itself either -- argument
// validation happens inside libsxmc_b200, which reports the reference's message with
// SXMC_EINVAL; that becomes `throw pdfz::Error(msg)` (by value, as the reference throws).
#include <sxmc/pdfz.h>

#include <cmath>
#include <stdexcept>

#include "sxmc_b200.h"

namespace pdfz {

static void check(int rc) {
  if (rc == SXMC_OK) return;
  if (rc == SXMC_EINVAL) throw Error(sxmc_last_error());
  throw std::runtime_error(sxmc_last_error());
}

static uint64_t g_random_seed = 0;
static uint64_t g_random_draws = 0;      // one Philox stream per RandomSample call

// ---- Eval ------------------------------------------------------------------------------

Eval::Eval(const std::vector<float>& /*samples*/, int _nfields, int _nobservables,
           const std::vector<double>& _lower, const std::vector<double>& _upper, unsigned _dataset)
    : nfields(_nfields), nobservables(_nobservables), lower(_lower), upper(_upper), dataset(_dataset),
      pdf_buffer(nullptr), pdf_offset(0), pdf_stride(1), norm_buffer(nullptr), norm_offset(0),
      param_buffer(nullptr), param_offset(0), param_stride(1), nsystematics(0), h_(nullptr) {
  // the checks of src/pdfz.cpp:64-82 run inside sxmc_hist_create (EvalHist ctor), which owns the samples
}

Eval::~Eval() {}

void Eval::SetPDFValueBuffer(hemi::Array<float>* output, int offset, int stride) {   // src/pdfz.cpp:105-110
  pdf_buffer = output;
  pdf_offset = offset;
  pdf_stride = stride;
}

void Eval::SetNormalizationBuffer(hemi::Array<unsigned int>* norm, int offset) {     // :112-116
  norm_buffer = norm;
  norm_offset = offset;
}

void Eval::SetParameterBuffer(hemi::Array<double>* params, int offset, int stride) { // :118-124
  param_buffer = params;
  param_offset = offset;
  param_stride = stride;
}

void Eval::AddSystematic(const Systematic& syst) {                                   // :126-174
  int obs = 0, extra = 0;
  hemi::Array<short>* pars = nullptr;
  int type = SXMC_SYST_SHIFT;
  if (const ShiftSystematic* s = dynamic_cast<const ShiftSystematic*>(&syst)) {
    type = SXMC_SYST_SHIFT; obs = s->obs; pars = s->pars;
  } else if (const ScaleSystematic* s = dynamic_cast<const ScaleSystematic*>(&syst)) {
    type = SXMC_SYST_SCALE; obs = s->obs; pars = s->pars;
  } else if (const CosThetaScaleSystematic* s = dynamic_cast<const CosThetaScaleSystematic*>(&syst)) {
    type = SXMC_SYST_CTSCALE; obs = s->obs; pars = s->pars;
  } else if (const ResolutionScaleSystematic* s = dynamic_cast<const ResolutionScaleSystematic*>(&syst)) {
    type = SXMC_SYST_RESOLUTION_SCALE; obs = s->obs; extra = s->true_obs; pars = s->pars;
  } else {
    throw Error("Unknown systematic type");                                          // :170
  }
  if (!pars) throw Error("Systematic without a parameter index array");
  check(sxmc_hist_add_systematic(h_, type, obs, extra, (int)pars->size(), pars->readOnlyHostPtr()));
  nsystematics += 1;
}

// ---- EvalHist --------------------------------------------------------------------------

EvalHist::EvalHist(const std::vector<float>& samples, int _nfields, int _nobservables,
                   const std::vector<double>& _lower, const std::vector<double>& _upper,
                   const std::vector<int>& _nbins, unsigned _dataset, bool /*optimize*/)
    : Eval(samples, _nfields, _nobservables, _lower, _upper, _dataset), nbins(_nbins), total_nbins(0) {
  check(sxmc_hist_create(samples.data(), samples.size(), _nfields, _nobservables, _lower.data(),
                         (int)_lower.size(), _upper.data(), (int)_upper.size(), _nbins.data(),
                         (int)_nbins.size(), _dataset, &h_));
  total_nbins = sxmc_hist_total_nbins(h_);
}

EvalHist::~EvalHist() {
  if (h_) sxmc_hist_destroy(h_);
}

void EvalHist::SetEvalPoints(const std::vector<float>& points) {                     // src/pdfz.cpp:245-302
  check(sxmc_hist_set_eval_points(h_, points.data(), points.size()));
}

// The reference autotunes launch shapes on the first call (src/pdfz.cpp:622-814); here the
// only first-use work is building the bin-sorted column layout, which Optimize() forces.
void EvalHist::Optimize() { check(sxmc_hist_prepare(h_)); }
void EvalHist::OptimizeBin() { check(sxmc_hist_prepare(h_)); }
void EvalHist::OptimizeEval() {}

void EvalHist::EvalAsync(bool do_eval_pdf) {                                          // :441-488
  if (!norm_buffer) throw Error("EvalHist::EvalAsync: no normalization buffer set");
  if (nsystematics > 0 && !param_buffer) throw Error("EvalHist::EvalAsync: no parameter buffer set");
  const double* params = param_buffer ? param_buffer->readOnlyDevicePtr() : nullptr;
  check(sxmc_hist_set_param_buffer(h_, params, param_offset, param_stride));
  check(sxmc_hist_set_norm_buffer(h_, norm_buffer->writeOnlyDevicePtr(), norm_offset));
  const bool lut = do_eval_pdf && pdf_buffer != nullptr;
  check(sxmc_hist_set_pdf_buffer(h_, lut ? pdf_buffer->writeOnlyDevicePtr() : nullptr, pdf_offset, pdf_stride));
  check(sxmc_hist_eval_async(h_, lut ? 1 : 0));
}

void EvalHist::EvalFinished() {                                                       // :491-495
  check(sxmc_hist_eval_finished(h_));
}

HistogramData EvalHist::CreateHistogram() {                                           // :497-592
  if (nobservables > 3) throw Error("Cannot EvalHist::CreateHistogram for dimensions greater than 3!");
  EvalAsync(false);
  EvalFinished();
  HistogramData out;
  out.nbins = nbins;
  out.lower = lower;
  out.upper = upper;
  out.norm = norm_buffer->readOnlyHostPtr()[norm_offset];
  std::vector<unsigned> counts((size_t)total_nbins);
  check(sxmc_hist_get_bins(h_, counts.data()));
  const double bin_volume = sxmc_hist_bin_volume(h_);
  out.contents.assign((size_t)total_nbins, 0.0);
  if (out.norm > 0)
    for (int i = 0; i < total_nbins; ++i) out.contents[i] = counts[i] / bin_volume / out.norm;
  return out;
}

HistogramData EvalHist::CreateHistogramProjection(int observable_index) {             // :596-620
  if (observable_index >= nobservables)
    throw Error("EvalHist::CreateHistogramProjection: Invalid observable index!");
  // the reference stops at "Not implemented" and returns NULL; here: the marginal density
  HistogramData full = CreateHistogram();
  HistogramData out;
  out.nbins.assign(1, nbins[observable_index]);
  out.lower.assign(1, lower[observable_index]);
  out.upper.assign(1, upper[observable_index]);
  out.norm = full.norm;
  out.contents.assign((size_t)nbins[observable_index], 0.0);
  std::vector<int> stride(nobservables, 1);
  for (int d = nobservables - 2; d >= 0; --d) stride[d] = stride[d + 1] * nbins[d + 1];
  double other_width = 1.0;
  for (int d = 0; d < nobservables; ++d)
    if (d != observable_index) other_width *= (upper[d] - lower[d]) / nbins[d];
  for (int i = 0; i < total_nbins; ++i)
    out.contents[(i / stride[observable_index]) % nbins[observable_index]] += full.contents[i] * other_width;
  return out;
}

HistogramData EvalHist::DefaultHistogram() {                                          // :926-940
  hemi::Array<unsigned> norms(1, true);
  hemi::Array<double> zeros(SXMC_MAX_PARAMETERS, true);
  zeros.writeOnlyHostPtr();
  hemi::Array<unsigned>* old_norm = norm_buffer;
  hemi::Array<double>* old_par = param_buffer;
  const int old_no = norm_offset, old_po = param_offset, old_ps = param_stride;
  SetNormalizationBuffer(&norms);
  SetParameterBuffer(&zeros);
  HistogramData out = CreateHistogram();
  // the reference leaves the object pointing at its dead stack temporaries (src/pdfz.cpp:928-936);
  // restoring the caller's buffers is the same contract minus the dangling pointers
  norm_buffer = old_norm; norm_offset = old_no;
  param_buffer = old_par; param_offset = old_po; param_stride = old_ps;
  return out;
}

void EvalHist::SetRandomSeed(uint64_t seed) {
  g_random_seed = seed;
  g_random_draws = 0;
}

int EvalHist::RandomSample(std::vector<float>& events, double nexpected, std::vector<double>& syst_vals,
                           std::vector<float>& uppers, std::vector<float>& lowers, bool poisson,
                           unsigned _dataset) {                                       // :817-922
  if (nobservables > 3) throw Error("Cannot EvalHist::CreateHistogram for dimensions greater than 3!");
  // The generator multiplies nexpected by the efficiency norm/n_mc (make_fake_dataset,
  // src/generator.cpp:30-39); RandomSample itself samples `nexpected` events, so n_mc is set to
  // the in-domain count: a first pass measures it.
  sxmc_hist* hs[1] = {h_};
  unsigned n_mc = (unsigned)sxmc_hist_nsamples(h_);
  sxmc_generator* g = nullptr;
  const int nlim = (int)uppers.size();
  check(sxmc_generator_create(hs, 1, &nexpected, &n_mc, syst_vals.data(), (int)syst_vals.size(),
                              nlim ? uppers.data() : nullptr, nlim ? lowers.data() : nullptr, nlim, &g));
  double eff = 0;
  int rc = sxmc_generator_efficiency(g, &eff);
  sxmc_generator_destroy(g);
  check(rc);
  const unsigned in_domain = (unsigned)std::llround(eff * n_mc);
  if (in_domain == 0) return 0;
  check(sxmc_generator_create(hs, 1, &nexpected, &in_domain, syst_vals.data(), (int)syst_vals.size(),
                              nlim ? uppers.data() : nullptr, nlim ? lowers.data() : nullptr, nlim, &g));
  unsigned observed = 0;
  size_t nfloats = 0;
  const uint64_t stream = g_random_draws++;
  rc = sxmc_generator_sample(g, poisson ? 1 : 0, g_random_seed, stream, nullptr, 0, &observed, &nfloats);
  if (rc == SXMC_OK && nfloats) {
    const size_t old = events.size();
    events.resize(old + nfloats);
    rc = sxmc_generator_sample(g, poisson ? 1 : 0, g_random_seed, stream, events.data() + old, nfloats,
                               &observed, &nfloats);
    for (size_t i = old + nobservables; i < events.size(); i += nobservables + 1) events[i] = (float)_dataset;
  }
  sxmc_generator_destroy(g);
  check(rc);
  return (int)observed;
}

int EvalHist::RandomSample(std::vector<float>& events, double nexpected, bool poisson, unsigned _dataset) {
  std::vector<double> syst_vals(SXMC_MAX_PARAMETERS, 0.0);       // src/pdfz.h:530-536: all systematics at zero
  std::vector<float> none_u, none_l;
  return RandomSample(events, nexpected, syst_vals, none_u, none_l, poisson, _dataset);
}

void EvalHist::GetSamples(std::vector<float>& sv) {                                   // src/pdfz.h:542-556
  const size_t nevents = sxmc_hist_nsamples(h_);
  std::vector<float> all(nevents * (size_t)nfields);
  check(sxmc_hist_get_samples(h_, all.data()));
  const size_t old = sv.size();
  const size_t ncols = (size_t)nobservables + 1;
  sv.resize(old + nevents * ncols);
  for (size_t i = 0; i < nevents; ++i) {
    for (int j = 0; j < nobservables; ++j) sv[old + i * ncols + j] = all[i * nfields + j];
    sv[old + i * ncols + nobservables] = (float)dataset;
  }
}

}  // namespace pdfz
