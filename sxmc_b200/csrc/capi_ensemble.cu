// capi_ensemble.cu -- the extern "C" layer of libsxmc_b200.so, part 3: what an ensemble of fake
// experiments needs around the chains (src/sxmc.cpp:62-134): the device fake-data generator
// (make_fake_dataset, src/generator.cpp:10-48), the Contour estimator as device reductions
// (src/error_estimators/contour.cpp:18-69) and the chain batch (up to 148 fits per cooperative launch).
#include "capi_internal.h"

using namespace sxmc;

extern "C" {

int sxmc_generator_destroy(sxmc_generator* g) {
  if (!g) return SXMC_OK;
  if (g->stream) cudaStreamSynchronize(g->stream);
  cudaFree(g->bins_dev); cudaFree(g->cum_dev); cudaFree(g->norms_dev); cudaFree(g->params_dev);
  cudaFree(g->sigs_dev); cudaFree(g->counts_dev); cudaFree(g->offsets_dev); cudaFree(g->events_dev);
  if (g->counts_host) cudaFreeHost(g->counts_host);
  g->batch.release();
  if (g->stream) cudaStreamDestroy(g->stream);
  delete g;
  return SXMC_OK;
}

int sxmc_generator_create(sxmc_hist* const* hs, int n, const double* nexpected, const unsigned* n_mc,
                          const double* syst_vals, int nsyst_vals, const float* uppers,
                          const float* lowers, int nlimits, sxmc_generator** out) {
  if (!hs || !out || n <= 0 || n > 256) SXMC_INVAL("sxmc_generator_create: 1..256 signal PDFs required");
  *out = nullptr;
  SXMC_TRY(ensure_init());
  const sxmc_hist* h0 = hs[0];
  if (h0->nobs > 3) SXMC_INVAL("Cannot EvalHist::CreateHistogram for dimensions greater than 3!");   // src/pdfz.cpp:497-499
  for (int j = 0; j < n; ++j) {
    const sxmc_hist* h = hs[j];
    bool same = h->nobs == h0->nobs && h->total_nbins == h0->total_nbins;
    for (int o = 0; same && o < h->nobs; ++o)
      same = h->lower[o] == h0->lower[o] && h->upper[o] == h0->upper[o] && h->nbins[o] == h0->nbins[o];
    if (!same) SXMC_INVAL("sxmc_generator_create: all signal PDFs must share the observables' binning");
  }
  if (nlimits != 0 && nlimits < h0->nobs) SXMC_INVAL("sxmc_generator_create: limits must cover every observable");
  sxmc_generator* g = new sxmc_generator;
  g->n = n; g->nobs = h0->nobs; g->total_nbins = h0->total_nbins;
  g->hs.assign(hs, hs + n);
  g->nexpected.assign(nexpected, nexpected + n);
  g->n_mc.assign(n_mc, n_mc + n);
  g->norms.resize(n);
  memset(&g->geo, 0, sizeof g->geo);
  g->geo.nobs = h0->nobs; g->geo.total_nbins = h0->total_nbins; g->geo.nlimits = nlimits ? h0->nobs : 0;
  for (int o = 0; o < h0->nobs; ++o) {
    g->geo.lower[o] = h0->lower[o];
    g->geo.width[o] = (h0->upper[o] - h0->lower[o]) / h0->nbins[o];
    g->geo.stride[o] = h0->stride[o];
    if (nlimits) { g->geo.lim_lo[o] = lowers[o]; g->geo.lim_hi[o] = uppers[o]; }
  }
  const size_t nb = (size_t)n * (size_t)std::max(1, g->total_nbins);
  cudaError_t e = cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(&g->bins_dev, sizeof(unsigned) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&g->cum_dev, sizeof(unsigned) * nb);
  if (e == cudaSuccess) e = cudaMalloc(&g->norms_dev, sizeof(unsigned) * n);
  if (e == cudaSuccess) e = cudaMalloc(&g->params_dev, sizeof(double) * std::max(1, nsyst_vals));
  if (e == cudaSuccess) e = cudaMalloc(&g->sigs_dev, sizeof(GenSignal) * n);
  if (e == cudaSuccess) e = cudaMalloc(&g->counts_dev, sizeof(unsigned long long) * (n + 1));
  if (e == cudaSuccess) e = cudaMalloc(&g->offsets_dev, sizeof(unsigned long long) * (n + 1));
  if (e == cudaSuccess) e = cudaMallocHost(&g->counts_host, sizeof(unsigned long long) * (n + 1));
  if (e == cudaSuccess) e = cudaMemsetAsync(g->bins_dev, 0, sizeof(unsigned) * nb, g->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(g->norms_dev, 0, sizeof(unsigned) * n, g->stream);
  if (e == cudaSuccess && nsyst_vals > 0)
    e = cudaMemcpyAsync(g->params_dev, syst_vals, sizeof(double) * nsyst_vals, cudaMemcpyHostToDevice, g->stream);
  if (e != cudaSuccess) { sxmc_generator_destroy(g); return cuda_fail(e, "sxmc_generator_create", __FILE__, __LINE__); }
  // one batched evaluation of every PDF at the systematics' means: the EvalAsync(false) of
  // get_efficiency (src/signal.cpp:192-194) and of CreateHistogram (src/pdfz.cpp:503-504)
  int rc = g->batch.reserve(n);
  for (int j = 0; j < n && !rc; ++j) {
    for (const SystDesc& d : hs[j]->systs)
      for (int k = 0; k < d.npars; ++k)
        if (d.pidx[k] >= nsyst_vals) { set_error("sxmc_generator_create: systematic coefficient index beyond syst_vals"); rc = SXMC_EINVAL; }
    if (!rc) rc = fill_job(hs[j], g->batch.host[j], g->bins_dev + (size_t)j * g->total_nbins, g->norms_dev + j,
                           g->params_dev, 1);
  }
  if (!rc) {
    g->batch.n_ep = 0;
    plan_launch(g->batch, n, g_tune);
    rc = upload_batch(g->batch, n, g->stream);
  }
  if (!rc) rc = launch_bin(g->batch.launch, g->stream);
  for (int j = 0; j < n && !rc; ++j)
    rc = launch_cum(g->bins_dev + (size_t)j * g->total_nbins, g->total_nbins, g->cum_dev + (size_t)j * g->total_nbins,
                    g->stream);
  if (!rc) {
    std::vector<GenSignal> sg(n);
    for (int j = 0; j < n; ++j) {
      sg[j].cum = g->cum_dev + (size_t)j * g->total_nbins;
      sg[j].norm = g->norms_dev + j;
      sg[j].nexpected = nexpected[j];
      sg[j].n_mc = n_mc[j];
      sg[j].dataset = hs[j]->dataset;
    }
    e = cudaMemcpyAsync(g->sigs_dev, sg.data(), sizeof(GenSignal) * n, cudaMemcpyHostToDevice, g->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(g->norms.data(), g->norms_dev, sizeof(unsigned) * n, cudaMemcpyDeviceToHost, g->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g->stream);
    if (e != cudaSuccess) rc = cuda_fail(e, "sxmc_generator_create", __FILE__, __LINE__);
  }
  if (rc) { sxmc_generator_destroy(g); return rc; }
  *out = g;
  return SXMC_OK;
}

int sxmc_generator_efficiency(sxmc_generator* g, double* eff_out) {
  if (!g || !eff_out) SXMC_INVAL("sxmc_generator_efficiency: NULL argument");
  for (int j = 0; j < g->n; ++j) eff_out[j] = 1.0 * g->norms[j] / (double)g->n_mc[j];   // src/signal.cpp:196
  return SXMC_OK;
}

// Draw one dataset into the generator's device buffer.  *nevents_out = rows generated.
static int generator_draw(sxmc_generator* g, int poisson, uint64_t seed, uint64_t stream_id,
                          unsigned long long* nevents_out) {
  // capacity: mean + 10 sigma + slack of the total expectation; regrown (and redrawn) if short
  double mean_total = 0;
  for (int j = 0; j < g->n; ++j) mean_total += g->nexpected[j] * (1.0 * g->norms[j] / (double)g->n_mc[j]);
  unsigned long long want = (unsigned long long)(mean_total + 10.0 * std::sqrt(std::max(mean_total, 1.0)) + 64.0);
  for (int attempt = 0; attempt < 2; ++attempt) {
    if (g->cap_events < want) {
      cudaFree(g->events_dev);
      g->events_dev = nullptr;
      g->cap_events = 0;
      SXMC_CUDA(cudaMalloc(&g->events_dev, sizeof(float) * want * (size_t)(g->nobs + 1)));
      g->cap_events = want;
    }
    SXMC_TRY(launch_generate(g->sigs_dev, g->n, g->geo, poisson, seed, stream_id, g->counts_dev, g->offsets_dev,
                             g->events_dev, g->cap_events, g->stream));
    SXMC_CUDA(cudaMemcpyAsync(g->counts_host, g->counts_dev, sizeof(unsigned long long) * (g->n + 1),
                              cudaMemcpyDeviceToHost, g->stream));
    SXMC_CUDA(cudaStreamSynchronize(g->stream));
    if (g->counts_host[g->n] <= g->cap_events) break;
    want = g->counts_host[g->n];      // counter-based RNG: the redraw yields the same counts
  }
  g->last_total = g->counts_host[g->n];
  *nevents_out = g->last_total;
  return SXMC_OK;
}

int sxmc_generator_sample(sxmc_generator* g, int poisson, uint64_t seed, uint64_t stream_id,
                          float* events_out, size_t max_floats, unsigned* observed_out,
                          size_t* nfloats_out) {
  if (!g || !nfloats_out) SXMC_INVAL("sxmc_generator_sample: NULL argument");
  unsigned long long ne = 0;
  SXMC_TRY(generator_draw(g, poisson, seed, stream_id, &ne));
  const size_t nf = (size_t)ne * (size_t)(g->nobs + 1);
  *nfloats_out = nf;
  if (observed_out) for (int j = 0; j < g->n; ++j) observed_out[j] = (unsigned)g->counts_host[j];
  if (events_out) {
    if (nf > max_floats) { set_error("sxmc_generator_sample: events_out too small (see *nfloats_out)"); return SXMC_EINVAL; }
    if (nf) SXMC_CUDA(cudaMemcpy(events_out, g->events_dev, sizeof(float) * nf, cudaMemcpyDeviceToHost));
  }
  return SXMC_OK;
}

// ---- Contour estimator ------------------------------------------------------------------------
int sxmc_rows_contour(const float* rows_dev, long nrows, int nparameters, float delta,
                      float* lower_out, float* upper_out, float* point_out, float* lmin_out,
                      long* ncontour_out) {
  if (!rows_dev || nrows <= 0 || nparameters <= 0) SXMC_INVAL("sxmc_rows_contour: no rows");
  SXMC_TRY(ensure_init());
  const int np = nparameters;
  const int NL = CONTOUR_MAX_LEVELS;
  const size_t words = 1 + 2 * NL + 2 * (size_t)NL * np;   // lmin | counts (u64 = 2 words) | mins | maxs
  unsigned* buf = nullptr;
  SXMC_CUDA(cudaMalloc(&buf, sizeof(unsigned) * (words + 1)));
  std::vector<unsigned> init(words + 1, 0u);
  // layout (words): [0] lmin, [1] pad, [2 .. 2+2NL) counts, then mins, then maxs
  const size_t off_counts = 2, off_min = 2 + 2 * (size_t)NL, off_max = off_min + (size_t)NL * np;
  init[0] = 0x7f800000u;
  for (size_t i = 0; i < (size_t)NL * np; ++i) { init[off_min + i] = 0x7f800000u; init[off_max + i] = 0xff800000u; }
  cudaError_t e = cudaMemcpy(buf, init.data(), sizeof(unsigned) * (words + 1), cudaMemcpyHostToDevice);
  int rc = SXMC_OK;
  float lmin = 0;
  if (e == cudaSuccess) rc = launch_rows_min_nll(rows_dev, nrows, np, (float*)buf, 0);
  if (e == cudaSuccess && !rc) e = cudaMemcpy(&lmin, buf, sizeof(float), cudaMemcpyDeviceToHost);
  // "likelihood+" << -lmin << "<" << delta : both floats pass through an ostream (6 significant
  // digits) and are parsed back as doubles by TTreeFormula (src/likelihood.cpp:91-92)
  auto through_ostream = [](float v) { char t[64]; snprintf(t, sizeof t, "%g", (double)v); return strtod(t, nullptr); };
  const double off = through_ostream(-lmin);
  ContourLevels lv;
  lv.n = NL;
  lv.thr[0] = through_ostream(delta);                         // the contour itself
  float dnll = 0.13f;                                         // src/error_estimators/contour.cpp:43-54
  for (int l = 1; l < NL; ++l) { lv.thr[l] = std::min(through_ostream(dnll), lv.thr[0]); dnll *= 5; }
  // (the point-estimate selection runs on the contour points, so its threshold never exceeds delta's)
  if (e == cudaSuccess && !rc)
    rc = launch_rows_extents(rows_dev, nrows, np, off, lv, (unsigned long long*)(buf + off_counts),
                             (float*)(buf + off_min), (float*)(buf + off_max), 0);
  std::vector<unsigned> res(words + 1);
  if (e == cudaSuccess && !rc) e = cudaMemcpy(res.data(), buf, sizeof(unsigned) * (words + 1), cudaMemcpyDeviceToHost);
  cudaFree(buf);
  if (e != cudaSuccess) return cuda_fail(e, "sxmc_rows_contour", __FILE__, __LINE__);
  if (rc) return rc;
  const unsigned long long* counts = (const unsigned long long*)(res.data() + off_counts);
  const float* mins = (const float*)(res.data() + off_min);
  const float* maxs = (const float*)(res.data() + off_max);
  if (counts[0] == 0) { set_error("sxmc_rows_contour: no contour points"); return SXMC_ESTATE; }   // assert at likelihood.cpp:97
  int lp = -1;
  for (int l = 1; l < NL; ++l) if (counts[l] >= 1) { lp = l; break; }
  if (lp < 0) lp = 0;
  for (int k = 0; k < np; ++k) {
    if (lower_out) lower_out[k] = mins[k];
    if (upper_out) upper_out[k] = maxs[k];
    // (GetMinimum + GetMaximum) / 2 on doubles, stored to a float member
    if (point_out) point_out[k] = (float)(((double)mins[(size_t)lp * np + k] + (double)maxs[(size_t)lp * np + k]) / 2);
  }
  if (lmin_out) *lmin_out = lmin;
  if (ncontour_out) *ncontour_out = (long)counts[0];
  return SXMC_OK;
}

int sxmc_chain_contour(sxmc_chain* c, float delta, float* lower_out, float* upper_out, float* point_out,
                       float* lmin_out, long* ncontour_out) {
  if (!c) SXMC_INVAL("sxmc_chain_contour: NULL chain");
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  const long n = (long)std::min(c->nrows, c->row_cap);
  return sxmc_rows_contour(c->rows_dev, n, c->nparameters, delta, lower_out, upper_out, point_out, lmin_out,
                           ncontour_out);
}

}  // extern "C"

// =========================================================================================
// Chain batch: many chains over the same PDFs in one cooperative launch (batch_kernels.cu)
// =========================================================================================
struct BatchSlot {
  uint64_t seed = 0, stream_id = 0;
  int* read_bins = nullptr; size_t rb_cap = 0;       // [nrb][rb_cap]
  size_t nevents = 0;
  float* rows = nullptr; long long row_cap = 0;
  double* tape_n = nullptr; size_t n_tape_n = 0;
  double* tape_u = nullptr; size_t n_tape_u = 0;
  double* nll_trace = nullptr; double* prop_trace = nullptr; unsigned trace_steps = 0;
  long long nrows = 0, accepted = 0;
};

struct sxmc_batch {
  Tuning tune;                          // kernel selection knobs, as they stood when the batch was created
  int nsources = 0, nsignals = 0, nparameters = 0, nobs = 0, nsystematics = 0;
  bool systematics_fixed = true;
  std::vector<double> means, sigmas, nexpected;
  std::vector<int> fixed;
  std::vector<unsigned> n_mc;
  std::vector<short> source_id;
  std::vector<sxmc_hist*> pdfs;
  int total_nbins = 0, max_chains = 0;
  bool shared_rb = true;            // every signal reads the same dataset: one read_bins array per slot
  size_t hist_words = 0;
  cudaStream_t stream = nullptr;
  ChainState* st_dev = nullptr;
  ChainState* st_host = nullptr;    // pinned [max_chains]
  unsigned* hist_dev = nullptr;
  NllSignal* sigs_dev = nullptr;
  long long* nevents_dev = nullptr;
  unsigned* sync_dev = nullptr;
  unsigned* sync_host = nullptr;
  float* points_dev = nullptr; size_t points_cap = 0;
  std::vector<BatchSlot> slots;
  Batch jobs;
  bool trace = false;
  float loop_ms = 0;
  int last_n = 0;
  unsigned long long* stamps_dev = nullptr;   // profiling hook
  bool mono = false;                // 1-D monotone program on sorted, NaN-free samples
  unsigned* mult_dev = nullptr;     // [max_chains][total_nbins] events per bin, when the grouped lookup applies
  MonoPlan* mono_plan = nullptr;    // built on the first run (host partition of the sorted samples)
  bool mono_plan_failed = false;
  int last_path = 0;                // 1 = monotone kernel, 2 = sample-by-sample kernel
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

extern "C" {

int sxmc_batch_destroy(sxmc_batch* b) {
  if (!b) return SXMC_OK;
  if (b->stream) cudaStreamSynchronize(b->stream);
  for (BatchSlot& s : b->slots) {
    cudaFree(s.read_bins); cudaFree(s.rows); cudaFree(s.tape_n); cudaFree(s.tape_u);
    cudaFree(s.nll_trace); cudaFree(s.prop_trace);
  }
  cudaFree(b->st_dev); cudaFree(b->hist_dev); cudaFree(b->sigs_dev); cudaFree(b->nevents_dev);
  cudaFree(b->sync_dev); cudaFree(b->points_dev); cudaFree(b->stamps_dev); cudaFree(b->mult_dev);
  if (b->st_host) cudaFreeHost(b->st_host);
  if (b->sync_host) cudaFreeHost(b->sync_host);
  b->jobs.release();
  delete b->mono_plan;
  if (b->ev0) cudaEventDestroy(b->ev0);
  if (b->ev1) cudaEventDestroy(b->ev1);
  if (b->stream) cudaStreamDestroy(b->stream);
  delete b;
  return SXMC_OK;
}

int sxmc_batch_create(const sxmc_chain_config* cfg, int max_chains, sxmc_batch** out) {
  if (!cfg || !out) SXMC_INVAL("sxmc_batch_create: NULL argument");
  *out = nullptr;
  if (max_chains < 1 || max_chains > BATCH_MAX_CHAINS) SXMC_INVAL("sxmc_batch_create: 1..148 chains per batch");
  if (cfg->nparameters <= 0 || cfg->nparameters > SXMC_MAX_PARAMETERS)
    SXMC_INVAL("sxmc_batch_create: nparameters out of range (SXMC_MAX_PARAMETERS)");
  if (cfg->nsignals <= 0 || cfg->nsignals > 128) SXMC_INVAL("sxmc_batch_create: nsignals out of range (1..128)");
  if (cfg->nsources <= 0 || cfg->nsources > cfg->nparameters) SXMC_INVAL("sxmc_batch_create: nsources out of range");
  SXMC_TRY(ensure_init());
  sxmc_batch* b = new sxmc_batch;
  b->tune = g_tune;
  b->nsources = cfg->nsources; b->nsignals = cfg->nsignals; b->nparameters = cfg->nparameters;
  b->nobs = cfg->nobservables; b->nsystematics = cfg->nsystematics;
  b->systematics_fixed = cfg->systematics_fixed != 0;
  b->means.assign(cfg->parameter_means, cfg->parameter_means + cfg->nparameters);
  b->sigmas.assign(cfg->parameter_sigma, cfg->parameter_sigma + cfg->nparameters);
  b->fixed.assign(cfg->parameter_fixed, cfg->parameter_fixed + cfg->nparameters);
  b->pdfs.assign(cfg->pdfs, cfg->pdfs + cfg->nsignals);
  b->nexpected.assign(cfg->nexpected, cfg->nexpected + cfg->nsignals);
  b->n_mc.assign(cfg->n_mc, cfg->n_mc + cfg->nsignals);
  b->source_id.assign(cfg->source_id, cfg->source_id + cfg->nsignals);
  b->total_nbins = b->pdfs[0]->total_nbins;
  b->max_chains = max_chains;
  for (int j = 0; j < b->nsignals; ++j) {
    if (b->pdfs[j]->total_nbins != b->total_nbins || b->pdfs[j]->nobs != b->nobs) {
      delete b;
      SXMC_INVAL("sxmc_batch_create: all signal PDFs must share the observables' binning");
    }
    if (b->source_id[j] < 0 || b->source_id[j] >= b->nsources) { delete b; SXMC_INVAL("sxmc_batch_create: source_id out of range"); }
    if (b->pdfs[j]->dataset != b->pdfs[0]->dataset) b->shared_rb = false;
  }
  b->hist_words = (size_t)b->nsignals * ((size_t)b->total_nbins + 1);
  b->slots.resize(max_chains);
  cudaError_t e = cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(&b->st_dev, sizeof(ChainState) * max_chains);
  if (e == cudaSuccess) e = cudaMallocHost(&b->st_host, sizeof(ChainState) * max_chains);
  if (e == cudaSuccess) e = cudaMalloc(&b->hist_dev, sizeof(unsigned) * b->hist_words * max_chains);
  if (e == cudaSuccess) e = cudaMalloc(&b->sigs_dev, sizeof(NllSignal) * b->nsignals * max_chains);
  if (e == cudaSuccess) e = cudaMalloc(&b->nevents_dev, sizeof(long long) * max_chains);
  if (e == cudaSuccess) e = cudaMalloc(&b->sync_dev, sizeof(unsigned) * 4);
  if (e == cudaSuccess) e = cudaMallocHost(&b->sync_host, sizeof(unsigned) * 4);
  if (e == cudaSuccess) e = cudaEventCreate(&b->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&b->ev1);
  // grouped lookup: one dataset for all signals and a (signal, bin) table that fits on chip
  if (e == cudaSuccess && b->shared_rb && (size_t)b->nsignals * b->total_nbins <= (size_t)BATCH_PDF_TABLE)
    e = cudaMalloc(&b->mult_dev, sizeof(unsigned) * (size_t)b->total_nbins * max_chains);
  if (e != cudaSuccess) { sxmc_batch_destroy(b); return cuda_fail(e, "sxmc_batch_create", __FILE__, __LINE__); }
  // the binning jobs (columns, systematic programs); outputs are addressed by the kernel itself
  int rc = b->jobs.reserve(b->nsignals);
  for (int j = 0; j < b->nsignals && !rc; ++j)
    rc = fill_job(b->pdfs[j], b->jobs.host[j], nullptr, nullptr, nullptr, 1);
  if (!rc) { b->jobs.n_ep = 0; plan_launch(b->jobs, b->nsignals, b->tune); }
  // monotone endpoint path (batch_kernels.cu): one observable in one column, every systematic a
  // SHIFT / SCALE / CTSCALE of it, samples sorted by it, no NaN / -inf among them
  if (!rc) {
    const BinLaunch& L = b->jobs.launch;
    bool mono = L.fast && L.nobs == 1 && L.nslots == 1;
    for (int j = 0; j < b->nsignals && mono && !rc; ++j) {
      sxmc_hist* h = b->pdfs[j];
      if (!h->sort || !h->prepared || h->primary != 0) mono = false;
      for (const SystDesc& d : h->systs)
        if (d.type == SXMC_SYST_RESOLUTION_SCALE || d.obs != 0 || d.npars > 1) mono = false;
      if (mono) {
        bool bad = false;
        rc = column_has_nan_or_neginf(h->cols_dev, h->col_pitch, &bad, b->stream);
        if (bad) mono = false;
      }
    }
    b->mono = mono && !rc;
  }
  if (rc) { sxmc_batch_destroy(b); return rc; }
  *out = b;
  return SXMC_OK;
}

static int batch_slot(sxmc_batch* b, int slot, BatchSlot** out) {
  if (!b || slot < 0 || slot >= b->max_chains) SXMC_INVAL("sxmc_batch: slot out of range");
  *out = &b->slots[slot];
  return SXMC_OK;
}

int sxmc_batch_set_seed(sxmc_batch* b, int slot, uint64_t seed, uint64_t stream_id) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  s->seed = seed; s->stream_id = stream_id;
  return SXMC_OK;
}

int sxmc_batch_set_tapes(sxmc_batch* b, int slot, const double* normals, size_t n_normals,
                         const double* uniforms, size_t n_uniforms) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  cudaFree(s->tape_n); cudaFree(s->tape_u);
  s->tape_n = nullptr; s->tape_u = nullptr; s->n_tape_n = 0; s->n_tape_u = 0;
  if (normals && uniforms) {
    SXMC_CUDA(cudaMalloc(&s->tape_n, sizeof(double) * std::max<size_t>(1, n_normals)));
    SXMC_CUDA(cudaMalloc(&s->tape_u, sizeof(double) * std::max<size_t>(1, n_uniforms)));
    SXMC_CUDA(cudaMemcpy(s->tape_n, normals, sizeof(double) * n_normals, cudaMemcpyHostToDevice));
    SXMC_CUDA(cudaMemcpy(s->tape_u, uniforms, sizeof(double) * n_uniforms, cudaMemcpyHostToDevice));
    s->n_tape_n = n_normals; s->n_tape_u = n_uniforms;
  }
  return SXMC_OK;
}

// SetEvalPoints of every PDF for one slot, from DEVICE rows (src/pdfz.cpp:245-302)
static int batch_slot_points(sxmc_batch* b, BatchSlot* s, const float* points_dev, size_t nevents) {
  const int nrb = b->shared_rb ? 1 : b->nsignals;
  if (s->rb_cap < nevents || !s->read_bins) {
    cudaFree(s->read_bins);
    s->read_bins = nullptr;
    s->rb_cap = std::max<size_t>(nevents + nevents / 8 + 64, 256);
    SXMC_CUDA(cudaMalloc(&s->read_bins, sizeof(int) * s->rb_cap * nrb));
  }
  s->nevents = nevents;
  for (int j = 0; j < nrb; ++j) {
    PrepGeo g;
    fill_prep_geo(b->pdfs[j], g);
    SXMC_TRY(launch_read_bins(points_dev, (long long)nevents, b->nobs, g, b->pdfs[j]->dataset,
                              s->read_bins + (size_t)j * s->rb_cap, b->stream));
  }
  if (b->mult_dev)
    SXMC_TRY(launch_event_mult(s->read_bins, (long long)nevents, b->total_nbins,
                               b->mult_dev + (size_t)(s - b->slots.data()) * b->total_nbins, b->stream));
  return SXMC_OK;
}

int sxmc_batch_set_data(sxmc_batch* b, int slot, const float* data, size_t data_nfloats) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  if (data_nfloats % (size_t)(b->nobs + 1) != 0)
    SXMC_INVAL("Number of entries in evaluation points array not divisible by number of observables.");
  if (b->points_cap < data_nfloats) {
    SXMC_CUDA(cudaStreamSynchronize(b->stream));
    cudaFree(b->points_dev);
    b->points_dev = nullptr;
    b->points_cap = data_nfloats + data_nfloats / 4 + 1024;
    SXMC_CUDA(cudaMalloc(&b->points_dev, sizeof(float) * b->points_cap));
  }
  if (data_nfloats)
    SXMC_CUDA(cudaMemcpyAsync(b->points_dev, data, sizeof(float) * data_nfloats, cudaMemcpyHostToDevice, b->stream));
  SXMC_TRY(batch_slot_points(b, s, b->points_dev, data_nfloats / (size_t)(b->nobs + 1)));
  SXMC_CUDA(cudaStreamSynchronize(b->stream));   // points_dev is reused by the next slot
  return SXMC_OK;
}

int sxmc_batch_generate_data(sxmc_batch* b, int slot, sxmc_generator* g, int poisson, uint64_t seed,
                             uint64_t stream_id, unsigned* observed_out) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  if (!g || g->nobs != b->nobs) SXMC_INVAL("sxmc_batch_generate_data: generator does not match the fit");
  unsigned long long ne = 0;
  SXMC_TRY(generator_draw(g, poisson, seed, stream_id, &ne));   // synchronises the generator's stream
  if (observed_out) for (int j = 0; j < g->n; ++j) observed_out[j] = (unsigned)g->counts_host[j];
  SXMC_TRY(batch_slot_points(b, s, g->events_dev, (size_t)ne));
  SXMC_CUDA(cudaStreamSynchronize(b->stream));   // the generator's buffer is overwritten by the next draw
  return SXMC_OK;
}

int sxmc_batch_enable_trace(sxmc_batch* b, int on) { b->trace = on != 0; return SXMC_OK; }

// MCMC::operator() (src/mcmc.cpp:143-387) for slots [0, nchains) at once.
int sxmc_batch_run(sxmc_batch* b, int nchains, unsigned nsteps, float burnin_fraction, int debug_mode) {
  if (!b || nchains < 1 || nchains > b->max_chains) SXMC_INVAL("sxmc_batch_run: nchains out of range");
  const int np = b->nparameters, ns = b->nsignals;
  // monotone fits: edge search over sorted samples (batch_mono_kernels.cu); others: the
  // sample-by-sample batch kernel (batch_kernels.cu)
  bool use_mono = b->mono && !b->tune.no_mono && !b->mono_plan_failed;
  if (use_mono && !b->mono_plan) {
    std::vector<unsigned long long> nsamp(ns);
    for (int j = 0; j < ns; ++j) nsamp[j] = b->pdfs[j]->nsamples;
    b->mono_plan = new MonoPlan;
    if (plan_chain_batch_mono(b->jobs.launch, nsamp.data(), b->max_chains, b->mono_plan) != SXMC_OK) {
      delete b->mono_plan;
      b->mono_plan = nullptr;
      b->mono_plan_failed = true;
      use_mono = false;
    }
  }
  BatchPlan plan;
  if (!use_mono) SXMC_TRY(plan_chain_batch(b->jobs.launch, nchains, &plan));
  std::vector<NllSignal> sg((size_t)nchains * ns);
  std::vector<long long> nev(nchains);
  for (int e = 0; e < nchains; ++e) {
    BatchSlot& s = b->slots[e];
    if (!s.read_bins) { set_error("sxmc_batch_run: a slot has no data (sxmc_batch_set_data)"); return SXMC_ESTATE; }
    if (s.row_cap < (long long)nsteps || !s.rows) {
      cudaFree(s.rows);
      s.rows = nullptr;
      s.row_cap = std::max<long long>(1, nsteps);
      SXMC_CUDA(cudaMalloc(&s.rows, sizeof(float) * (size_t)s.row_cap * (np + 1)));
    }
    if (b->trace && s.trace_steps < nsteps) {
      cudaFree(s.nll_trace); cudaFree(s.prop_trace);
      s.nll_trace = nullptr; s.prop_trace = nullptr;
      SXMC_CUDA(cudaMalloc(&s.nll_trace, sizeof(double) * std::max(1u, nsteps)));
      SXMC_CUDA(cudaMalloc(&s.prop_trace, sizeof(double) * std::max(1u, nsteps) * np));
      s.trace_steps = nsteps;
    }
    ChainState& st = b->st_host[e];
    init_state_common(st, np, b->nsources, ns, b->means.data(), b->sigmas.data(), b->fixed.data(), nsteps,
                      burnin_fraction, debug_mode);
    st.seed = s.seed; st.stream_id = s.stream_id;
    st.tape_normals = s.tape_n; st.n_normals = s.n_tape_n;
    st.tape_uniforms = s.tape_u; st.n_uniforms = s.n_tape_u;
    st.rows = s.rows; st.row_cap = s.row_cap;
    st.nll_trace = b->trace ? s.nll_trace : nullptr;
    st.prop_trace = b->trace ? s.prop_trace : nullptr;
    for (int j = 0; j < ns; ++j) {
      NllSignal& q = sg[(size_t)e * ns + j];
      memset(&q, 0, sizeof q);
      q.read_bins = s.read_bins + (b->shared_rb ? 0 : (size_t)j * s.rb_cap);
      q.nexpected = b->nexpected[j];
      q.bin_volume = b->pdfs[j]->bin_volume;
      q.n_mc = b->n_mc[j];
      q.total_nbins = b->total_nbins;
      q.source_id = b->source_id[j];
    }
    nev[e] = (long long)s.nevents;
  }
  SXMC_CUDA(cudaMemcpyAsync(b->st_dev, b->st_host, sizeof(ChainState) * nchains, cudaMemcpyHostToDevice, b->stream));
  SXMC_CUDA(cudaMemcpyAsync(b->sigs_dev, sg.data(), sizeof(NllSignal) * sg.size(), cudaMemcpyHostToDevice, b->stream));
  SXMC_CUDA(cudaMemcpyAsync(b->nevents_dev, nev.data(), sizeof(long long) * nchains, cudaMemcpyHostToDevice, b->stream));
  SXMC_CUDA(cudaMemsetAsync(b->hist_dev, 0, sizeof(unsigned) * b->hist_words * nchains, b->stream));
  SXMC_CUDA(cudaMemsetAsync(b->sync_dev, 0, sizeof(unsigned) * 4, b->stream));
  BatchParams Q;
  memset(&Q, 0, sizeof Q);
  Q.st = b->st_dev; Q.sigs = b->sigs_dev; Q.hist = b->hist_dev; Q.hist_words = b->hist_words;
  Q.nevents = b->nevents_dev; Q.sync = b->sync_dev;
  Q.nchains = nchains; Q.ns = ns; Q.np = np; Q.nsources = b->nsources;
  Q.nsteps = nsteps; Q.do_init = 1;
  Q.rebin = (b->nsystematics > 0 && !b->systematics_fixed) ? 1 : 0;   // src/mcmc.cpp:264
  Q.stamps = b->stamps_dev;
  Q.mult = b->mult_dev;
  Q.mono = (b->mono && !b->tune.no_mono) ? 1 : 0;
  SXMC_CUDA(cudaEventRecord(b->ev0, b->stream));
  if (use_mono) SXMC_TRY(launch_chain_batch_mono(*b->mono_plan, Q, b->stream));
  else SXMC_TRY(launch_chain_batch(plan, Q, b->stream));
  b->last_path = use_mono ? 1 : 2;
  SXMC_CUDA(cudaEventRecord(b->ev1, b->stream));
  SXMC_CUDA(cudaMemcpyAsync(b->sync_host, b->sync_dev, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost, b->stream));
  SXMC_CUDA(cudaMemcpyAsync(b->st_host, b->st_dev, sizeof(ChainState) * nchains, cudaMemcpyDeviceToHost, b->stream));
  SXMC_CUDA(cudaStreamSynchronize(b->stream));
  if (b->sync_host[3]) { set_error("chain batch kernel timed out at a grid barrier"); return SXMC_ESTATE; }
  SXMC_CUDA(cudaEventElapsedTime(&b->loop_ms, b->ev0, b->ev1));
  for (int e = 0; e < nchains; ++e) {
    b->slots[e].nrows = b->st_host[e].nrows;
    b->slots[e].accepted = b->st_host[e].accepted;
  }
  b->last_n = nchains;
  return SXMC_OK;
}

int sxmc_batch_nrows(sxmc_batch* b, int slot, long* nrows_out, long* accepted_out) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  if (nrows_out) *nrows_out = (long)std::min(s->nrows, s->row_cap);
  if (accepted_out) *accepted_out = (long)s->accepted;
  return SXMC_OK;
}

int sxmc_batch_read(sxmc_batch* b, int slot, float* rows_out, long max_rows) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  const long long n = std::min<long long>(std::min(s->nrows, s->row_cap), max_rows);
  if (n > 0)
    SXMC_CUDA(cudaMemcpy(rows_out, s->rows, sizeof(float) * (size_t)n * (b->nparameters + 1), cudaMemcpyDeviceToHost));
  return SXMC_OK;
}

int sxmc_batch_read_trace(sxmc_batch* b, int slot, double* nll_trace_out, double* proposed_out, unsigned nsteps) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  if (!s->nll_trace || nsteps > s->trace_steps) { set_error("sxmc_batch_read_trace: trace not enabled"); return SXMC_ESTATE; }
  SXMC_CUDA(cudaMemcpy(nll_trace_out, s->nll_trace, sizeof(double) * nsteps, cudaMemcpyDeviceToHost));
  SXMC_CUDA(cudaMemcpy(proposed_out, s->prop_trace, sizeof(double) * nsteps * b->nparameters, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}

int sxmc_batch_contour(sxmc_batch* b, int slot, float delta, float* lower_out, float* upper_out,
                       float* point_out, float* lmin_out, long* ncontour_out) {
  BatchSlot* s; SXMC_TRY(batch_slot(b, slot, &s));
  const long n = (long)std::min(s->nrows, s->row_cap);
  return sxmc_rows_contour(s->rows, n, b->nparameters, delta, lower_out, upper_out, point_out, lmin_out,
                           ncontour_out);
}

int sxmc_batch_loop_ms(sxmc_batch* b, float* ms_out) { *ms_out = b->loop_ms; return SXMC_OK; }
int sxmc_batch_last_path(sxmc_batch* b, int* path_out, int* max_edges_out) {
  *path_out = b->last_path;
  if (max_edges_out) *max_edges_out = b->mono_plan ? b->mono_plan->max_edges : 0;
  return SXMC_OK;
}

int sxmc_batch_phase_stamps(sxmc_batch* b, int enable, unsigned long long* out148x8) {
  if (enable && !b->stamps_dev) {
    SXMC_CUDA(cudaMalloc(&b->stamps_dev, sizeof(unsigned long long) * (148 * 8 + 16)));
    SXMC_CUDA(cudaMemset(b->stamps_dev, 0, sizeof(unsigned long long) * (148 * 8 + 16)));
  }
  if (out148x8 && b->stamps_dev)   // + 16 words of per-chain clock deltas of one warp (monotone kernel)
    SXMC_CUDA(cudaMemcpy(out148x8, b->stamps_dev, sizeof(unsigned long long) * (148 * 8 + 16), cudaMemcpyDeviceToHost));
  if (!enable && b->stamps_dev) { cudaFree(b->stamps_dev); b->stamps_dev = nullptr; }
  return SXMC_OK;
}

}  // extern "C"
