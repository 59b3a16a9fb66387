// capi_hist.cu -- the extern "C" layer of libsxmc_b200.so, part 1 (declared in include/sxmc_b200.h):
// library and error state, device buffers (the device half of hemi::Array), kernel-selection knobs,
// and the pdfz::EvalHist handles: argument validation with the reference's pdfz::Error messages,
// layout preparation, launch planning, batched evaluation.
#include "capi_internal.h"

namespace sxmc {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d in `%s`", (int)e, cudaGetErrorString(e), file,
           line, what);
  g_error = buf;
  return SXMC_ECUDA;
}

Tuning g_tune;
int g_timing = 0;
float g_last_timed[2] = {-1.f, -1.f};   // bin, nll of the most recent timed pass of ANY chain (sxmc_last_kernel_ms)
static bool g_init_done = false;

int ensure_init() {
  if (g_init_done) return SXMC_OK;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    set_error("no usable CUDA device: libsxmc_b200 has no CPU fallback");
    cudaGetLastError();
    return SXMC_ECUDA;
  }
  int dev = 0, major = 0;
  SXMC_CUDA(cudaGetDevice(&dev));
  SXMC_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) {
    set_error("libsxmc_b200 is built for sm_100a (Blackwell B200) only");
    return SXMC_ECUDA;
  }
  g_init_done = true;
  return SXMC_OK;
}


}  // namespace sxmc

using namespace sxmc;

namespace sxmc {

void fill_prep_geo(const sxmc_hist* h, PrepGeo& g) {
  memset(&g, 0, sizeof g);
  for (int i = 0; i < h->nobs; ++i) {
    g.lower[i] = h->lower[i];
    g.upper[i] = h->upper[i];
    g.scale[i] = h->nbins[i] / (h->upper[i] - h->lower[i]);   // src/pdfz.cpp:365-368
    g.nbins[i] = h->nbins[i];
    g.stride[i] = h->stride[i];
  }
  g.total_nbins = h->total_nbins;
}

static int wanted_primary(const sxmc_hist* h) {
  if (!h->systs.empty() && h->systs[0].obs < h->nfields) return h->systs[0].obs;
  return 0;
}


// Undo the tiled layout: the caller's rows, in their original order, back in aos_dev.
static int recover_rows(sxmc_hist* h) {
  const unsigned long long n = h->nsamples;
  if (n > 0) {
    SXMC_CUDA(cudaMalloc(&h->aos_dev, sizeof(float) * n * h->nfields));
    SXMC_TRY(unprepare_to_aos(h->cols_dev, h->col_pitch, h->perm_dev, n, h->nfields, h->aos_dev, h->stream));
    SXMC_CUDA(cudaStreamSynchronize(h->stream));
  }
  cudaFree(h->cols_dev); h->cols_dev = nullptr;
  cudaFree(h->perm_dev); h->perm_dev = nullptr;
  h->prepared = false;
  return SXMC_OK;
}

// Build (or rebuild) the sorted, tiled SoA layout.
int ensure_prepared(sxmc_hist* h) {
  const int primary = wanted_primary(h);
  if (h->prepared && (!h->sort || primary == h->primary)) return SXMC_OK;
  const unsigned long long n = h->nsamples;
  if (h->prepared) SXMC_TRY(recover_rows(h));   // systematics changed the sort key
  h->nwtiles = (n + WT_SAMPLES - 1) / WT_SAMPLES;
  h->col_pitch = h->nwtiles * WT_SAMPLES;
  if (h->col_pitch > 0) {
    SXMC_CUDA(cudaMalloc(&h->cols_dev, sizeof(float) * h->col_pitch * h->nfields));
    SXMC_CUDA(cudaMalloc(&h->perm_dev, sizeof(unsigned) * n));
  }
  PrepGeo g;
  fill_prep_geo(h, g);
  int rc = prepare_layout(h->aos_dev, n, h->nfields, h->nobs, g, primary, h->sort, h->cols_dev,
                          h->col_pitch, h->perm_dev, h->stream);
  if (rc) return rc;
  cudaFree(h->aos_dev);
  h->aos_dev = nullptr;
  h->primary = primary;
  h->prepared = true;
  return SXMC_OK;
}

// Fill one BinJob for handle h writing into (bins, norm) and reading (params, stride).
int fill_job(sxmc_hist* h, BinJob& job, unsigned* bins, unsigned* norm, const double* params,
                    long long param_stride) {
  int rc = ensure_prepared(h);
  if (rc) return rc;
  memset(&job, 0, sizeof job);
  for (int f = 0; f < h->nfields; ++f) job.cols[f] = h->cols_dev + (size_t)f * h->col_pitch;
  job.nwtiles = h->nwtiles;
  job.bins = bins;
  job.norm = norm;
  job.params = params;
  job.param_stride = param_stride;
  job.total_nbins = h->total_nbins;
  job.nobs = h->nobs;
  job.nfields = h->nfields;
  job.nsyst = (int)h->systs.size();
  for (int i = 0; i < h->nobs; ++i) {
    job.lower[i] = h->lower[i];
    job.upper[i] = h->upper[i];
    job.scale[i] = h->nbins[i] / (h->upper[i] - h->lower[i]);
    job.stride[i] = h->stride[i];
  }
  // register slots: observables first, then every other field a systematic touches
  int slot_of[SXMC_MAX_NFIELDS];
  for (int f = 0; f < SXMC_MAX_NFIELDS; ++f) slot_of[f] = -1;
  int ns = 0;
  bool overflow = false;
  for (int o = 0; o < h->nobs; ++o) {
    if (ns < MAX_SLOTS) { job.slot_field[ns] = o; slot_of[o] = ns; }
    else overflow = true;
    ++ns;
  }
  for (size_t k = 0; k < h->systs.size(); ++k) {
    SystDesc d = h->systs[k];
    const int need[2] = {d.obs, d.type == SXMC_SYST_RESOLUTION_SCALE ? d.extra : d.obs};
    for (int t = 0; t < 2; ++t) {
      if (slot_of[need[t]] < 0) {
        if (ns < MAX_SLOTS) { job.slot_field[ns] = need[t]; slot_of[need[t]] = ns; }
        else overflow = true;
        ++ns;
      }
    }
    d.obs_slot = (short)std::max(0, slot_of[d.obs]);
    d.extra_slot = (short)(d.type == SXMC_SYST_RESOLUTION_SCALE ? std::max(0, slot_of[d.extra]) : 0);
    job.syst[k] = d;
  }
  job.nslots = overflow ? MAX_SLOTS + 1 : ns;
  return SXMC_OK;
}

// Four-row chunks (bin_fast_plan_big) are taken where they were measured to win: launches that stream
// a gigabyte or more into global-memory histograms of moderate size -- workload D on one GPU, 299 ->
// 275 us per step.  Not below (half of D: 153.0 -> 154.6 us, an eighth: 58.3 -> 61.1), not with a
// shared-memory histogram (C: 195 -> 205 us: two stages of four rows pipeline worse than four of two
// at 16 warps), not with 200 000 bins (311 -> 322 us: 12 warps hide the reductions' latency worse).
// profiles/r02_bin_chunk_rows_ab.txt
static const unsigned long long BIG_CHUNK_MIN_BYTES = 1ull << 30;
static const int BIG_CHUNK_MAX_BINS = 65536;

// Decide the launch shape for a filled batch of n jobs.
void plan_launch(Batch& b, int n, const Tuning& T) {
  BinLaunch& L = b.launch;
  memset(&L, 0, sizeof L);
  L.jobs_host = b.host;
  L.jobs_dev = b.dev;
  L.prefix_dev = b.prefix_dev;
  L.njobs = n;
  L.cta_split = T.cta_split;
  bool uniform = true;
  int max_bins = 0, max_slots = 0;
  bool poly = false;
  for (int j = 0; j < n; ++j) {
    const BinJob& job = b.host[j];
    max_bins = std::max(max_bins, job.total_nbins);
    max_slots = std::max(max_slots, job.nslots);
    const BinJob& j0 = b.host[0];
    if (job.nobs != j0.nobs || job.total_nbins != j0.total_nbins) uniform = false;
    for (int o = 0; uniform && o < job.nobs; ++o)
      if (job.lower[o] != j0.lower[o] || job.upper[o] != j0.upper[o] || job.scale[o] != j0.scale[o] ||
          job.stride[o] != j0.stride[o])
        uniform = false;
    for (int k = 0; k < job.nsyst; ++k)
      if (job.syst[k].npars > 1) poly = true;   // per-sample polynomials go through the generic kernel
    if (job.nsyst > 4) poly = true;             // FAST_MAX_OPS register-resident systematics
  }
  L.max_total_nbins = max_bins;
  L.nobs = n ? b.host[0].nobs : 0;
  L.nslots = max_slots;
  L.smem_hist = max_bins <= T.smem_hist_max_bins;
  L.blocks_per_sm = T.blocks_per_sm;
  L.fast = n > 0 && uniform && !poly && !T.force_generic && L.nobs <= 3 &&
           bin_fast_supported(L.nobs, max_slots);
  if (L.fast) {
    // pipeline stages share the shared memory with the privatised histogram; when the
    // histogram leaves no room for them it stays in global memory (L2 atomics)
    if (!(L.smem_hist && bin_fast_plan(max_slots, max_bins, BIN_STATIC_RESERVE, &L.rows_per_stage,
                                       &L.nstages, &L.stage_off, &L.smem_bytes))) {
      L.smem_hist = false;
      bin_fast_plan(max_slots, 0, BIN_STATIC_RESERVE, &L.rows_per_stage, &L.nstages, &L.stage_off,
                    &L.smem_bytes);
    }
    L.nwarps = FAST_WARPS;
    // launches that stream hundreds of megabytes: chunks of four rows (bin_fast_plan_big)
    unsigned long long stream_bytes = 0;
    for (int j = 0; j < n; ++j) stream_bytes += b.host[j].nwtiles * (unsigned long long)WT_SAMPLES * 4ull * max_slots;
    const bool want_big = T.chunk_form == 2 || (T.chunk_form == 0 && !L.smem_hist && stream_bytes >= BIG_CHUNK_MIN_BYTES &&
                                                max_bins <= BIG_CHUNK_MAX_BINS);
    if (want_big) {
      int nw = 0, R = 0, S = 0; unsigned off = 0; size_t sm = 0;
      if (bin_fast_plan_big(max_slots, L.smem_hist ? max_bins : 0, BIN_STATIC_RESERVE, &nw, &R, &S, &off, &sm)) {
        L.nwarps = nw; L.rows_per_stage = R; L.nstages = S; L.stage_off = off; L.smem_bytes = sm;
      }
    }
    // every job of a fast launch uses the same slot count (columns it does not need are
    // simply re-reads of observable 0)
    for (int j = 0; j < n; ++j) {
      BinJob& job = b.host[j];
      for (int s = job.nslots; s < max_slots; ++s) job.slot_field[s] = 0;
      job.nslots = max_slots;
    }
    for (int o = 0; o < L.nobs; ++o) {
      L.lower[o] = b.host[0].lower[o]; L.upper[o] = b.host[0].upper[o];
      L.scale[o] = b.host[0].scale[o]; L.stride[o] = b.host[0].stride[o];
    }
    // the systematic program, if every job runs the same one and a kernel was built for it
    L.prog_nops = -1;
    L.prog_codes = 0;
    if (!T.force_dynamic) {
      int nops = b.host[0].nsyst;
      unsigned codes = 0;
      bool same = true;
      for (int j = 0; j < n && same; ++j) {
        const BinJob& job = b.host[j];
        if (job.nsyst != nops) { same = false; break; }
        unsigned cj = 0;
        for (int k = 0; k < job.nsyst; ++k)
          cj |= (unsigned)(job.syst[k].type | (job.syst[k].obs_slot << 2) | (job.syst[k].extra_slot << 5)) << (8 * k);
        if (j == 0) codes = cj;
        else if (cj != codes) same = false;
      }
      if (same && bin_fast_has_program(L.nobs, max_slots, nops, codes)) {
        L.prog_nops = nops;
        L.prog_codes = codes;
      } else if (same && nops >= 1 && nops <= FAST_MAX_OPS && !T.no_jit &&
                 jit_bin_fast_kernel(L.nobs, max_slots, L.smem_hist, nops, codes)) {
        // not a program the build instantiated: compiled for this fit now (NVRTC, cached), so that
        // the steps never wait for it
        L.prog_nops = nops;
        L.prog_codes = codes;
        L.jit = true;
      }
    }
  }
  // work units: chunks of rows_per_stage rows (fast) or block-tiles of BIN_WARPS warp-tiles
  unsigned long long total = 0;
  b.prefix_host[0] = 0;
  for (int j = 0; j < n; ++j) {
    const BinJob& job = b.host[j];
    if (L.fast) total += job.nwtiles * (unsigned long long)(WT_ROWS / L.rows_per_stage);
    else total += (job.nwtiles + BIN_WARPS - 1) / BIN_WARPS;
    b.prefix_host[j + 1] = total;
  }
  L.total_units = total;
}

int upload_batch(Batch& b, int n, cudaStream_t stream) {
  SXMC_CUDA(cudaMemcpyAsync(b.dev, b.host, sizeof(BinJob) * n, cudaMemcpyHostToDevice, stream));
  SXMC_CUDA(cudaMemcpyAsync(b.prefix_dev, b.prefix_host, sizeof(unsigned long long) * (n + 1),
                            cudaMemcpyHostToDevice, stream));
  if (b.n_ep > 0)
    SXMC_CUDA(cudaMemcpyAsync(b.ep_dev, b.ep_host, sizeof(EvalPdfJob) * b.n_ep, cudaMemcpyHostToDevice, stream));
  return SXMC_OK;
}

}  // namespace sxmc

extern "C" {

const char* sxmc_last_error(void) { return g_error.c_str(); }
const char* sxmc_version(void) { return "sxmc_b200 0.1 (sm_100a)"; }

int sxmc_init(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    set_error("no usable CUDA device: libsxmc_b200 has no CPU fallback");
    return SXMC_ECUDA;
  }
  SXMC_CUDA(cudaSetDevice(device));
  g_init_done = false;
  return ensure_init();
}

int sxmc_device_synchronize(void) {
  SXMC_TRY(ensure_init());
  SXMC_CUDA(cudaDeviceSynchronize());
  return SXMC_OK;
}

int sxmc_buf_alloc(size_t bytes, void** dev_out) {
  SXMC_TRY(ensure_init());
  SXMC_CUDA(cudaMalloc(dev_out, bytes ? bytes : 1));
  return SXMC_OK;
}
int sxmc_buf_free(void* dev) { if (dev) SXMC_CUDA(cudaFree(dev)); return SXMC_OK; }
int sxmc_buf_upload(void* dev, const void* host, size_t bytes) {
  if (bytes) SXMC_CUDA(cudaMemcpy(dev, host, bytes, cudaMemcpyHostToDevice));
  return SXMC_OK;
}
int sxmc_buf_download(void* host, const void* dev, size_t bytes) {
  if (bytes) SXMC_CUDA(cudaMemcpy(host, dev, bytes, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_buf_zero(void* dev, size_t bytes) {
  if (bytes) SXMC_CUDA(cudaMemset(dev, 0, bytes));
  return SXMC_OK;
}
int sxmc_host_alloc(size_t bytes, void** host_out) {
  SXMC_TRY(ensure_init());
  SXMC_CUDA(cudaMallocHost(host_out, bytes ? bytes : 1));
  return SXMC_OK;
}
int sxmc_host_free(void* host) { if (host) SXMC_CUDA(cudaFreeHost(host)); return SXMC_OK; }

int sxmc_set_bin_tuning(int smem_hist_max_bins, int blocks_per_sm, int force_generic) {
  // the defaults of handles created from now on (existing handles keep the knobs they were created with)
  if (smem_hist_max_bins >= 0) g_tune.smem_hist_max_bins = std::min(smem_hist_max_bins, 56 * 1024);
  g_tune.blocks_per_sm = blocks_per_sm;
  g_tune.force_generic = force_generic & 1;
  g_tune.force_dynamic = (force_generic >> 1) & 1;
  g_tune.no_mono = (force_generic >> 2) & 1;
  g_tune.p2p_form = (force_generic >> 3) & 3;
  g_tune.no_group = (force_generic >> 5) & 1;
  g_tune.no_pdl = (force_generic >> 6) & 1;
  g_tune.no_jit = (force_generic >> 7) & 1;
  g_tune.cta_split = (force_generic >> 8) & 3;
  g_tune.chunk_form = (force_generic >> 10) & 3;
  return SXMC_OK;
}
int sxmc_jit_compile_check(int nobs, int nslots, int smem_hist, int nops, unsigned codes, size_t* cubin_bytes_out) {
  return jit_compile_check(nobs, nslots, smem_hist, nops, codes, cubin_bytes_out);
}
int sxmc_nll_debug_stamps(int enable, unsigned long long* out1202x8) {
  static unsigned long long* dev = nullptr;
  const size_t bytes = sizeof(unsigned long long) * 1202 * 8;
  if (enable && !dev) {
    SXMC_CUDA(cudaMalloc(&dev, bytes));
    SXMC_CUDA(cudaMemset(dev, 0, bytes));
  }
  if (out1202x8 && dev) {
    SXMC_CUDA(cudaDeviceSynchronize());
    SXMC_CUDA(cudaMemcpy(out1202x8, dev, bytes, cudaMemcpyDeviceToHost));
  }
  if (!enable && dev) { cudaFree(dev); dev = nullptr; }
  set_nll_xdbg(enable ? dev : nullptr);
  return SXMC_OK;
}
int sxmc_enable_kernel_timing(int on) { g_timing = on; return SXMC_OK; }
int sxmc_bin_debug_stamps(int enable, unsigned long long* out256x32) {
  static unsigned long long* dev = nullptr;
  const size_t bytes = sizeof(unsigned long long) * 256 * 32;
  if (enable && !dev) {
    SXMC_CUDA(cudaMalloc(&dev, bytes));
    SXMC_CUDA(cudaMemset(dev, 0, bytes));
  }
  if (out256x32 && dev) {
    SXMC_CUDA(cudaDeviceSynchronize());
    SXMC_CUDA(cudaMemcpy(out256x32, dev, bytes, cudaMemcpyDeviceToHost));
  }
  if (!enable && dev) { cudaFree(dev); dev = nullptr; }
  set_bin_xdbg(enable ? dev : nullptr);
  return SXMC_OK;
}
int sxmc_last_kernel_ms(float* bin_ms, float* nll_ms) {
  if (bin_ms) *bin_ms = g_last_timed[0];
  if (nll_ms) *nll_ms = g_last_timed[1];
  return SXMC_OK;
}

// ---------------------------------------------------------------------------------------
// EvalHist
// ---------------------------------------------------------------------------------------
int sxmc_hist_create(const float* samples, size_t nfloats, int nfields, int nobservables,
                     const double* lower, int nlower, const double* upper, int nupper,
                     const int* nbins, int nnbins, unsigned dataset, sxmc_hist** out) {
  if (!out) SXMC_INVAL("sxmc_hist_create: out is NULL");
  *out = nullptr;
  // Eval::Eval checks, in the reference's order and wording (src/pdfz.cpp:63-81)
  if (nfields <= 0 || nfloats % (size_t)nfields != 0)
    SXMC_INVAL("Length of samples array is not divisible by number of fields.");
  if (nobservables == 0) SXMC_INVAL("Number of observables in PDF is zero.");
  if (nobservables > nfields)
    SXMC_INVAL("Number of observables cannot be greater than number of fields.");
  if (nupper != nobservables)
    SXMC_INVAL("Number of upper bounds must be same as number of observables.");
  if (nlower != nobservables)
    SXMC_INVAL("Number of lower bounds must be same as number of observables.");
  // EvalHist::EvalHist checks (src/pdfz.cpp:189-195)
  if (nnbins != nobservables)
    SXMC_INVAL("Size of nbins array must be same as number of observables.");
  if (nfields > SXMC_MAX_NFIELDS)
    SXMC_INVAL("Exceeded maximum number of fields per sample. Edit MAX_NFIELDS in pdfz.cpp to fix this!");
  if (nobservables < 0) SXMC_INVAL("Number of observables is negative.");

  sxmc_hist* h = new sxmc_hist;
  h->tune = g_tune;
  h->nfields = nfields;
  h->nobs = nobservables;
  h->dataset = dataset;
  h->lower.assign(lower, lower + nobservables);
  h->upper.assign(upper, upper + nobservables);
  h->nbins.assign(nbins, nbins + nobservables);
  h->stride.assign(nobservables, 1);
  h->bin_volume = 1.0f;   // src/pdfz.cpp:201-204
  for (int i = 0; i < nobservables; ++i) h->bin_volume *= (upper[i] - lower[i]) / nbins[i];
  for (int i = nobservables - 2; i >= 0; --i) h->stride[i] = nbins[i + 1] * h->stride[i + 1];
  long long total = (long long)h->stride[0] * nbins[0];
  for (int i = 0; i < nobservables; ++i) if (nbins[i] < 0) total = 0;
  if (total == 0) { delete h; SXMC_INVAL("Cannot make histogram with zero bins."); }
  if (total > 0x7fffffffLL) { delete h; SXMC_INVAL("Histogram has more than 2^31-1 bins."); }
  h->total_nbins = (int)total;
  h->nsamples = nfloats / (size_t)nfields;

  int rc = ensure_init();
  if (rc) { delete h; return rc; }
  cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(&h->bins_dev, sizeof(unsigned) * (size_t)h->total_nbins);
  if (e == cudaSuccess) e = cudaMemset(h->bins_dev, 0, sizeof(unsigned) * (size_t)h->total_nbins);
  if (e == cudaSuccess && nfloats > 0) {
    e = cudaMalloc(&h->aos_dev, sizeof(float) * nfloats);
    if (e == cudaSuccess) e = cudaMemcpy(h->aos_dev, samples, sizeof(float) * nfloats, cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&h->ev1);
  if (e != cudaSuccess) { sxmc_hist_destroy(h); return cuda_fail(e, "sxmc_hist_create", __FILE__, __LINE__); }
  *out = h;
  return SXMC_OK;
}

int sxmc_hist_destroy(sxmc_hist* h) {
  if (!h) return SXMC_OK;
  if (h->stream) cudaStreamSynchronize(h->stream);
  cudaFree(h->aos_dev); cudaFree(h->cols_dev); cudaFree(h->perm_dev);
  cudaFree(h->bins_dev); cudaFree(h->read_bins_dev);
  h->batch.release();
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return SXMC_OK;
}

int sxmc_hist_add_systematic(sxmc_hist* h, int type, int obs, int extra_field, int npars,
                             const short* pars) {
  if (!h) SXMC_INVAL("sxmc_hist_add_systematic: NULL handle");
  if (type < 0 || type > 3) SXMC_INVAL("Unknown systematic type");   // src/pdfz.cpp:170
  if ((int)h->systs.size() >= SXMC_MAX_SYST) SXMC_INVAL("Too many systematics on one PDF (SXMC_MAX_SYST).");
  if (npars < 0 || npars > SXMC_MAX_NPARS) SXMC_INVAL("Too many coefficients in one systematic (SXMC_MAX_NPARS).");
  if (obs < 0 || obs >= h->nfields) SXMC_INVAL("Systematic observable index is not a field of the samples.");
  if (type == SXMC_SYST_RESOLUTION_SCALE && (extra_field < 0 || extra_field >= h->nfields))
    SXMC_INVAL("Systematic truth-field index is not a field of the samples.");
  SystDesc d;
  memset(&d, 0, sizeof d);
  d.type = (short)type;
  d.obs = (short)obs;
  d.extra = (short)(type == SXMC_SYST_RESOLUTION_SCALE ? extra_field : 0);
  d.npars = (short)npars;
  for (int i = 0; i < npars; ++i) d.pidx[i] = pars[i];
  h->systs.push_back(d);
  return SXMC_OK;
}

// SetEvalPoints from points already on the device, enqueued on `stream` without a host sync
// (sxmc_chain_set_data uploads the data events once for all PDFs of a fit).
}  // extern "C"
namespace sxmc {
int hist_set_eval_points_dev(sxmc_hist* h, const float* pts_dev, size_t np, cudaStream_t stream) {
  if (np > h->read_bins_cap || !h->read_bins_dev) {
    cudaFree(h->read_bins_dev);
    h->read_bins_dev = nullptr;
    h->read_bins_cap = 0;
    SXMC_CUDA(cudaMalloc(&h->read_bins_dev, sizeof(int) * (np ? np : 1)));
    h->read_bins_cap = np ? np : 1;
  }
  h->npoints = np;
  if (np == 0) return SXMC_OK;
  PrepGeo g;
  fill_prep_geo(h, g);
  return launch_read_bins(pts_dev, (long long)np, h->nobs, g, h->dataset, h->read_bins_dev, stream);
}
}  // namespace sxmc
extern "C" {

int sxmc_hist_set_eval_points(sxmc_hist* h, const float* points, size_t nfloats) {
  if (!h) SXMC_INVAL("sxmc_hist_set_eval_points: NULL handle");
  if (nfloats % (size_t)(h->nobs + 1) != 0)
    SXMC_INVAL("Number of entries in evaluation points array not divisible by number of observables.");
  const size_t np = nfloats / (size_t)(h->nobs + 1);
  float* pts = nullptr;
  if (np) SXMC_CUDA(cudaMalloc(&pts, sizeof(float) * nfloats));
  cudaError_t e = np ? cudaMemcpyAsync(pts, points, sizeof(float) * nfloats, cudaMemcpyHostToDevice, h->stream) : cudaSuccess;
  int rc = SXMC_OK;
  if (e == cudaSuccess) {
    rc = hist_set_eval_points_dev(h, pts, np, h->stream);
    e = cudaStreamSynchronize(h->stream);
  }
  cudaFree(pts);
  if (e != cudaSuccess) return cuda_fail(e, "sxmc_hist_set_eval_points", __FILE__, __LINE__);
  return rc;
}

int sxmc_hist_set_pdf_buffer(sxmc_hist* h, float* dev, long offset, long stride) {
  h->pdf_dev = dev; h->pdf_offset = offset; h->pdf_stride = stride; return SXMC_OK;
}
int sxmc_hist_set_norm_buffer(sxmc_hist* h, unsigned* dev, long offset) {
  h->norm_dev = dev; h->norm_offset = offset; return SXMC_OK;
}
int sxmc_hist_set_param_buffer(sxmc_hist* h, const double* dev, long offset, long stride) {
  h->param_dev = dev; h->param_offset = offset; h->param_stride = stride; return SXMC_OK;
}

int sxmc_hist_set_layout(sxmc_hist* h, int sort) {
  const bool want = sort != 0;
  if (h->prepared && h->sort != want) SXMC_TRY(recover_rows(h));
  h->sort = want;
  return SXMC_OK;
}
int sxmc_hist_prepare(sxmc_hist* h) { return ensure_prepared(h); }

static int hists_eval_on(sxmc_hist* const* hs, int n, int do_eval_pdf, cudaStream_t stream, Batch& b) {
  SXMC_TRY(ensure_init());
  SXMC_TRY(b.reserve(n));
  b.n_ep = 0;
  b.max_points = 0;
  for (int j = 0; j < n; ++j) {
    sxmc_hist* h = hs[j];
    if (!h->norm_dev) { set_error("EvalAsync: no normalization buffer bound"); return SXMC_ESTATE; }
    if (!h->systs.empty() && !h->param_dev) { set_error("EvalAsync: no parameter buffer bound"); return SXMC_ESTATE; }
    SXMC_TRY(fill_job(h, b.host[j], h->bins_dev, h->norm_dev + h->norm_offset,
                      h->param_dev ? h->param_dev + h->param_offset : nullptr, h->param_stride));
    if (do_eval_pdf && h->read_bins_dev && h->npoints > 0) {   // :474 guard
      if (!h->pdf_dev) { set_error("EvalAsync: no PDF value buffer bound"); return SXMC_ESTATE; }
      EvalPdfJob& e = b.ep_host[b.n_ep++];
      e.read_bins = h->read_bins_dev;
      e.bins = h->bins_dev;
      e.norm = h->norm_dev + h->norm_offset;
      e.out = h->pdf_dev + h->pdf_offset;
      e.stride = h->pdf_stride;
      e.npoints = (long long)h->npoints;
      e.bin_volume = h->bin_volume;
      e.total_nbins = h->total_nbins;
      b.max_points = std::max(b.max_points, e.npoints);
    }
  }
  plan_launch(b, n, hs[0]->tune);
  SXMC_TRY(upload_batch(b, n, stream));
  SXMC_TRY(launch_zero_jobs(b.dev, n, stream));   // zero_hist (src/pdfz.cpp:334-346), all PDFs at once
  SXMC_TRY(launch_bin(b.launch, stream));
  SXMC_TRY(launch_eval_pdf(b.ep_dev, b.n_ep, b.max_points, stream));
  return SXMC_OK;
}

int sxmc_hist_eval_async(sxmc_hist* h, int do_eval_pdf) {
  if (!h) SXMC_INVAL("sxmc_hist_eval_async: NULL handle");
  sxmc_hist* one[1] = {h};
  return hists_eval_on(one, 1, do_eval_pdf, h->stream, h->batch);
}
int sxmc_hist_eval_finished(sxmc_hist* h) {
  SXMC_CUDA(cudaStreamSynchronize(h->stream));
  return SXMC_OK;
}
int sxmc_hists_eval_async(sxmc_hist* const* hs, int n, int do_eval_pdf) {
  if (n <= 0) return SXMC_OK;
  // handles evaluated individually may still be running on their own streams
  for (int j = 1; j < n; ++j) SXMC_CUDA(cudaStreamSynchronize(hs[j]->stream));
  return hists_eval_on(hs, n, do_eval_pdf, hs[0]->stream, hs[0]->batch);
}

int sxmc_hist_total_nbins(const sxmc_hist* h) { return h->total_nbins; }
double sxmc_hist_bin_volume(const sxmc_hist* h) { return h->bin_volume; }
size_t sxmc_hist_nsamples(const sxmc_hist* h) { return h->nsamples; }
size_t sxmc_hist_neval_points(const sxmc_hist* h) { return h->npoints; }
int sxmc_hist_get_bins(sxmc_hist* h, unsigned* host_out) {
  SXMC_CUDA(cudaStreamSynchronize(h->stream));
  SXMC_CUDA(cudaMemcpy(host_out, h->bins_dev, sizeof(unsigned) * (size_t)h->total_nbins, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_hist_get_read_bins(sxmc_hist* h, int* host_out) {
  if (h->npoints) SXMC_CUDA(cudaMemcpy(host_out, h->read_bins_dev, sizeof(int) * h->npoints, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_hist_get_samples(sxmc_hist* h, float* host_out) {
  const size_t nf = h->nsamples * (size_t)h->nfields;
  if (nf == 0) return SXMC_OK;
  if (!h->prepared) {
    SXMC_CUDA(cudaMemcpy(host_out, h->aos_dev, sizeof(float) * nf, cudaMemcpyDeviceToHost));
    return SXMC_OK;
  }
  float* tmp = nullptr;
  SXMC_CUDA(cudaMalloc(&tmp, sizeof(float) * nf));
  int rc = unprepare_to_aos(h->cols_dev, h->col_pitch, h->perm_dev, h->nsamples, h->nfields, tmp, h->stream);
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (e == cudaSuccess) e = cudaMemcpy(host_out, tmp, sizeof(float) * nf, cudaMemcpyDeviceToHost);
  cudaFree(tmp);
  if (e != cudaSuccess) return cuda_fail(e, "sxmc_hist_get_samples", __FILE__, __LINE__);
  return rc;
}

}  // extern "C"
