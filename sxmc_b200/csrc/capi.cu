// capi.cu -- the extern "C" layer of libsxmc_b200.so (declared in include/sxmc_b200.h):
// handles, argument validation with the reference's pdfz::Error messages, layout
// preparation, batched launches, the device-resident MCMC loop (CUDA graph) and NCCL glue.
#include "common.cuh"
#include "kernels.h"

#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

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

static int g_smem_hist_max_bins = 8192;
static int g_blocks_per_sm = 0;
static int g_force_generic = 0;
static int g_force_dynamic = 0;
static int g_p2p_form = 0;       // peer-memory all-reduce: 0 automatic (two-shot above 2 ranks), 1 one-shot, 2 two-shot
static int g_no_mono = 0;        // chain batch: never take the monotone endpoint path (tests)
static int g_no_jit = 0;         // programs outside the build's table keep the run-time dispatching kernel
static int g_no_pdl = 0;         // steps without programmatic dependent launch (measurement knob)
static int g_no_group = 0;       // lookup: never group the events by bin (the multi-dataset form, on request)
static long long g_persistent_max_bytes = 64ll << 20;   // steps streaming less than this run persistent   // fast kernel with run-time systematic dispatch only
static int g_timing = 0;
static float g_last_bin_ms = -1.f, g_last_nll_ms = -1.f;
static bool g_init_done = false;

static int ensure_init() {
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

// A batch of binning jobs staged in pinned host memory and mirrored on the device.
struct Batch {
  BinJob* host = nullptr;          // pinned
  BinJob* dev = nullptr;
  unsigned long long* prefix_host = nullptr;   // pinned
  unsigned long long* prefix_dev = nullptr;
  EvalPdfJob* ep_host = nullptr;   // pinned
  EvalPdfJob* ep_dev = nullptr;
  int cap = 0;
  BinLaunch launch;
  long long max_points = 0;
  int n_ep = 0;

  int reserve(int n) {
    if (n <= cap) return SXMC_OK;
    release();
    SXMC_CUDA(cudaMallocHost(&host, sizeof(BinJob) * n));
    SXMC_CUDA(cudaMalloc(&dev, sizeof(BinJob) * n));
    SXMC_CUDA(cudaMallocHost(&prefix_host, sizeof(unsigned long long) * (n + 1)));
    SXMC_CUDA(cudaMalloc(&prefix_dev, sizeof(unsigned long long) * (n + 1)));
    SXMC_CUDA(cudaMallocHost(&ep_host, sizeof(EvalPdfJob) * n));
    SXMC_CUDA(cudaMalloc(&ep_dev, sizeof(EvalPdfJob) * n));
    cap = n;
    return SXMC_OK;
  }
  void release() {
    if (host) cudaFreeHost(host);
    if (dev) cudaFree(dev);
    if (prefix_host) cudaFreeHost(prefix_host);
    if (prefix_dev) cudaFree(prefix_dev);
    if (ep_host) cudaFreeHost(ep_host);
    if (ep_dev) cudaFree(ep_dev);
    host = nullptr; dev = nullptr; prefix_host = nullptr; prefix_dev = nullptr;
    ep_host = nullptr; ep_dev = nullptr; cap = 0;
  }
};

}  // namespace sxmc

using namespace sxmc;

struct sxmc_hist {
  int nfields = 0, nobs = 0;
  unsigned dataset = 0;
  std::vector<double> lower, upper;
  std::vector<int> nbins, stride;
  int total_nbins = 0;
  double bin_volume = 0;
  size_t nsamples = 0;
  float* aos_dev = nullptr;             // caller's rows until the tiled layout is built
  float* cols_dev = nullptr;            // nfields columns of col_pitch floats
  unsigned long long col_pitch = 0;     // nwtiles * WT_SAMPLES
  unsigned long long nwtiles = 0;
  unsigned* perm_dev = nullptr;         // sorted position -> original row
  bool prepared = false;
  bool sort = true;
  int primary = -1;
  unsigned* bins_dev = nullptr;
  int* read_bins_dev = nullptr;
  size_t npoints = 0;
  size_t read_bins_cap = 0;             // allocated ints (grow-only: the pointer is baked into chain tables / graphs)
  float* pdf_dev = nullptr; long pdf_offset = 0, pdf_stride = 1;
  unsigned* norm_dev = nullptr; long norm_offset = 0;
  const double* param_dev = nullptr; long param_offset = 0, param_stride = 1;
  std::vector<SystDesc> systs;
  cudaStream_t stream = nullptr;
  Batch batch;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

namespace sxmc {

static void fill_prep_geo(const sxmc_hist* h, PrepGeo& g) {
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

#define SXMC_TRY(expr) do { int _rc = (expr); if (_rc) return _rc; } while (0)
#define SXMC_INVAL(msg) do { set_error(msg); return SXMC_EINVAL; } while (0)

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
static int ensure_prepared(sxmc_hist* h) {
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
static int fill_job(sxmc_hist* h, BinJob& job, unsigned* bins, unsigned* norm, const double* params,
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

// Decide the launch shape for a filled batch of n jobs.
static void plan_launch(Batch& b, int n) {
  BinLaunch& L = b.launch;
  memset(&L, 0, sizeof L);
  L.jobs_host = b.host;
  L.jobs_dev = b.dev;
  L.prefix_dev = b.prefix_dev;
  L.njobs = n;
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
  L.smem_hist = max_bins <= g_smem_hist_max_bins;
  L.blocks_per_sm = g_blocks_per_sm;
  L.fast = n > 0 && uniform && !poly && !g_force_generic && L.nobs <= 3 &&
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
    if (!g_force_dynamic) {
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
      } else if (same && nops >= 1 && nops <= FAST_MAX_OPS && !g_no_jit &&
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

static int upload_batch(Batch& b, int n, cudaStream_t stream) {
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
  if (smem_hist_max_bins >= 0) g_smem_hist_max_bins = std::min(smem_hist_max_bins, 56 * 1024);
  g_blocks_per_sm = blocks_per_sm;
  g_force_generic = force_generic & 1;
  g_force_dynamic = (force_generic >> 1) & 1;
  g_no_mono = (force_generic >> 2) & 1;
  g_p2p_form = (force_generic >> 3) & 3;
  g_no_group = (force_generic >> 5) & 1;
  g_no_pdl = (force_generic >> 6) & 1;
  g_no_jit = (force_generic >> 7) & 1;
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
  if (bin_ms) *bin_ms = g_last_bin_ms;
  if (nll_ms) *nll_ms = g_last_nll_ms;
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
static int hist_set_eval_points_dev(sxmc_hist* h, const float* pts_dev, size_t np, cudaStream_t stream) {
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
  plan_launch(b, n);
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

// =========================================================================================
// NCCL glue (dlopen'd so that single-GPU use has no NCCL dependency)
// =========================================================================================
namespace sxmc {
typedef struct { char internal[128]; } nccl_unique_id;
typedef void* nccl_comm_t;
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(nccl_unique_id*) = nullptr;
  int (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int) = nullptr;
  int (*CommDestroy)(nccl_comm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
static int load_nccl() {
  if (g_nccl.lib) return SXMC_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* lib = nullptr;
  for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
  if (!lib) { set_error(std::string("cannot load libnccl: ") + dlerror()); return SXMC_ENCCL; }
  g_nccl.GetUniqueId = (int (*)(nccl_unique_id*))dlsym(lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(nccl_comm_t*, int, nccl_unique_id, int))dlsym(lib, "ncclCommInitRank");
  g_nccl.CommDestroy = (int (*)(nccl_comm_t))dlsym(lib, "ncclCommDestroy");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t))dlsym(lib, "ncclAllReduce");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) {
    set_error("libnccl lacks a required symbol");
    return SXMC_ENCCL;
  }
  g_nccl.lib = lib;
  return SXMC_OK;
}
static int nccl_fail(int rc, const char* what) {
  set_error(std::string("NCCL error in ") + what + ": " +
            (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?"));
  return SXMC_ENCCL;
}
constexpr int NCCL_UINT32 = 3;   // ncclUint32
constexpr int NCCL_SUM = 0;      // ncclSum
}  // namespace sxmc

struct sxmc_comm {
  sxmc::nccl_comm_t comm = nullptr;
  int nranks = 1, rank = 0;
};

struct sxmc_chain {
  int nsources = 0, nsignals = 0, nparameters = 0, nobs = 0, nsystematics = 0;
  bool systematics_fixed = true;
  std::vector<double> means, sigmas, nexpected;
  std::vector<int> fixed;
  std::vector<unsigned> n_mc;
  std::vector<short> source_id;
  std::vector<sxmc_hist*> pdfs;
  int total_nbins = 0;
  uint64_t seed = 0, stream_id = 0;
  cudaStream_t stream = nullptr;
  ChainState* st_dev = nullptr;
  ChainState* st_host = nullptr;      // pinned
  unsigned* hist_buf[2] = {nullptr, nullptr};   // [ns*total_nbins bins][ns norms]: what the lookup reads
  unsigned* part_buf[2] = {nullptr, nullptr};   // what the binning kernel writes: == hist_buf, or this rank's
                                                // partials inside its peer window (sample sharding over P2P)
  // peer-memory all-reduce (p2p_kernels.cu)
  unsigned* p2p_window = nullptr;               // [2][p2p_pitch] partials + [2][p2p_pitch] sums + [2][P2P_MAX_RANKS] flags, cudaIpc-exported
  unsigned* hist_own[2] = {nullptr, nullptr};   // the cudaMalloc'd hist_buf (hist_buf points into the window once connected)
  size_t p2p_pitch = 0;
  void* p2p_peer[P2P_MAX_RANKS] = {nullptr};    // opened peer windows (own slot = p2p_window)
  int p2p_nranks = 0, p2p_rank = 0;
  unsigned* p2p_state = nullptr;                // [0] seq, [1] ticket, [2] abort
  unsigned* p2p_state_host = nullptr;
  size_t hist_words = 0;
  Batch batch[2];
  NllSignal* sigs_dev[2] = {nullptr, nullptr};
  double* partials = nullptr;
  unsigned* ticket = nullptr;
  float* lut_dev = nullptr;
  size_t lut_cap = 0;                   // floats (grow-only)
  double* ptable_dev = nullptr;         // [nsignals][total_nbins] PDF values of the step (per-event lookup)
  bool ptable_valid = false;            // holds the values of the histograms the fixed-systematics steps read
  float* data_dev = nullptr;            // the data events, uploaded once per sxmc_chain_set_data
  size_t data_cap = 0;                  // floats (grow-only)
  std::vector<const void*> table_sig;   // what the uploaded tables and the captured graph bake: pointers and nevents
  std::vector<const void*> layout_sig;  // the part of it the monotone plan depends on (sample columns, systematics)
  size_t nevents = 0;
  int nll_grid = 1;
  bool have_data = false;
  int last_buf = 0;
  float* rows_dev = nullptr; long long row_cap = 0;
  double* nll_trace_dev = nullptr; double* prop_trace_dev = nullptr; unsigned trace_steps = 0;
  bool trace = false;
  double* tape_n_dev = nullptr; size_t n_tape_n = 0;
  double* tape_u_dev = nullptr; size_t n_tape_u = 0;
  sxmc_comm* comm = nullptr;
  int mode = 0;
  float loop_ms = 0;
  long launches = 0;
  long long nrows = 0, accepted = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  unsigned steps_done = 0, steps_planned = 0;
  cudaGraphExec_t graph_exec = nullptr;   // GRAPH_STEPS steps; rebuilt when the tables change
  long graph_launches = 0;
  void* flush_buf = nullptr; size_t flush_bytes = 0;
  double* bcast_dev = nullptr;      // persistent kernel: proposed-vector mailbox [2][MAX_PARAMETERS]
  unsigned* sync_dev = nullptr;     // persistent kernel: barrier counters, go, abort
  unsigned* sync_host = nullptr;    // pinned readback of the above
  long long step_bytes = 0;         // algorithmic bytes one binning pass streams
  bool used_persistent = false;
  // monotone one-chain path (batch_mono_kernels.cu with a batch of one)
  MonoPlan* mono_plan = nullptr;
  bool mono_checked = false;
  unsigned* mult_dev = nullptr;     // events per bin (grouped lookup), when every signal reads one dataset
  bool mult_fits_table = false;     // ... and the (signal, bin) table fits the batch kernels' shared memory
  long long* nevents_dev = nullptr;
  int last_path = 0;                // 1 graph/stream, 2 persistent, 4 monotone
  float last_comm_ms = 0;           // all-reduce time inside the last advance_timed (sum over its steps)
  unsigned long long* stamps_dev = nullptr;   // persistent kernel phase stamps (profiling hook)
  std::vector<cudaEvent_t> tev;
};

namespace sxmc {

static void chain_drop_graph(sxmc_chain* c) {
  if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
}

static int chain_build_tables(sxmc_chain* c) {
  const int ns = c->nsignals;
  c->step_bytes = 0;
  for (int b = 0; b < 2; ++b) {
    SXMC_TRY(c->batch[b].reserve(ns));
    unsigned* bins = c->hist_buf[b];
    unsigned* norms = c->hist_buf[b] + (size_t)ns * c->total_nbins;
    unsigned* pbins = c->part_buf[b];
    unsigned* pnorms = c->part_buf[b] + (size_t)ns * c->total_nbins;
    std::vector<NllSignal> sg(ns);
    for (int j = 0; j < ns; ++j) {
      sxmc_hist* h = c->pdfs[j];
      SXMC_TRY(fill_job(h, c->batch[b].host[j], pbins + (size_t)j * c->total_nbins, pnorms + j,
                        c->st_dev->proposed + c->nsources, 1));
      sg[j].bins = bins + (size_t)j * c->total_nbins;
      sg[j].norm = norms + j;
      sg[j].read_bins = h->read_bins_dev;
      sg[j].nexpected = c->nexpected[j];
      sg[j].bin_volume = h->bin_volume;
      sg[j].n_mc = c->n_mc[j];
      sg[j].total_nbins = h->total_nbins;
      sg[j].source_id = c->source_id[j];
    }
    c->batch[b].n_ep = 0;
    plan_launch(c->batch[b], ns);
    if (b == 0)
      for (int j = 0; j < ns; ++j)
        c->step_bytes += (long long)c->batch[0].host[j].nwtiles * WT_SAMPLES * 4ll *
                         std::max(1, c->batch[0].host[j].nslots);
    SXMC_TRY(upload_batch(c->batch[b], ns, c->stream));
    SXMC_CUDA(cudaMemcpyAsync(c->sigs_dev[b], sg.data(), sizeof(NllSignal) * ns, cudaMemcpyHostToDevice, c->stream));
  }
  SXMC_CUDA(cudaStreamSynchronize(c->stream));    // `sg` and the pinned staging tables may be reused after this
  return SXMC_OK;
}

// in_step: the launch follows the previous step's lookup kernel in the stream -- its prologue may
// overlap that kernel's tail (programmatic dependent launch)
static int chain_bin(sxmc_chain* c, int b, cudaEvent_t after_bin = nullptr, bool in_step = false) {
  BinLaunch BL = c->batch[b].launch;
  BL.pdl = in_step && !g_no_pdl;
  SXMC_TRY(launch_bin(BL, c->stream));
  c->launches += 1;
  if (after_bin) SXMC_CUDA(cudaEventRecord(after_bin, c->stream));
  if (c->p2p_nranks > 1) {
    P2PParams W;
    memset(&W, 0, sizeof W);
    for (int p = 0; p < c->p2p_nranks; ++p) {
      W.part[p] = (const unsigned*)c->p2p_peer[p];
      W.flags[p] = (unsigned*)c->p2p_peer[p] + 4 * c->p2p_pitch;
    }
    W.sum = c->hist_buf[b];
    W.sum_offset = (2 + (size_t)b) * c->p2p_pitch;
    W.twoshot = g_p2p_form ? (g_p2p_form == 2) : (c->p2p_nranks > 2);
    W.seq = c->p2p_state; W.ticket = c->p2p_state + 1; W.abort_flag = c->p2p_state + 2;
    W.words = c->hist_words; W.pitch = c->p2p_pitch; W.parity = b;
    W.nranks = c->p2p_nranks; W.rank = c->p2p_rank;
    SXMC_TRY(launch_hist_allreduce_p2p(W, c->stream));
    c->launches += 1;
  } else if (c->comm && c->comm->nranks > 1) {
    int rc = g_nccl.AllReduce(c->hist_buf[b], c->hist_buf[b], c->hist_words, NCCL_UINT32, NCCL_SUM,
                              c->comm->comm, c->stream);
    if (rc) return nccl_fail(rc, "ncclAllReduce");
    c->launches += 1;
  }
  return SXMC_OK;
}

static int chain_nll(sxmc_chain* c, int b, int mode, bool zero_other, float* lut) {
  NllLaunch L;
  L.sigs_dev = c->sigs_dev[b];
  L.norms_dev = c->hist_buf[b] + (size_t)c->nsignals * c->total_nbins;
  L.nsignals = c->nsignals;
  L.nparameters = c->nparameters;
  L.nsources = c->nsources;
  L.nevents = (long long)c->nevents;
  L.state = c->st_dev;
  L.partials = c->partials;
  L.ticket = c->ticket;
  L.mode = mode;
  L.lut = lut;
  L.zero_ptr = zero_other ? c->part_buf[1 - b] : nullptr;
  L.zero_words = zero_other ? c->hist_words : 0;
  L.grid = c->nll_grid;
  L.mult = c->mult_dev;
  L.mult_nbins = c->total_nbins;
  L.ptable = nullptr;
  // a step that re-binned: the lookup's prologue overlaps the binning kernel's tail; the early clear
  // of the other histogram buffer only when no peer can still be reading it
  L.pdl = mode == NLL_MODE_STEP && zero_other && !g_no_pdl;
  L.zero_early = L.pdl && !((c->comm && c->comm->nranks > 1) || c->p2p_nranks > 1);
  // events looked up one by one (signals of several datasets): eval_pdf once per (signal, bin) first
  if (c->ptable_dev && !c->mult_dev && !lut && mode == NLL_MODE_STEP && c->nevents * (size_t)c->nsignals >= 4096) {
    if (zero_other || !c->ptable_valid) {        // zero_other == the step re-binned
      SXMC_TRY(launch_pdf_table(L.sigs_dev, L.norms_dev, c->nsignals, c->total_nbins, c->ptable_dev, c->stream));
      c->launches += 1;
      c->ptable_valid = !zero_other;
    }
    L.ptable = c->ptable_dev;
  }
  SXMC_TRY(launch_nll(L, c->stream));
  c->launches += 1;
  return SXMC_OK;
}

}  // namespace sxmc

extern "C" {

int sxmc_hists_eval_host(sxmc_hist* const* hs, int n, const double* params, int nparams,
                         long param_offset, unsigned* norms_out, float* pdf_out) {
  if (n <= 0) return SXMC_OK;
  SXMC_TRY(ensure_init());
  cudaStream_t stream = hs[0]->stream;
  double* p_dev = nullptr;
  unsigned* n_dev = nullptr;
  float* pdf_dev = nullptr;
  size_t npts = hs[0]->npoints;
  for (int j = 1; j < n; ++j) npts = std::max(npts, hs[j]->npoints);
  cudaError_t e = cudaMalloc(&p_dev, sizeof(double) * std::max(1, nparams));
  if (e == cudaSuccess) e = cudaMalloc(&n_dev, sizeof(unsigned) * n);
  if (e == cudaSuccess && pdf_out && npts) e = cudaMalloc(&pdf_dev, sizeof(float) * npts * n);
  int rc = SXMC_OK;
  if (e == cudaSuccess && nparams > 0)
    e = cudaMemcpyAsync(p_dev, params, sizeof(double) * nparams, cudaMemcpyHostToDevice, stream);
  if (e == cudaSuccess) {
    for (int j = 0; j < n; ++j) {
      sxmc_hist_set_param_buffer(hs[j], p_dev, param_offset, 1);
      sxmc_hist_set_norm_buffer(hs[j], n_dev, j);
      sxmc_hist_set_pdf_buffer(hs[j], pdf_dev, (long)(j * npts), 1);
    }
    rc = sxmc_hists_eval_async(hs, n, pdf_dev != nullptr);
    if (!rc) e = cudaMemcpyAsync(norms_out, n_dev, sizeof(unsigned) * n, cudaMemcpyDeviceToHost, stream);
    if (!rc && e == cudaSuccess && pdf_dev)
      e = cudaMemcpyAsync(pdf_out, pdf_dev, sizeof(float) * npts * n, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  }
  for (int j = 0; j < n; ++j) {   // the temporaries die here: unbind them
    sxmc_hist_set_param_buffer(hs[j], nullptr, 0, 1);
    sxmc_hist_set_norm_buffer(hs[j], nullptr, 0);
    sxmc_hist_set_pdf_buffer(hs[j], nullptr, 0, 1);
  }
  cudaFree(p_dev); cudaFree(n_dev); cudaFree(pdf_dev);
  if (e != cudaSuccess) return cuda_fail(e, "sxmc_hists_eval_host", __FILE__, __LINE__);
  return rc;
}

int sxmc_nll_eval(const float* lut_dev, size_t nevents, size_t nsignals, const double* pars_dev,
                  size_t nparameters, size_t nsources, const double* means_dev,
                  const double* sigmas_dev, const double* nexpected_dev, const unsigned* n_mc_dev,
                  const short* source_id_dev, const unsigned* norms_dev, double* nll_out_host,
                  double* event_sum_out_host) {
  SXMC_TRY(ensure_init());
  double* out2 = nullptr;
  SXMC_CUDA(cudaMalloc(&out2, 2 * sizeof(double)));
  int rc = launch_nll_lut(lut_dev, nevents, nsignals, pars_dev, nparameters, nsources, means_dev,
                          sigmas_dev, nexpected_dev, n_mc_dev, source_id_dev, norms_dev, out2, 0);
  double h[2] = {0, 0};
  cudaError_t e = cudaMemcpy(h, out2, sizeof h, cudaMemcpyDeviceToHost);
  cudaFree(out2);
  if (e != cudaSuccess) return cuda_fail(e, "sxmc_nll_eval", __FILE__, __LINE__);
  if (nll_out_host) *nll_out_host = h[0];
  if (event_sum_out_host) *event_sum_out_host = h[1];
  return rc;
}

// ---------------------------------------------------------------------------------------
// MCMC
// ---------------------------------------------------------------------------------------
int sxmc_chain_create(const sxmc_chain_config* cfg, sxmc_chain** out) {
  if (!cfg || !out) SXMC_INVAL("sxmc_chain_create: NULL argument");
  *out = nullptr;
  if (cfg->nparameters <= 0 || cfg->nparameters > SXMC_MAX_PARAMETERS)
    SXMC_INVAL("sxmc_chain_create: nparameters out of range (SXMC_MAX_PARAMETERS)");
  if (cfg->nsignals <= 0 || cfg->nsignals > 128) SXMC_INVAL("sxmc_chain_create: nsignals out of range (1..128)");
  if (cfg->nsources <= 0 || cfg->nsources > cfg->nparameters) SXMC_INVAL("sxmc_chain_create: nsources out of range");
  SXMC_TRY(ensure_init());
  sxmc_chain* c = new sxmc_chain;
  c->nsources = cfg->nsources; c->nsignals = cfg->nsignals; c->nparameters = cfg->nparameters;
  c->nobs = cfg->nobservables; c->nsystematics = cfg->nsystematics;
  c->systematics_fixed = cfg->systematics_fixed != 0;
  c->means.assign(cfg->parameter_means, cfg->parameter_means + cfg->nparameters);
  c->sigmas.assign(cfg->parameter_sigma, cfg->parameter_sigma + cfg->nparameters);
  c->fixed.assign(cfg->parameter_fixed, cfg->parameter_fixed + cfg->nparameters);
  c->pdfs.assign(cfg->pdfs, cfg->pdfs + cfg->nsignals);
  c->nexpected.assign(cfg->nexpected, cfg->nexpected + cfg->nsignals);
  c->n_mc.assign(cfg->n_mc, cfg->n_mc + cfg->nsignals);
  c->source_id.assign(cfg->source_id, cfg->source_id + cfg->nsignals);
  c->total_nbins = c->pdfs[0]->total_nbins;
  for (int j = 0; j < c->nsignals; ++j) {
    if (c->pdfs[j]->total_nbins != c->total_nbins || c->pdfs[j]->nobs != c->nobs) {
      delete c;
      SXMC_INVAL("sxmc_chain_create: all signal PDFs must share the observables' binning");
    }
    if (c->source_id[j] < 0 || c->source_id[j] >= c->nsources) {
      delete c;
      SXMC_INVAL("sxmc_chain_create: source_id out of range");
    }
  }
  c->hist_words = (size_t)c->nsignals * ((size_t)c->total_nbins + 1);
  cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(&c->st_dev, sizeof(ChainState));
  if (e == cudaSuccess) e = cudaMallocHost(&c->st_host, sizeof(ChainState));
  for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
    const size_t padded = (c->hist_words + 31) & ~(size_t)31;     // whole 128-byte lines (staged peer exchange copies uint4)
    e = cudaMalloc(&c->hist_buf[b], sizeof(unsigned) * padded);
    if (e == cudaSuccess) e = cudaMemset(c->hist_buf[b], 0, sizeof(unsigned) * padded);
    c->part_buf[b] = c->hist_buf[b];
    c->hist_own[b] = c->hist_buf[b];
    if (e == cudaSuccess) e = cudaMalloc(&c->sigs_dev[b], sizeof(NllSignal) * c->nsignals);
  }
  if (e == cudaSuccess) e = cudaMalloc(&c->partials, sizeof(double) * 148 * 8);
  if (e == cudaSuccess) e = cudaMalloc(&c->ticket, sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMalloc(&c->bcast_dev, sizeof(double) * 2 * SXMC_MAX_PARAMETERS);
  if (e == cudaSuccess) e = cudaMalloc(&c->sync_dev, sizeof(unsigned) * 4);
  if (e == cudaSuccess) e = cudaMallocHost(&c->sync_host, sizeof(unsigned) * 4);
  if (e == cudaSuccess) e = cudaMemset(c->ticket, 0, sizeof(unsigned));
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev1);
  if (e != cudaSuccess) { sxmc_chain_destroy(c); return cuda_fail(e, "sxmc_chain_create", __FILE__, __LINE__); }
  *out = c;
  return SXMC_OK;
}

int sxmc_chain_destroy(sxmc_chain* c) {
  if (!c) return SXMC_OK;
  if (c->stream) cudaStreamSynchronize(c->stream);
  chain_drop_graph(c);
  cudaFree(c->flush_buf);
  cudaFree(c->bcast_dev); cudaFree(c->sync_dev); cudaFree(c->stamps_dev);
  cudaFree(c->mult_dev); cudaFree(c->nevents_dev); cudaFree(c->data_dev);
  for (int p = 0; p < c->p2p_nranks; ++p)
    if (p != c->p2p_rank && c->p2p_peer[p]) cudaIpcCloseMemHandle(c->p2p_peer[p]);
  cudaFree(c->p2p_window); cudaFree(c->p2p_state);
  if (c->p2p_state_host) cudaFreeHost(c->p2p_state_host);
  delete c->mono_plan;
  if (c->sync_host) cudaFreeHost(c->sync_host);
  for (auto& e : c->tev) cudaEventDestroy(e);
  cudaFree(c->st_dev);
  if (c->st_host) cudaFreeHost(c->st_host);
  for (int b = 0; b < 2; ++b) { cudaFree(c->hist_own[b]); cudaFree(c->sigs_dev[b]); c->batch[b].release(); }
  cudaFree(c->partials); cudaFree(c->ticket); cudaFree(c->lut_dev); cudaFree(c->rows_dev); cudaFree(c->ptable_dev);
  cudaFree(c->nll_trace_dev); cudaFree(c->prop_trace_dev); cudaFree(c->tape_n_dev); cudaFree(c->tape_u_dev);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return SXMC_OK;
}

int sxmc_chain_set_seed(sxmc_chain* c, uint64_t seed, uint64_t stream_id) {
  c->seed = seed; c->stream_id = stream_id; return SXMC_OK;
}

int sxmc_chain_set_tapes(sxmc_chain* c, const double* normals, size_t n_normals,
                         const double* uniforms, size_t n_uniforms) {
  cudaFree(c->tape_n_dev); cudaFree(c->tape_u_dev);
  c->tape_n_dev = nullptr; c->tape_u_dev = nullptr; c->n_tape_n = 0; c->n_tape_u = 0;
  if (normals && uniforms) {
    SXMC_CUDA(cudaMalloc(&c->tape_n_dev, sizeof(double) * std::max<size_t>(1, n_normals)));
    SXMC_CUDA(cudaMalloc(&c->tape_u_dev, sizeof(double) * std::max<size_t>(1, n_uniforms)));
    SXMC_CUDA(cudaMemcpy(c->tape_n_dev, normals, sizeof(double) * n_normals, cudaMemcpyHostToDevice));
    SXMC_CUDA(cudaMemcpy(c->tape_u_dev, uniforms, sizeof(double) * n_uniforms, cudaMemcpyHostToDevice));
    c->n_tape_n = n_normals; c->n_tape_u = n_uniforms;
  }
  return SXMC_OK;
}

int sxmc_chain_set_comm(sxmc_chain* c, sxmc_comm* comm) { chain_drop_graph(c); c->comm = comm; return SXMC_OK; }
int sxmc_chain_set_mode(sxmc_chain* c, int mode) { c->mode = mode; return SXMC_OK; }
int sxmc_chain_phase_stamps(sxmc_chain* c, int enable, unsigned long long* out64x8) {
  if (enable && !c->stamps_dev) {
    SXMC_CUDA(cudaMalloc(&c->stamps_dev, sizeof(unsigned long long) * 2048));
    SXMC_CUDA(cudaMemset(c->stamps_dev, 0, sizeof(unsigned long long) * 2048));
  }
  if (out64x8 && c->stamps_dev)
    SXMC_CUDA(cudaMemcpy(out64x8, c->stamps_dev, sizeof(unsigned long long) * 2048, cudaMemcpyDeviceToHost));
  if (!enable && c->stamps_dev) { cudaFree(c->stamps_dev); c->stamps_dev = nullptr; }
  return SXMC_OK;
}
int sxmc_chain_enable_trace(sxmc_chain* c, int on) { c->trace = on != 0; return SXMC_OK; }

int sxmc_chain_set_data(sxmc_chain* c, const float* data, size_t data_nfloats) {
  if (data_nfloats % (size_t)(c->nobs + 1) != 0)
    SXMC_INVAL("Number of entries in evaluation points array not divisible by number of observables.");
  c->nevents = data_nfloats / (size_t)(c->nobs + 1);
  // One upload of the data events for all PDFs, then p->SetEvalPoints(data) for every pdf
  // (src/mcmc.cpp:235) as launches on the chain's stream.  Buffers only ever grow, so that a fit
  // re-run on new data of the same size keeps its tables and its captured step graph.
  if (data_nfloats > c->data_cap || !c->data_dev) {
    cudaFree(c->data_dev);
    c->data_dev = nullptr;
    c->data_cap = 0;
    SXMC_CUDA(cudaMalloc(&c->data_dev, sizeof(float) * std::max<size_t>(1, data_nfloats)));
    c->data_cap = std::max<size_t>(1, data_nfloats);
  }
  if (data_nfloats)
    SXMC_CUDA(cudaMemcpyAsync(c->data_dev, data, sizeof(float) * data_nfloats, cudaMemcpyHostToDevice, c->stream));
  for (int j = 0; j < c->nsignals; ++j) SXMC_TRY(hist_set_eval_points_dev(c->pdfs[j], c->data_dev, c->nevents, c->stream));
  c->nll_grid = nll_grid_for((long long)c->nevents);
  const size_t lut_need = std::max<size_t>(1, c->nevents * c->nsignals);
  if (lut_need > c->lut_cap || !c->lut_dev) {
    cudaFree(c->lut_dev);
    c->lut_dev = nullptr;
    c->lut_cap = 0;
    SXMC_CUDA(cudaMalloc(&c->lut_dev, sizeof(float) * lut_need));
    c->lut_cap = lut_need;
  }
  // grouped lookup: events per bin, when all signals read one dataset
  bool shared = true;
  for (int j = 1; j < c->nsignals; ++j) shared = shared && c->pdfs[j]->dataset == c->pdfs[0]->dataset;
  c->mult_fits_table = (size_t)c->nsignals * c->total_nbins <= (size_t)BATCH_PDF_TABLE;
  if (shared && !g_no_group) {
    if (!c->mult_dev) SXMC_CUDA(cudaMalloc(&c->mult_dev, sizeof(unsigned) * (size_t)c->total_nbins));
    SXMC_TRY(launch_event_mult(c->pdfs[0]->read_bins_dev, (long long)c->nevents, c->total_nbins, c->mult_dev, c->stream));
  } else {
    if (c->mult_dev) { cudaFree(c->mult_dev); c->mult_dev = nullptr; }
    if (!c->ptable_dev)     // per-event lookup: the step's PDF values per (signal, bin)
      SXMC_CUDA(cudaMalloc(&c->ptable_dev, sizeof(double) * (size_t)c->nsignals * (size_t)c->total_nbins));
  }
  if (!c->nevents_dev) SXMC_CUDA(cudaMalloc(&c->nevents_dev, sizeof(long long)));
  const long long ne = (long long)c->nevents;
  SXMC_CUDA(cudaMemcpyAsync(c->nevents_dev, &ne, sizeof ne, cudaMemcpyHostToDevice, c->stream));
  // everything the uploaded tables and the captured graph depend on
  std::vector<const void*> sig, lsig;
  sig.push_back((const void*)(uintptr_t)c->nevents);
  sig.push_back(c->lut_dev); sig.push_back(c->mult_dev);
  sig.push_back(c->hist_buf[0]); sig.push_back(c->hist_buf[1]); sig.push_back(c->part_buf[0]); sig.push_back(c->part_buf[1]);
  lsig.push_back((const void*)(uintptr_t)(g_smem_hist_max_bins * 8 + g_force_generic * 2 + g_force_dynamic));
  for (int j = 0; j < c->nsignals; ++j) {
    sxmc_hist* h = c->pdfs[j];
    SXMC_TRY(ensure_prepared(h));
    sig.push_back(h->read_bins_dev);
    lsig.push_back(h->cols_dev);
    lsig.push_back((const void*)(uintptr_t)(h->systs.size() * 16 + (size_t)h->primary + 1));
  }
  sig.insert(sig.end(), lsig.begin(), lsig.end());
  if (sig != c->table_sig) {
    chain_drop_graph(c);   // the captured launches bake nevents and the table pointers
    SXMC_TRY(chain_build_tables(c));
    c->table_sig = sig;
  }
  if (lsig != c->layout_sig) {
    c->layout_sig = lsig;
    c->mono_checked = false;
    delete c->mono_plan;
    c->mono_plan = nullptr;
  }
  SXMC_CUDA(cudaStreamSynchronize(c->stream));   // `data` may be freed by the caller after this returns
  c->have_data = true;
  return SXMC_OK;
}

}  // extern "C"

namespace sxmc {

// Host-side initial state: the block of src/mcmc.cpp:152-228.
static void init_state_common(ChainState& s, int np, int nsources, int nsignals, const double* means,
                              const double* sigmas, const int* fixed, unsigned nsteps, float burnin_fraction,
                              int debug_mode) {
  memset(&s, 0, sizeof s);
  int nfloat = 0;
  for (int i = 0; i < np; ++i) {
    s.current[i] = means[i];
    s.proposed[i] = means[i];
    s.means[i] = means[i];
    s.sigmas[i] = sigmas[i];
    s.fixed[i] = fixed[i];
    nfloat += fixed[i] ? 0 : 1;
  }
  const float scale_factor = 2.4 * 2.4 / nfloat;   // Haario, 2001 (src/mcmc.cpp:199)
  s.scale_factor = scale_factor;
  for (int i = 0; i < np; ++i) {
    if (fixed[i]) { s.jump_width[i] = -1; continue; }
    float mean = means[i];
    float sigma = sigmas[i];
    float width = 0.1;
    if (sigma > 0) {
      width = sigma;
    } else if (i < nsignals) {     // sic: nsignals, not nsources (src/mcmc.cpp:217)
      float m = std::max(mean, (float)10);
      width = sqrt(m) / m;
    } else {
      width = sqrt(std::max(mean, (float)1));
    }
    s.jump_width[i] = 0.1 * width * scale_factor;
  }
  s.nparameters = np;
  s.nsources = nsources;
  s.nsignals = nsignals;
  s.debug_mode = debug_mode;
  s.step = 0;
  s.burnin_steps = nsteps * burnin_fraction;        // src/mcmc.cpp:152
}

static void chain_init_state(sxmc_chain* c, ChainState& s, unsigned nsteps, float burnin_fraction,
                             int debug_mode) {
  init_state_common(s, c->nparameters, c->nsources, c->nsignals, c->means.data(), c->sigmas.data(),
                    c->fixed.data(), nsteps, burnin_fraction, debug_mode);
  s.seed = c->seed;
  s.stream_id = c->stream_id;
  s.tape_normals = c->tape_n_dev; s.n_normals = c->n_tape_n;
  s.tape_uniforms = c->tape_u_dev; s.n_uniforms = c->n_tape_u;
  s.rows = c->rows_dev;
  s.row_cap = c->row_cap;
  s.nll_trace = c->trace ? c->nll_trace_dev : nullptr;
  s.prop_trace = c->trace ? c->prop_trace_dev : nullptr;
}

static int chain_push_state(sxmc_chain* c) {
  SXMC_CUDA(cudaMemcpyAsync(c->st_dev, c->st_host, sizeof(ChainState), cudaMemcpyHostToDevice, c->stream));
  return SXMC_OK;
}
static int chain_pull_state(sxmc_chain* c) {
  SXMC_CUDA(cudaMemcpyAsync(c->st_host, c->st_dev, sizeof(ChainState), cudaMemcpyDeviceToHost, c->stream));
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  return SXMC_OK;
}

// after a synchronisation: did a peer fail to show up in a peer-memory exchange?
static int chain_p2p_check(sxmc_chain* c) {
  if (c->p2p_nranks <= 1) return SXMC_OK;
  SXMC_CUDA(cudaMemcpy(c->p2p_state_host, c->p2p_state, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost));
  if (c->p2p_state_host[2]) {
    set_error("peer-memory histogram exchange timed out waiting for another rank");
    return SXMC_ESTATE;
  }
  return SXMC_OK;
}

// Fixed systematics: every step reads the initial evaluation (buffer 0), so the PDF-value table of
// the per-event lookup is computed once, outside any graph capture.
static int chain_fixed_table(sxmc_chain* c, bool rebin) {
  if (rebin || c->mult_dev || !c->ptable_dev || c->ptable_valid) return SXMC_OK;
  if (c->nevents * (size_t)c->nsignals < 4096) return SXMC_OK;
  SXMC_TRY(launch_pdf_table(c->sigs_dev[0], c->hist_buf[0] + (size_t)c->nsignals * c->total_nbins, c->nsignals,
                            c->total_nbins, c->ptable_dev, c->stream));
  c->launches += 1;
  c->ptable_valid = true;
  return SXMC_OK;
}

// one MCMC step enqueued on the chain's stream: [bin -> (allreduce)] -> lookup+nll+decide+propose
static int chain_enqueue_step(sxmc_chain* c, bool rebin, int b) {
  if (rebin) SXMC_TRY(chain_bin(c, b, nullptr, true));
  SXMC_TRY(chain_nll(c, b, NLL_MODE_STEP, rebin, nullptr));
  return SXMC_OK;
}

}  // namespace sxmc

extern "C" {

int sxmc_chain_begin(sxmc_chain* c, const float* data, size_t data_nfloats, unsigned nsteps,
                     float burnin_fraction, int debug_mode) {
  if (!c) SXMC_INVAL("sxmc_chain_begin: NULL chain");
  SXMC_TRY(sxmc_chain_set_data(c, data, data_nfloats));
  const int np = c->nparameters;
  // chain buffer: every step fits (the reference drains a sync_interval-sized buffer instead)
  if (c->row_cap < (long long)nsteps || !c->rows_dev) {
    cudaFree(c->rows_dev);
    c->rows_dev = nullptr;
    c->row_cap = std::max<long long>(1, nsteps);
    SXMC_CUDA(cudaMalloc(&c->rows_dev, sizeof(float) * (size_t)c->row_cap * (np + 1)));
  }
  if (c->trace && c->trace_steps < nsteps) {
    cudaFree(c->nll_trace_dev); cudaFree(c->prop_trace_dev);
    c->nll_trace_dev = nullptr; c->prop_trace_dev = nullptr;
    SXMC_CUDA(cudaMalloc(&c->nll_trace_dev, sizeof(double) * std::max(1u, nsteps)));
    SXMC_CUDA(cudaMalloc(&c->prop_trace_dev, sizeof(double) * std::max(1u, nsteps) * np));
    c->trace_steps = nsteps;
  }
  chain_init_state(c, *c->st_host, nsteps, burnin_fraction, debug_mode);
  SXMC_TRY(chain_push_state(c));
  c->launches = 0;
  c->loop_ms = 0;
  c->steps_done = 0;
  c->steps_planned = nsteps;

  // initial evaluation at the means (src/mcmc.cpp:233-252): always bins, whatever is fixed
  for (int b = 0; b < 2; ++b) {
    SXMC_CUDA(cudaMemsetAsync(c->hist_buf[b], 0, sizeof(unsigned) * c->hist_words, c->stream));
    if (c->part_buf[b] != c->hist_buf[b])
      SXMC_CUDA(cudaMemsetAsync(c->part_buf[b], 0, sizeof(unsigned) * c->hist_words, c->stream));
  }
  SXMC_TRY(chain_bin(c, 0));
  SXMC_TRY(chain_nll(c, 0, NLL_MODE_INIT, false, nullptr));
  c->ptable_valid = false;
  c->last_buf = 0;
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  return chain_p2p_check(c);
}

// Advance the chain by nsteps (the body of the loop at src/mcmc.cpp:261-378), device-resident.
int sxmc_chain_advance(sxmc_chain* c, unsigned nsteps) {
  if (!c || !c->have_data) { set_error("sxmc_chain_advance: sxmc_chain_begin was not called"); return SXMC_ESTATE; }
  if ((unsigned long long)c->steps_done + nsteps > c->steps_planned) {
    set_error("sxmc_chain_advance: more steps than sxmc_chain_begin announced");
    return SXMC_ESTATE;
  }
  const bool rebin = c->nsystematics > 0 && !c->systematics_fixed;   // src/mcmc.cpp:264
  const bool has_comm = (c->comm && c->comm->nranks > 1) || c->p2p_nranks > 1;
  SXMC_TRY(chain_fixed_table(c, rebin));
  const unsigned GRAPH_STEPS = 16;   // even: the histogram double-buffer parity is preserved
  unsigned left = nsteps;
  // Small fits: the whole segment in one cooperative launch (chain_kernels.cu).  Chosen when a
  // step is too short for per-kernel launches to be hidden (the samples fit in L2).
  c->used_persistent = false;
  const BinLaunch& BL = c->batch[0].launch;
  const bool persistent_ok = BL.fast && BL.njobs <= FAST_MAX_JOBS && !has_comm && c->nsignals <= 128 &&
                             c->pdfs.size() && nsteps > 0;
  // Monotone fits (one observable, shift / scale / ctscale only): the edge-search kernel of
  // batch_mono_kernels.cu with a batch of one chain.
  if ((c->mode == 0 || c->mode == 4) && !has_comm && nsteps > 0 && c->pdfs.size()) {
    if (!c->mono_checked) {
      c->mono_checked = true;
      bool mono = BL.fast && BL.nobs == 1 && BL.nslots == 1;
      for (int j = 0; j < c->nsignals && mono; ++j) {
        sxmc_hist* h = c->pdfs[j];
        if (!h->sort || !h->prepared || h->primary != 0) mono = false;
        for (const SystDesc& d : h->systs)
          if (d.type == SXMC_SYST_RESOLUTION_SCALE || d.obs != 0 || d.npars > 1) mono = false;
        bool bad = false;
        if (mono) SXMC_TRY(column_has_nan_or_neginf(h->cols_dev, h->col_pitch, &bad, c->stream));
        if (bad) mono = false;
      }
      if (mono) {
        std::vector<unsigned long long> nsamp(c->nsignals);
        for (int j = 0; j < c->nsignals; ++j) nsamp[j] = c->pdfs[j]->nsamples;
        c->mono_plan = new MonoPlan;
        if (plan_chain_batch_mono(BL, nsamp.data(), 1, c->mono_plan) != SXMC_OK) { delete c->mono_plan; c->mono_plan = nullptr; }
      }
    }
    if (c->mono_plan && !g_no_mono) {
      BatchParams Q;
      memset(&Q, 0, sizeof Q);
      Q.st = c->st_dev; Q.sigs = c->sigs_dev[0]; Q.hist = c->hist_buf[0]; Q.hist_words = c->hist_words;
      Q.nevents = c->nevents_dev; Q.sync = c->sync_dev;
      Q.nchains = 1; Q.ns = c->nsignals; Q.np = c->nparameters; Q.nsources = c->nsources;
      Q.nsteps = nsteps; Q.do_init = 0; Q.rebin = rebin ? 1 : 0;
      Q.mult = c->mult_fits_table ? c->mult_dev : nullptr; Q.mono = 1;
      if (rebin) SXMC_CUDA(cudaMemsetAsync(c->hist_buf[0], 0, sizeof(unsigned) * c->hist_words, c->stream));
      else if (c->last_buf != 0) {   // fixed systematics: the lookups read the initial evaluation
        SXMC_CUDA(cudaMemcpyAsync(c->hist_buf[0], c->hist_buf[c->last_buf], sizeof(unsigned) * c->hist_words,
                                  cudaMemcpyDeviceToDevice, c->stream));
      }
      SXMC_CUDA(cudaMemsetAsync(c->sync_dev, 0, sizeof(unsigned) * 4, c->stream));
      SXMC_CUDA(cudaEventRecord(c->ev0, c->stream));
      SXMC_TRY(launch_chain_batch_mono(*c->mono_plan, Q, c->stream));
      SXMC_CUDA(cudaEventRecord(c->ev1, c->stream));
      SXMC_CUDA(cudaMemcpyAsync(c->sync_host, c->sync_dev, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost, c->stream));
      SXMC_CUDA(cudaStreamSynchronize(c->stream));
      if (c->sync_host[3]) { set_error("monotone chain kernel timed out at a grid barrier"); return SXMC_ESTATE; }
      float ms = 0;
      SXMC_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
      c->loop_ms += ms;
      c->launches += 1;
      c->steps_done += nsteps;
      c->last_buf = 0;
      c->last_path = 4;
      return SXMC_OK;
    }
    if (c->mode == 4) { set_error("sxmc_chain_set_mode(4): the fit is not monotone (or does not fit on chip)"); return SXMC_ESTATE; }
  }
  const bool want_persistent = c->mode == 2 || (c->mode == 0 && c->step_bytes <= g_persistent_max_bytes);
  if (persistent_ok && want_persistent) {
    PersistParams Q;
    memset(&Q, 0, sizeof Q);
    Q.st = c->st_dev;
    Q.sigs = c->sigs_dev[0];
    Q.hist_buf[0] = c->hist_buf[0];
    Q.hist_buf[1] = c->hist_buf[1];
    Q.hist_words = c->hist_words;
    Q.partials = c->partials;
    Q.bcast = c->bcast_dev;
    Q.sync = c->sync_dev;
    Q.nevents = (long long)c->nevents;
    Q.ns = c->nsignals; Q.np = c->nparameters; Q.nsources = c->nsources;
    Q.nsteps = nsteps;
    Q.step0 = c->steps_done;
    Q.rebin = rebin ? 1 : 0;
    Q.stamps = c->stamps_dev;
    SXMC_CUDA(cudaMemsetAsync(c->sync_dev, 0, sizeof(unsigned) * 4, c->stream));
    SXMC_CUDA(cudaEventRecord(c->ev0, c->stream));
    int rc = launch_chain_persistent(BL, Q, c->stream);
    if (rc == SXMC_OK) {
      SXMC_CUDA(cudaEventRecord(c->ev1, c->stream));
      SXMC_CUDA(cudaMemcpyAsync(c->sync_host, c->sync_dev, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost, c->stream));
      SXMC_CUDA(cudaStreamSynchronize(c->stream));
      if (c->sync_host[3]) { set_error("persistent chain kernel timed out at a grid barrier"); return SXMC_ESTATE; }
      float ms = 0;
      SXMC_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
      c->loop_ms += ms;
      c->launches += 1;
      c->steps_done += nsteps;
      if (rebin && c->steps_done > 0) c->last_buf = (int)(c->steps_done & 1);
      c->used_persistent = true;
      c->last_path = 2;
      return SXMC_OK;
    }
    if (c->mode == 2) return rc;     // explicitly requested: report why it cannot run
    cudaGetLastError();              // automatic: fall through to the graph path
  }
  auto plain_step = [&]() -> int {
    const unsigned i = c->steps_done;
    int rc = chain_enqueue_step(c, rebin, rebin ? (int)((i + 1) & 1) : 0);
    if (!rc) { c->steps_done += 1; left -= 1; }
    return rc;
  };
  if ((c->mode == 0 || c->mode == 3) && left >= 2 * GRAPH_STEPS && !c->graph_exec) {
    // CUDA graph of GRAPH_STEPS consecutive steps starting at an even step index
    cudaGraph_t graph = nullptr;
    const long before = c->launches;
    SXMC_CUDA(cudaStreamBeginCapture(c->stream, has_comm ? cudaStreamCaptureModeRelaxed
                                                          : cudaStreamCaptureModeThreadLocal));
    int rc = SXMC_OK;
    for (unsigned k = 0; k < GRAPH_STEPS && !rc; ++k) rc = chain_enqueue_step(c, rebin, rebin ? (int)((k + 1) & 1) : 0);
    cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    c->graph_launches = c->launches - before;
    c->launches = before;
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamEndCapture", __FILE__, __LINE__);
    e = cudaGraphInstantiate(&c->graph_exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { c->graph_exec = nullptr; return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__); }
  }
  SXMC_CUDA(cudaEventRecord(c->ev0, c->stream));
  if ((c->mode == 0 || c->mode == 3) && c->graph_exec) {
    if ((c->steps_done & 1u) && left > 0) SXMC_TRY(plain_step());
    while (left >= GRAPH_STEPS) {
      SXMC_CUDA(cudaGraphLaunch(c->graph_exec, c->stream));
      c->launches += c->graph_launches;
      c->steps_done += GRAPH_STEPS;
      left -= GRAPH_STEPS;
    }
  }
  while (left > 0) SXMC_TRY(plain_step());
  if (rebin && c->steps_done > 0) c->last_buf = (int)(c->steps_done & 1);   // step n-1 used buffer n&1
  SXMC_CUDA(cudaEventRecord(c->ev1, c->stream));
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  float ms = 0;
  SXMC_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  c->loop_ms += ms;
  c->last_path = 1;
  return chain_p2p_check(c);
}

// Same steps as sxmc_chain_advance, issued as plain stream launches with a CUDA event around
// every kernel, and (flush_bytes > 0) an L2 flush -- a memset of flush_bytes -- before every
// step, outside the timed intervals.  Measurement only.
int sxmc_chain_advance_timed(sxmc_chain* c, unsigned nsteps, size_t flush_bytes, float* step_ms_sum,
                             float* bin_ms_sum, float* nll_ms_sum) {
  if (!c || !c->have_data) { set_error("sxmc_chain_advance_timed: sxmc_chain_begin was not called"); return SXMC_ESTATE; }
  if ((unsigned long long)c->steps_done + nsteps > c->steps_planned) {
    set_error("sxmc_chain_advance_timed: more steps than sxmc_chain_begin announced");
    return SXMC_ESTATE;
  }
  const bool rebin = c->nsystematics > 0 && !c->systematics_fixed;
  SXMC_TRY(chain_fixed_table(c, rebin));
  if (flush_bytes > c->flush_bytes) {
    cudaFree(c->flush_buf);
    c->flush_buf = nullptr;
    c->flush_bytes = 0;
    SXMC_CUDA(cudaMalloc(&c->flush_buf, flush_bytes));
    c->flush_bytes = flush_bytes;
  }
  const unsigned CHUNK = 64;
  if (c->tev.empty()) {
    c->tev.resize(4 * CHUNK);
    for (auto& e : c->tev) SXMC_CUDA(cudaEventCreate(&e));
  }
  double t_step = 0, t_bin = 0, t_nll = 0, t_comm = 0;
  unsigned left = nsteps;
  while (left > 0) {
    const unsigned n = std::min(left, CHUNK);
    for (unsigned k = 0; k < n; ++k) {
      const unsigned i = c->steps_done;
      const int b = rebin ? (int)((i + 1) & 1) : 0;
      if (flush_bytes) SXMC_CUDA(cudaMemsetAsync(c->flush_buf, (int)(i & 0xff), flush_bytes, c->stream));
      SXMC_CUDA(cudaEventRecord(c->tev[3 * k], c->stream));
      if (rebin) SXMC_TRY(chain_bin(c, b, c->tev[3 * CHUNK + k], true));
      else SXMC_CUDA(cudaEventRecord(c->tev[3 * CHUNK + k], c->stream));
      SXMC_CUDA(cudaEventRecord(c->tev[3 * k + 1], c->stream));
      SXMC_TRY(chain_nll(c, b, NLL_MODE_STEP, rebin, nullptr));
      SXMC_CUDA(cudaEventRecord(c->tev[3 * k + 2], c->stream));
      c->steps_done += 1;
    }
    SXMC_CUDA(cudaStreamSynchronize(c->stream));
    for (unsigned k = 0; k < n; ++k) {
      float a = 0, b2 = 0;
      SXMC_CUDA(cudaEventElapsedTime(&a, c->tev[3 * k], c->tev[3 * k + 1]));
      SXMC_CUDA(cudaEventElapsedTime(&b2, c->tev[3 * k + 1], c->tev[3 * k + 2]));
      float cm = 0;     // binning kernel end -> all-reduce end (0 without a communicator)
      SXMC_CUDA(cudaEventElapsedTime(&cm, c->tev[3 * CHUNK + k], c->tev[3 * k + 1]));
      t_comm += cm;
      t_bin += a - cm; t_nll += b2; t_step += a + b2;      // bin = the binning launch alone
    }
    left -= n;
  }
  if (rebin && c->steps_done > 0) c->last_buf = (int)(c->steps_done & 1);
  c->loop_ms += (float)t_step;
  if (step_ms_sum) *step_ms_sum = (float)t_step;
  if (bin_ms_sum) *bin_ms_sum = (float)t_bin;
  if (nll_ms_sum) *nll_ms_sum = (float)t_nll;
  c->last_comm_ms = (float)t_comm;
  return chain_p2p_check(c);
}

int sxmc_chain_end(sxmc_chain* c, long* nrows_out) {
  if (!c) SXMC_INVAL("sxmc_chain_end: NULL chain");
  SXMC_TRY(chain_pull_state(c));
  c->nrows = c->st_host->nrows;
  c->accepted = c->st_host->accepted;
  if (nrows_out) *nrows_out = (long)std::min(c->nrows, c->row_cap);
  return SXMC_OK;
}

int sxmc_chain_run(sxmc_chain* c, const float* data, size_t data_nfloats, unsigned nsteps,
                   float burnin_fraction, int debug_mode, unsigned sync_interval, long* nrows_out) {
  (void)sync_interval;
  if (!c) SXMC_INVAL("sxmc_chain_run: NULL chain");
  SXMC_TRY(sxmc_chain_begin(c, data, data_nfloats, nsteps, burnin_fraction, debug_mode));
  SXMC_TRY(sxmc_chain_advance(c, nsteps));
  return sxmc_chain_end(c, nrows_out);
}

int sxmc_chain_read(sxmc_chain* c, float* rows_out, long max_rows) {
  long long n = std::min<long long>(std::min(c->nrows, c->row_cap), max_rows);
  if (n > 0)
    SXMC_CUDA(cudaMemcpy(rows_out, c->rows_dev, sizeof(float) * (size_t)n * (c->nparameters + 1),
                         cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_chain_accepted(sxmc_chain* c, long* accepted_out) { *accepted_out = (long)c->accepted; return SXMC_OK; }
int sxmc_chain_loop_ms(sxmc_chain* c, float* ms_out) { *ms_out = c->loop_ms; return SXMC_OK; }
int sxmc_chain_launches(sxmc_chain* c, long* launches_out) { *launches_out = c->launches; return SXMC_OK; }
int sxmc_chain_last_path(sxmc_chain* c, int* path_out) { *path_out = c->last_path; return SXMC_OK; }
int sxmc_chain_last_comm_ms(sxmc_chain* c, float* ms_out) { *ms_out = c->last_comm_ms; return SXMC_OK; }
int sxmc_chain_bin_kernel_kind(sxmc_chain* c, int* kind_out) {
  if (!c || !kind_out) SXMC_INVAL("sxmc_chain_bin_kernel_kind: NULL argument");
  if (!c->have_data) { set_error("sxmc_chain_bin_kernel_kind: no data set (the launch is planned with the data)"); return SXMC_ESTATE; }
  const BinLaunch& L = c->batch[0].launch;
  *kind_out = !L.fast ? 0 : (L.prog_nops < 0 ? 1 : (L.jit ? 3 : 2));
  return SXMC_OK;
}

int sxmc_chain_read_trace(sxmc_chain* c, double* nll_trace_out, double* proposed_out, unsigned nsteps) {
  if (!c->nll_trace_dev || nsteps > c->trace_steps) { set_error("trace was not enabled for that many steps"); return SXMC_ESTATE; }
  if (nll_trace_out) SXMC_CUDA(cudaMemcpy(nll_trace_out, c->nll_trace_dev, sizeof(double) * nsteps, cudaMemcpyDeviceToHost));
  if (proposed_out)
    SXMC_CUDA(cudaMemcpy(proposed_out, c->prop_trace_dev, sizeof(double) * nsteps * c->nparameters, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}

int sxmc_chain_nll_at(sxmc_chain* c, const double* pars, int rebin, double* nll_out, double* event_sum_out) {
  if (!c->have_data) { set_error("sxmc_chain_nll_at: no data set"); return SXMC_ESTATE; }
  if (!c->st_host->nparameters) {   // never run: build a neutral state
    chain_init_state(c, *c->st_host, 0, 0.f, 0);
    SXMC_TRY(chain_push_state(c));
    rebin = 1;
  }
  SXMC_CUDA(cudaMemcpyAsync(c->st_dev->proposed, pars, sizeof(double) * c->nparameters, cudaMemcpyHostToDevice, c->stream));
  const int b = c->last_buf;
  cudaEvent_t e0 = c->ev0, e1 = c->ev1;
  if (rebin) {
    SXMC_CUDA(cudaMemsetAsync(c->part_buf[b], 0, sizeof(unsigned) * c->hist_words, c->stream));
    if (g_timing) SXMC_CUDA(cudaEventRecord(e0, c->stream));
    SXMC_TRY(chain_bin(c, b));
    if (g_timing) {
      SXMC_CUDA(cudaEventRecord(e1, c->stream));
      SXMC_CUDA(cudaEventSynchronize(e1));
      SXMC_CUDA(cudaEventElapsedTime(&g_last_bin_ms, e0, e1));
    }
  }
  if (g_timing) SXMC_CUDA(cudaEventRecord(e0, c->stream));
  SXMC_TRY(chain_nll(c, b, NLL_MODE_EVAL_ONLY, false, c->lut_dev));
  if (g_timing) {
    SXMC_CUDA(cudaEventRecord(e1, c->stream));
    SXMC_CUDA(cudaEventSynchronize(e1));
    SXMC_CUDA(cudaEventElapsedTime(&g_last_nll_ms, e0, e1));
  }
  double out[2];
  SXMC_CUDA(cudaMemcpyAsync(out, &c->st_dev->last_nll, sizeof out, cudaMemcpyDeviceToHost, c->stream));
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  if (nll_out) *nll_out = out[0];
  if (event_sum_out) *event_sum_out = out[1];
  return SXMC_OK;
}

int sxmc_chain_get_norms(sxmc_chain* c, unsigned* norms_out) {
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(norms_out, c->hist_buf[c->last_buf] + (size_t)c->nsignals * c->total_nbins,
                       sizeof(unsigned) * c->nsignals, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_chain_get_lut(sxmc_chain* c, float* lut_out) {
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(lut_out, c->lut_dev, sizeof(float) * c->nevents * c->nsignals, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_chain_get_bins(sxmc_chain* c, unsigned* bins_out) {
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(bins_out, c->hist_buf[c->last_buf], sizeof(unsigned) * (size_t)c->nsignals * c->total_nbins,
                       cudaMemcpyDeviceToHost));
  return SXMC_OK;
}

// ---------------------------------------------------------------------------------------
// multi-GPU
// ---------------------------------------------------------------------------------------
int sxmc_comm_unique_id(void* id128_out) {
  SXMC_TRY(load_nccl());
  nccl_unique_id id;
  int rc = g_nccl.GetUniqueId(&id);
  if (rc) return nccl_fail(rc, "ncclGetUniqueId");
  memcpy(id128_out, &id, sizeof id);
  return SXMC_OK;
}
int sxmc_comm_create(const void* id128, int nranks, int rank, sxmc_comm** out) {
  SXMC_TRY(ensure_init());
  SXMC_TRY(load_nccl());
  nccl_unique_id id;
  memcpy(&id, id128, sizeof id);
  sxmc_comm* c = new sxmc_comm;
  c->nranks = nranks; c->rank = rank;
  int rc = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  if (rc) { delete c; return nccl_fail(rc, "ncclCommInitRank"); }
  *out = c;
  return SXMC_OK;
}
int sxmc_comm_destroy(sxmc_comm* comm) {
  if (!comm) return SXMC_OK;
  if (comm->comm) g_nccl.CommDestroy(comm->comm);
  delete comm;
  return SXMC_OK;
}
int sxmc_comm_allreduce_u32(sxmc_comm* comm, unsigned* dev, size_t count) {
  int rc = g_nccl.AllReduce(dev, dev, count, NCCL_UINT32, NCCL_SUM, comm->comm, 0);
  if (rc) return nccl_fail(rc, "ncclAllReduce");
  SXMC_CUDA(cudaStreamSynchronize(0));
  return SXMC_OK;
}

}  // extern "C"

// =========================================================================================
// Ensemble pieces: fake-data generator and the Contour estimator (ensemble_kernels.cu)
// =========================================================================================
struct sxmc_generator {
  int n = 0, nobs = 0, total_nbins = 0;
  std::vector<sxmc_hist*> hs;
  unsigned* bins_dev = nullptr;     // [n][total_nbins] histograms at the systematics' means
  unsigned* cum_dev = nullptr;      // [n][total_nbins] inclusive prefix sums
  unsigned* norms_dev = nullptr;    // [n]
  double* params_dev = nullptr;
  GenSignal* sigs_dev = nullptr;
  unsigned long long* counts_dev = nullptr;    // [n+1]
  unsigned long long* offsets_dev = nullptr;   // [n+1]
  unsigned long long* counts_host = nullptr;   // pinned [n+1]
  float* events_dev = nullptr; unsigned long long cap_events = 0;
  GenGeo geo;
  std::vector<unsigned> norms, n_mc;
  std::vector<double> nexpected;
  Batch batch;
  cudaStream_t stream = nullptr;
  unsigned long long last_total = 0;
};

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
    plan_launch(g->batch, n);
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
  if (!rc) { b->jobs.n_ep = 0; plan_launch(b->jobs, b->nsignals); }
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
  bool use_mono = b->mono && !g_no_mono && !b->mono_plan_failed;
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
  Q.mono = (b->mono && !g_no_mono) ? 1 : 0;
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


// =========================================================================================
// Sample sharding over NVLink peer memory: the chain's partial histograms live in a window the
// other ranks open with CUDA IPC; the per-step exchange is hist_allreduce_p2p_kernel.
// =========================================================================================
extern "C" {

int sxmc_chain_p2p_export(sxmc_chain* c, void* handle64_out) {
  if (!c || !handle64_out) SXMC_INVAL("sxmc_chain_p2p_export: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  if (!c->p2p_window) {
    c->p2p_pitch = (c->hist_words + 31) & ~(size_t)31;           // 128-byte aligned parities
    const size_t words = 4 * c->p2p_pitch + 2 * P2P_MAX_RANKS;
    SXMC_CUDA(cudaMalloc(&c->p2p_window, sizeof(unsigned) * words));
    SXMC_CUDA(cudaMemset(c->p2p_window, 0, sizeof(unsigned) * words));
    SXMC_CUDA(cudaMalloc(&c->p2p_state, sizeof(unsigned) * 16));      // + 4 x u64 phase clocks
    SXMC_CUDA(cudaMemset(c->p2p_state, 0, sizeof(unsigned) * 16));
    SXMC_CUDA(cudaMallocHost(&c->p2p_state_host, sizeof(unsigned) * 16));
  }
  cudaIpcMemHandle_t h;
  SXMC_CUDA(cudaIpcGetMemHandle(&h, c->p2p_window));
  memcpy(handle64_out, &h, 64);
  return SXMC_OK;
}

int sxmc_chain_p2p_connect(sxmc_chain* c, const void* handles, int nranks, int rank) {
  if (!c || !handles) SXMC_INVAL("sxmc_chain_p2p_connect: NULL argument");
  if (nranks < 1 || nranks > P2P_MAX_RANKS || rank < 0 || rank >= nranks)
    SXMC_INVAL("sxmc_chain_p2p_connect: 1..8 ranks of one box");
  if (!c->p2p_window) { set_error("sxmc_chain_p2p_connect: call sxmc_chain_p2p_export first"); return SXMC_ESTATE; }
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  chain_drop_graph(c);
  for (int p = 0; p < nranks; ++p) {
    if (p == rank) { c->p2p_peer[p] = c->p2p_window; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + 64 * (size_t)p, 64);
    SXMC_CUDA(cudaIpcOpenMemHandle(&c->p2p_peer[p], h, cudaIpcMemLazyEnablePeerAccess));
  }
  c->p2p_nranks = nranks;
  c->p2p_rank = rank;
  c->part_buf[0] = c->p2p_window;
  c->part_buf[1] = c->p2p_window + c->p2p_pitch;
  c->hist_buf[0] = c->p2p_window + 2 * c->p2p_pitch;      // peers store their slices of the sum here
  c->hist_buf[1] = c->p2p_window + 3 * c->p2p_pitch;
  c->table_sig.clear();
  if (c->have_data) SXMC_TRY(chain_build_tables(c));
  return SXMC_OK;
}

int sxmc_chain_p2p_status(sxmc_chain* c, unsigned* exchanges_out, int* aborted_out) {
  if (!c || !c->p2p_state) { if (exchanges_out) *exchanges_out = 0; if (aborted_out) *aborted_out = 0; return SXMC_OK; }
  SXMC_CUDA(cudaMemcpy(c->p2p_state_host, c->p2p_state, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost));
  if (exchanges_out) *exchanges_out = c->p2p_state_host[0];
  if (aborted_out) *aborted_out = (int)c->p2p_state_host[2];
  return SXMC_OK;
}

int sxmc_chain_p2p_clocks(sxmc_chain* c, double* ns_out3) {
  if (!c || !ns_out3) SXMC_INVAL("sxmc_chain_p2p_clocks: NULL argument");
  ns_out3[0] = ns_out3[1] = ns_out3[2] = 0;
  if (!c->p2p_state) return SXMC_OK;
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(c->p2p_state_host, c->p2p_state, sizeof(unsigned) * 16, cudaMemcpyDeviceToHost));
  const unsigned long long* clk = reinterpret_cast<const unsigned long long*>(c->p2p_state_host + 4);
  for (int i = 0; i < 3; ++i) ns_out3[i] = (double)clk[1 + i];
  return SXMC_OK;
}

}  // extern "C"
