// capi_internal.h -- what the translation units of the extern "C" layer share (capi_hist.cu: library,
// buffers, pdfz::EvalHist handles; capi_chain.cu: MCMC chains, NCCL glue, peer-memory windows;
// capi_ensemble.cu: fake-data generator, contour reductions, chain batches): the handle structs behind
// the opaque pointers of include/sxmc_b200.h, the kernel-selection knobs and the launch planning.
#pragma once
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

// Kernel-selection and measurement knobs (sxmc_set_bin_tuning).  The process-wide instance holds the
// DEFAULTS: every handle copies it when it is created, so two fits in one process can be tuned
// independently (set, create, set differently, create).
struct Tuning {
  int smem_hist_max_bins = 8192;   // histograms up to this many bins are privatised in shared memory
  int blocks_per_sm = 0;           // generic kernel: <= 0 occupancy maximum
  int force_generic = 0;           // the per-sample interpreter for everything
  int force_dynamic = 0;           // fast kernel with run-time systematic dispatch only
  int p2p_form = 0;                // peer-memory exchange: 0 automatic (fused into the rank-sharded lookup where the fit allows it, else two-shot above 2 ranks), 1 one-shot all-reduce kernel, 2 two-shot
  int no_mono = 0;                 // never take the monotone endpoint path (tests)
  int no_jit = 0;                  // programs outside the build's table keep the run-time dispatching kernel
  int cta_split = 0;               // measurement knob of the fast binning kernel: 0 production (chunks dealt to a CTA's warps round robin; global-memory histograms: warps independent of each other), 1 one contiguous range per warp inside per-CTA segments (the round-1 form), 2 global histograms: one contiguous range per warp of the grid, 3 global histograms: per-CTA segments, round robin
  int chunk_form = 0;              // fast binning pipeline: 0 automatic (four-row chunks for launches of >= 1 GiB into global-memory histograms of <= 65536 bins), 1 always two-row chunks, 2 four-row chunks wherever they fit (measurement knob)
  int no_pdl = 0;                  // steps without programmatic dependent launch (measurement knob)
  int no_group = 0;                // lookup: never group the events by bin (the multi-dataset form, on request)
  long long persistent_max_bytes = 64ll << 20;   // steps streaming less than this run persistent
};
extern Tuning g_tune;
extern int g_timing;               // sxmc_enable_kernel_timing (process-wide switch; results live on the handles)

typedef void* nccl_comm_t;         // ncclComm_t (NCCL is dlopen'd: capi_chain.cu)

#define SXMC_TRY(expr) do { int _rc = (expr); if (_rc) return _rc; } while (0)
#define SXMC_INVAL(msg) do { ::sxmc::set_error(msg); return SXMC_EINVAL; } while (0)

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

using namespace sxmc;   // (internal header: the handle structs below live at global scope, as the C ABI names them)

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
  Tuning tune;                          // kernel selection knobs, as they stood when the handle was created
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

struct sxmc_comm {
  sxmc::nccl_comm_t comm = nullptr;
  int nranks = 1, rank = 0;
};

struct sxmc_chain {
  Tuning tune;                          // kernel selection knobs, as they stood when the chain was created
  float last_bin_ms = -1.f, last_nll_ms = -1.f;   // sxmc_enable_kernel_timing: the last timed evaluate-only pass
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
  unsigned long long p2p_timeout_ns = 60ull * 1000000000ull;   // SXMC_P2P_TIMEOUT_S (default 60 s)
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


namespace sxmc {
int ensure_init();
void fill_prep_geo(const sxmc_hist* h, PrepGeo& g);
int ensure_prepared(sxmc_hist* h);
int fill_job(sxmc_hist* h, BinJob& job, unsigned* bins, unsigned* norm, const double* params, long long param_stride);
void plan_launch(Batch& b, int n, const Tuning& T);
int upload_batch(Batch& b, int n, cudaStream_t stream);
int hist_set_eval_points_dev(sxmc_hist* h, const float* pts_dev, size_t np, cudaStream_t stream);
void init_state_common(ChainState& s, int np, int nsources, int nsignals, const double* means, const double* sigmas,
                       const int* fixed, unsigned nsteps, float burnin_fraction, int debug_mode);
}  // namespace sxmc
