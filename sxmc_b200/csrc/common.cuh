// common.cuh -- shared device-side structures of libsxmc_b200 (sm_100a only).
#pragma once
#ifdef __CUDACC_RTC__
// NVRTC translation unit (jit.cu compiles bin_fast_kernel for a fit's own systematic program at run
// time): no host headers there.  The limits of include/sxmc_b200.h restated -- the normal build
// below checks that the two stay in step.
#define SXMC_SYST_SHIFT 0
#define SXMC_SYST_SCALE 1
#define SXMC_SYST_RESOLUTION_SCALE 2
#define SXMC_SYST_CTSCALE 3
#define SXMC_MAX_NFIELDS 10
#define SXMC_MAX_SYST 8
#define SXMC_MAX_NPARS 4
#define SXMC_MAX_PARAMETERS 256
#else
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

#include "../../include/sxmc_b200.h"
static_assert(SXMC_SYST_SHIFT == 0 && SXMC_SYST_SCALE == 1 && SXMC_SYST_RESOLUTION_SCALE == 2 && SXMC_SYST_CTSCALE == 3 &&
                  SXMC_MAX_NFIELDS == 10 && SXMC_MAX_SYST == 8 && SXMC_MAX_NPARS == 4 && SXMC_MAX_PARAMETERS == 256,
              "common.cuh: the NVRTC block above restates these");
#endif

namespace sxmc {

// ---- sample layout -------------------------------------------------------------------
// Samples of one PDF are stored as one float column per field (SoA).  Within a column,
// samples are grouped in warp-tiles of WT_SAMPLES = 32 lanes x LANE_RUN samples.  In the
// (bin-sorted) sample order, lane l of a warp-tile owns the LANE_RUN *consecutive* samples
// [l*LANE_RUN, (l+1)*LANE_RUN); in memory the tile is LANE_RUN/4 rows of 32 float4, row q
// holding samples 4q..4q+3 of every lane, so that one LDG.128 per lane reads 512
// contiguous bytes per warp while each thread walks a sorted run (its bin changes rarely,
// which is what makes the per-thread run-length aggregation of the histogram updates work).
constexpr int LANE_RUN = 32;
constexpr int WT_SAMPLES = 32 * LANE_RUN;       // 1024 samples per warp-tile
constexpr int WT_ROWS = LANE_RUN / 4;           // float4 rows per warp-tile
constexpr int BIN_THREADS = 256;
constexpr int BIN_WARPS = BIN_THREADS / 32;     // warp-tiles per block-tile

#ifndef SXMC_FAST_WARPS
#define SXMC_FAST_WARPS 16
#endif
constexpr int FAST_WARPS = SXMC_FAST_WARPS;     // warps of the persistent fast binning CTA
constexpr int FAST_THREADS = FAST_WARPS * 32;
// sorted position s -> offset inside a tiled column
__host__ __device__ __forceinline__ unsigned long long tiled_offset(unsigned long long s) {
  const unsigned long long wt = s / WT_SAMPLES;
  const unsigned r = (unsigned)(s % WT_SAMPLES);
  const unsigned lane = r / LANE_RUN, i = r % LANE_RUN;
  return ((wt * WT_ROWS + (i >> 2)) * 32ull + lane) * 4ull + (i & 3);
}

constexpr int FAST_MAX_STAGES = 8;              // bulk-copy pipeline stages per warp

constexpr int MAX_SLOTS = 6;                    // columns the fast kernel keeps in registers

struct SystDesc {
  short type;        // SXMC_SYST_*
  short obs;         // FIELD index of the observable it moves
  short extra;       // FIELD index of the truth value (RESOLUTION_SCALE)
  short npars;
  short pidx[SXMC_MAX_NPARS];   // global coefficient indices
  short obs_slot;    // register slot of obs / extra in the fast kernel
  short extra_slot;
};

// One PDF's binning work.  Lives in device memory; read with uniform loads.
struct BinJob {
  const float* cols[SXMC_MAX_NFIELDS];   // per FIELD column (tiled layout), nwtiles*1024 floats
  unsigned long long nwtiles;            // warp-tiles (samples padded with +inf observables)
  unsigned* bins;                        // [total_nbins] global histogram
  unsigned* norm;                        // in-domain count slot
  const double* params;                  // parameter buffer + offset
  long long param_stride;
  int total_nbins;
  int nobs;
  int nfields;
  int nsyst;
  int nslots;
  int slot_field[MAX_SLOTS];             // register slot -> field index (observables first)
  double lower[SXMC_MAX_NFIELDS];
  double upper[SXMC_MAX_NFIELDS];
  double scale[SXMC_MAX_NFIELDS];        // nbins / (upper - lower), src/pdfz.cpp:365-368
  int stride[SXMC_MAX_NFIELDS];          // row-major strides, src/pdfz.cpp:206-213
  SystDesc syst[SXMC_MAX_SYST];
};

// Per-launch description of the fast binning kernel, passed BY VALUE (kernel parameters live in
// the constant bank: a CTA can issue its first bulk copy without a single global load).
constexpr int FAST_MAX_JOBS = 32;
constexpr int FAST_MAX_OPS = 4;
struct FastGeo {
  double lo[3], hi[3], sc[3];
  int st[3];
  int total_nbins;
};
struct FastJob {
  const float* col[MAX_SLOTS];   // register slot -> column (tiled layout)
  unsigned* bins;
  unsigned* norm;
  const double* params;
  int param_stride;
  int nsyst;
  unsigned short code[FAST_MAX_OPS];   // type | obs_slot << 4 | extra_slot << 8
  short pidx[FAST_MAX_OPS];            // coefficient index of each systematic
  unsigned char npars[FAST_MAX_OPS];   // 0: the systematic has no coefficient (p = 0)
};
struct FastParams {
  unsigned long long prefix[FAST_MAX_JOBS + 1];   // chunk prefix sums
  FastJob job[FAST_MAX_JOBS];
  FastGeo geo;
  int njobs, R, S;
  unsigned stage_off;
  // job-aligned partition (small fits): CTAs [cta_first[j], cta_first[j+1]) work on job j only,
  // so no CTA pays the per-segment overhead twice; aligned == 0: even split of all chunks
  unsigned short cta_first[FAST_MAX_JOBS + 1];
  int aligned;
  int cta_split;                // partition form, 0 in production (Tuning::cta_split, capi_internal.h)
  int nwarps;                   // warps of the CTA the launch runs with (<= FAST_WARPS): tiles are dealt over these
  // profiling hook (sxmc_bin_debug_stamps): [grid][32] %globaltimer stamps, NULL in production.
  // Per CTA: 0 entry, 1 pipeline initialised, 2 first data arrived (warp 0), 3 segments done,
  // 4+3s+{0,1,2} of segment s < 4: parameters resolved / warp 0 out of its chunk loop / merged,
  // 11: %smid + 1, 16+w: warp w out of the chunk loop of the CTA's LAST segment, 31 exit.
  unsigned long long* xdbg;
};

// Arguments of the persistent chain kernel (chain_kernels.cu), by value.
struct ChainState;
struct NllSignal;
struct PersistParams {
  ChainState* st;
  const NllSignal* sigs;        // static per-signal fields (read_bins, nexpected, bin_volume, n_mc, source_id)
  unsigned* hist_buf[2];        // [nsignals][total_nbins] + norms[nsignals], double-buffered
  unsigned long long hist_words;
  double* partials;             // [grid] per-CTA event sums
  double* bcast;                // [2][SXMC_MAX_PARAMETERS] proposed vector by step parity
  unsigned* sync;               // [0] bin barrier, [1] lookup arrivals, [2] go, [3] abort
  long long nevents;
  int ns, np, nsources;
  unsigned nsteps;              // steps in this launch
  unsigned step0;               // chain index of the first one (histogram buffer parity)
  int rebin;
  unsigned long long* stamps;   // optional [64][8] phase timestamps of CTA 0 (ns), profiling only
};

// Arguments of the chain batch kernel (batch_kernels.cu), by value.
constexpr int BATCH_MAX_CHAINS = 148;   // one chain per SM in the lookup/tail phase
constexpr int BATCH_MAX_GROUP = 8;      // chains binned per shared-memory histogram pass
constexpr int BATCH_PDF_TABLE = 4096;   // (signal, bin) PDF values kept in shared memory by the grouped lookup
struct BatchParams {
  ChainState* st;               // [nchains]
  const NllSignal* sigs;        // [nchains][ns]: per-chain read_bins; bins/norm unused here
  unsigned* hist;               // [nchains][hist_words]: [ns][total_nbins] bins + norms[ns]
  unsigned long long hist_words;
  const long long* nevents;     // [nchains]
  unsigned* sync;               // [0] barrier counter, [3] abort
  int nchains, ns, np, nsources;
  unsigned nsteps;              // MCMC steps of every chain in this launch
  int do_init;                  // first super-step = initial evaluation + first proposal
  int rebin;
  int group;
  const unsigned* mult;         // optional [nchains][total_nbins] events per bin (grouped lookup), else NULL
  int mono;                     // 1-D monotone program on sorted samples: endpoint evaluation (batch_kernels.cu)
  unsigned long long* stamps;   // optional [grid][8] %globaltimer stamps of super-step 20 (profiling)
};

// Arguments of the monotone chain batch kernel (batch_mono_kernels.cu), by value: every CTA
// holds the sorted samples [s0, s0 + n) of one job as a plain float array in shared memory.
constexpr int MONO_MAX_CTAS = 160;
struct MonoJob {
  const float* col;                    // the observable's column (tiled layout)
  int nsyst;
  unsigned short code[FAST_MAX_OPS];   // type | obs_slot << 4 | extra_slot << 8 (as FastJob)
  short pidx[FAST_MAX_OPS];
  unsigned char npars[FAST_MAX_OPS];
};
struct MonoParams {
  FastGeo geo;
  MonoJob job[FAST_MAX_JOBS];
  unsigned s0[MONO_MAX_CTAS];
  unsigned n[MONO_MAX_CTAS];
  unsigned char jobof[MONO_MAX_CTAS];
  int njobs;
};

// Arguments of the peer-memory histogram all-reduce (p2p_kernels.cu), by value.
constexpr int P2P_MAX_RANKS = 8;
struct P2PParams {
  const unsigned* part[P2P_MAX_RANKS];   // every rank's partial-histogram window (own one included)
  unsigned* flags[P2P_MAX_RANKS];        // every rank's flag array [2][P2P_MAX_RANKS]: flags[p][r] = last step rank r
                                         // published to p ("partial complete"); [p][8 + r] = "my slice of the sum is written"
  unsigned* sum;                         // this rank's complete histogram (one-shot form)
  unsigned long long sum_offset;         // two-shot form: words from a window's start to the sum of this parity
  unsigned* seq;                         // this rank's step counter (private, advanced by the kernel)
  unsigned* ticket;
  unsigned* abort_flag;
  unsigned long long words;              // histogram words
  unsigned long long pitch;              // words between the two parities inside a window
  int parity;
  int nranks, rank;
  int twoshot;                           // reduce-scatter + all-gather (nranks > 2)
  unsigned long long peer_timeout_ns;    // wall-clock bound of a wait for the other ranks (they are other processes)
};

// One signal as seen by the lookup + event-likelihood kernel.
struct NllSignal;
struct ChainState;
// Arguments of the fused exchange + rank-sharded lookup (nll_sharded_kernels.cu), by value.
constexpr int P2P_PARTIALS_MAX = 148 * 8;  // doubles per window: one per logical block of the unsharded lookup grid
struct ShardNllParams {
  P2PParams W;                  // windows, flags, parity of this step's partial histograms
  const NllSignal* sigs;        // static per-signal fields (nexpected, bin_volume, n_mc, source_id)
  ChainState* st;
  const unsigned* mult;         // events per bin [nbins]
  unsigned* zero_ptr;           // partial buffer of the next step, cleared here
  unsigned long long zero_words;
  unsigned long long partials_offset;   // words from a window's start to its [P2P_PARTIALS_MAX] x {lo, tag, hi, tag}
  int ns, np, nsources, nbins, mode;
  int logical_grid, lb0;        // filled by launch_nll_sharded
};

struct NllSignal {
  const unsigned* bins;       // histogram of this step
  const unsigned* norm;       // its in-domain count
  const int* read_bins;       // per event flat bin, -1 outside, -2 other dataset
  double nexpected;
  double bin_volume;
  unsigned n_mc;
  int total_nbins;
  short source_id;
};

// Device-resident chain state (the hemi::Arrays of src/mcmc.cpp:159-232 in one struct).
struct ChainState {
  double current[SXMC_MAX_PARAMETERS];
  double proposed[SXMC_MAX_PARAMETERS];
  double means[SXMC_MAX_PARAMETERS];
  double sigmas[SXMC_MAX_PARAMETERS];
  float jump_width[SXMC_MAX_PARAMETERS];
  int fixed[SXMC_MAX_PARAMETERS];
  double stat_sx[SXMC_MAX_PARAMETERS];    // running sums over the rows since the last reset
  double stat_sxx[SXMC_MAX_PARAMETERS];
  double stat_n;
  double current_nll;
  double proposed_nll;
  double event_total;                     // persistent event_total_sum (NaN keeps the old one)
  float scale_factor;
  int nparameters, nsources, nsignals;
  int debug_mode;
  unsigned step;                          // next step index i of the loop
  unsigned burnin_steps;
  long long nrows;                        // rows in the ntuple since the last Reset()
  long long accepted;
  unsigned long long seed, stream_id;
  // tapes (test hook)
  const double* tape_normals; unsigned long long n_normals, i_normal;
  const double* tape_uniforms; unsigned long long n_uniforms, i_uniform;
  // outputs
  float* rows;                            // [row_cap][nparameters+1]
  long long row_cap;
  double* nll_trace;                      // optional [nsteps]
  double* prop_trace;                     // optional [nsteps][nparameters]
  double last_nll, last_event_sum;        // results of an evaluate-only pass
  // lookup_nll_kernel: the step's random numbers, drawn by block 0 while the other blocks look
  // up events, read by whichever block finishes last (Philox is stateless: counter = step)
  double pre_z[SXMC_MAX_PARAMETERS];
  double pre_u;
};

// ---- fake-data generator (ensemble_kernels.cu) ---------------------------------------------
struct GenSignal {
  const unsigned* cum;     // inclusive prefix sums of the histogram at the systematics' means
  const unsigned* norm;    // its in-domain count (get_efficiency)
  double nexpected;
  unsigned n_mc;
  unsigned dataset;
};
struct GenGeo {
  double lower[3], width[3];   // bin low edge = lower + k * width
  int stride[3];
  float lim_lo[3], lim_hi[3];  // observable limits of the rejection test (float vectors in the reference)
  int nobs, nlimits, total_nbins;
};
constexpr int CONTOUR_MAX_LEVELS = 12;
struct ContourLevels {
  double thr[CONTOUR_MAX_LEVELS];
  int n;
};

enum NllMode { NLL_MODE_STEP = 0, NLL_MODE_INIT = 1, NLL_MODE_EVAL_ONLY = 2 };

// eval_pdf's value, (float)(count / (norm * bin_volume)) (src/pdfz.cpp:429-432).  A zero count is the
// common case (a signal occupies a fraction of the bins), and a zero numerator sends CUDA's double
// division down its out-of-line slow path (ncu: 7 calls per warp in workload D's lookup): answer it
// directly -- 0 / bn is 0 for every bn except 0 and NaN, where it is NaN, as the division gives.
__device__ __forceinline__ float pdf_value(unsigned count, double bn) {
  if (count == 0u) return (bn == 0.0 || bn != bn) ? __int_as_float(0x7fc00000) : 0.0f;
  return (float)((double)count / bn);
}

// ---- programmatic dependent launch ---------------------------------------------------------
// A kernel launched with `pdl` may start while its predecessor in the stream is still running:
// everything it does before pdl_wait() must not depend on (or disturb) the predecessor; pdl_wait()
// returns once the predecessor has completed and its writes are visible.  pdl_trigger() lets the
// successor's blocks be scheduled as soon as this grid's blocks have all called it (or exited).
// Both are no-ops in a kernel launched without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

#ifndef __CUDACC_RTC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                     cudaStream_t stream, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#endif

#ifndef __CUDACC_RTC__
// error plumbing --------------------------------------------------------------------------
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define SXMC_CUDA(expr)                                                       \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return ::sxmc::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#endif

}  // namespace sxmc
