// kernels.h -- host-callable launchers of the sm_100a kernels (internal to libsxmc_b200).
#pragma once
#include "common.cuh"

namespace sxmc {

struct PrepGeo {     // histogram geometry passed by value to the small preparation kernels
  double lower[SXMC_MAX_NFIELDS], upper[SXMC_MAX_NFIELDS], scale[SXMC_MAX_NFIELDS];
  int nbins[SXMC_MAX_NFIELDS], stride[SXMC_MAX_NFIELDS];
  int total_nbins;
};

struct BinLaunch {
  const BinJob* jobs_host;   // the same table on the host (the fast kernel takes it by value)
  const BinJob* jobs_dev;
  const unsigned long long* prefix_dev;   // [njobs+1] prefix sums of work units: chunks of
                                          // rows_per_stage rows (fast) / 8-warp-tile block-tiles (generic)
  int njobs;
  unsigned long long total_units;
  bool fast;                 // register-resident kernel (uniform geometry, nobs<=3)
  bool smem_hist;            // shared-memory privatised histograms
  int nobs, nslots;
  int max_total_nbins;
  int blocks_per_sm;         // <=0: occupancy maximum (generic kernel)
  int cta_split;             // fast kernel partition, see Tuning::cta_split (capi_internal.h)
  bool pdl;                  // fast kernel: programmatic dependent launch on the previous kernel of the stream
  bool jit;                  // the program has no instantiation in the build: NVRTC-compiled kernel (jit.cu)
  int prog_nops;                 // >= 0: compile-time systematic program, < 0: run-time dispatch
  unsigned prog_codes;
  int rows_per_stage, nstages;   // fast kernel pipeline shape
  int nwarps;                    // fast kernel: warps per CTA (0 = FAST_WARPS)
  unsigned stage_off;            // byte offset of the stages in dynamic shared memory
  size_t smem_bytes;
  double lower[3], upper[3], scale[3];
  int stride[3];
};

struct EvalPdfJob {
  const int* read_bins;
  const unsigned* bins;
  const unsigned* norm;
  float* out;                // pdf buffer + offset
  long long stride;
  long long npoints;
  double bin_volume;
  int total_nbins;
};

bool bin_fast_supported(int nobs, int nslots);
bool bin_fast_has_program(int nobs, int nslots, int nops, unsigned codes);
bool bin_fast_plan_big(int nslots, int hist_bins, size_t static_reserve, int* nwarps_out, int* R_out, int* S_out,
                       unsigned* stage_off_out, size_t* smem_out);
bool bin_fast_plan(int nslots, int hist_bins, size_t static_reserve, int* R_out, int* S_out,
                   unsigned* stage_off_out, size_t* smem_out);
constexpr size_t BIN_STATIC_RESERVE = 4096;   // static shared memory of bin_fast_kernel
// FastParams of jobs [g0, g0+ng) of a planned launch; *total_out = their chunk count
// `grid` = the CTA count the kernel will be launched with (decides the job-aligned partition)
void build_fast_params(const BinLaunch& L, int g0, int ng, int grid, FastParams& P,
                       unsigned long long* total_out);
// chain_kernels.cu
int launch_chain_persistent(const BinLaunch& L, const PersistParams& Q, cudaStream_t stream);
int launch_bin(const BinLaunch& L, cudaStream_t stream);
// ---- jit.cu: bin_fast_kernel NVRTC-compiled for a program outside the build's table
void* jit_bin_fast_kernel(int nobs, int ns, bool smem, int nops, unsigned codes);   // nullptr: not available
int jit_launch_bin(void* fn, unsigned grid, size_t smem_bytes, cudaStream_t stream, bool pdl, const FastParams& P);
int jit_compile_check(int nobs, int ns, int smem, int nops, unsigned codes, size_t* cubin_bytes_out);
void set_bin_xdbg(unsigned long long* dev);   // [256][32] stamps of the next fast launches, or NULL
int launch_zero_jobs(const BinJob* jobs_dev, int njobs, cudaStream_t stream);
int prepare_layout(const float* aos_dev, unsigned long long n, int nfields, int nobs,
                   const PrepGeo& g, int primary, bool sort, float* cols_dev,
                   unsigned long long col_pitch, unsigned* perm_dev, cudaStream_t stream);
int unprepare_to_aos(const float* cols_dev, unsigned long long col_pitch, const unsigned* perm_dev,
                     unsigned long long n, int nfields, float* aos_dev, cudaStream_t stream);
int launch_read_bins(const float* points_dev, long long npoints, int nobs, const PrepGeo& g,
                     unsigned dataset, int* read_bins_dev, cudaStream_t stream);
int launch_eval_pdf(const EvalPdfJob* jobs_dev, int njobs, long long max_points, cudaStream_t stream);

// ---- nll_kernels.cu ----
struct NllLaunch {
  const NllSignal* sigs_dev;
  const unsigned* norms_dev;    // [nsignals] in-domain counts of this step (== sigs[j].norm)
  int nsignals, nparameters, nsources;
  long long nevents;
  ChainState* state;            // device
  double* partials;             // [grid] device
  unsigned* ticket;             // device, zero before first use (self-resetting)
  int mode;                     // NllMode
  float* lut;                   // optional materialised lut [nsignals][nevents] (signal-major)
  unsigned* zero_ptr;           // optional buffer to clear for the next step
  unsigned long long zero_words;
  int grid;                     // fixed for the life of a chain (deterministic summation order)
  const unsigned* mult;         // optional events per bin [mult_nbins]: grouped lookup (all signals one dataset)
  int mult_nbins;               // bins per PDF (also the row length of ptable)
  const double* ptable;         // optional [nsignals][mult_nbins] PDF values of this step (launch_pdf_table)
  bool pdl;                     // programmatic dependent launch on the previous kernel of the stream (the binning)
  bool zero_early;              // clear zero_ptr before waiting for that kernel (nobody else reads the buffer)
};
int nll_grid_for(long long nevents);
void set_nll_xdbg(unsigned long long* dev);   // [1202][8] stamps of the next lookup launches, or NULL
int launch_pdf_table(const NllSignal* sigs_dev, const unsigned* norms_dev, int nsignals, int nbins, double* table_dev,
                     cudaStream_t stream);
int launch_nll(const NllLaunch& L, cudaStream_t stream);
// the sample-sharded step's exchange fused into a rank-sharded grouped lookup (nll_sharded_kernels.cu)
int nll_sharded_logical_grid(int nbins);
int launch_nll_sharded(ShardNllParams Q, bool pdl, cudaStream_t stream);

// standalone MCMC::nll over a materialised lut (src/mcmc.cpp:390-415)
int launch_nll_lut(const float* lut, size_t ne, size_t ns, const double* pars, size_t npars,
                   size_t nsources, const double* means, const double* sigmas,
                   const double* nexpected, const unsigned* n_mc, const short* source_id,
                   const unsigned* norms, double* out2_dev, cudaStream_t stream);

// ---- batch_kernels.cu ----
struct BatchPlan {
  BinLaunch launch;   // shared-memory plan of the resident samples beside `group` histograms
  int group;
  int grid;
};
int plan_chain_batch(const BinLaunch& L, int nchains, BatchPlan* plan);
int launch_chain_batch(const BatchPlan& plan, const BatchParams& Q, cudaStream_t stream);
// ---- batch_mono_kernels.cu ----
struct MonoPlan {
  MonoParams params;
  int grid;
  size_t smem_bytes;
  int max_edges;          // largest number of nominal bin edges inside one CTA's range
  int prog_nops;
  unsigned prog_codes;
};
int plan_chain_batch_mono(const BinLaunch& L, const unsigned long long* nsamples, int nchains, MonoPlan* plan);
int launch_chain_batch_mono(const MonoPlan& plan, const BatchParams& Q, cudaStream_t stream);
int column_has_nan_or_neginf(const float* col_dev, unsigned long long n, bool* out, cudaStream_t stream);

// ---- p2p_kernels.cu ----
int launch_hist_allreduce_p2p(const P2PParams& W, cudaStream_t stream);

// ---- ensemble_kernels.cu ----
int launch_event_mult(const int* read_bins_dev, long long nevents, int tnb, unsigned* mult_dev, cudaStream_t stream);
int launch_cum(const unsigned* bins_dev, int nb, unsigned* cum_dev, cudaStream_t stream);
int launch_generate(const GenSignal* sigs_dev, int n, const GenGeo& geo, int poisson,
                    unsigned long long seed, unsigned long long stream_id, unsigned long long* counts_dev,
                    unsigned long long* offsets_dev, float* events_dev, unsigned long long cap_events,
                    cudaStream_t stream);
int launch_rows_min_nll(const float* rows_dev, long long nrows, int np, float* out_dev, cudaStream_t stream);
int launch_rows_extents(const float* rows_dev, long long nrows, int np, double off, const ContourLevels& lv,
                        unsigned long long* counts_dev, float* mins_dev, float* maxs_dev, cudaStream_t stream);

}  // namespace sxmc
