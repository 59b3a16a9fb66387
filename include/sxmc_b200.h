/* sxmc_b200.h -- C ABI of libsxmc_b200.so: the B200 (sm_100a) device layer that replaces
 * hemi underneath sxmc's per-step likelihood hot path.
 *
 * Every entry point cites the reference interface it replaces (paths relative to the
 * mastbaum/sxmc tree).  Plain pointers and sizes only; no C++ or torch types.  All
 * functions return 0 on success and a non-zero SXMC_E* code on failure;
 * sxmc_last_error() returns the message (argument-validation messages are the
 * reference's own pdfz::Error strings, so the C++ shim can rethrow them verbatim).
 *
 * There is no CPU fallback: every compute entry point fails with SXMC_ECUDA when no
 * CUDA device is usable.
 */
#ifndef SXMC_B200_H
#define SXMC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SXMC_OK 0
#define SXMC_EINVAL 1  /* argument validation failed (a pdfz::Error in the reference) */
#define SXMC_ECUDA 2   /* CUDA runtime / driver failure, or no device */
#define SXMC_ESTATE 3  /* call sequence error (e.g. eval before buffers are bound) */
#define SXMC_ENCCL 4   /* NCCL failure */
#define SXMC_ENOTSMALL 5 /* sxmc_batch_*: the fit's samples do not fit on chip (or its shape has no batch
                          * kernel): run its chains one at a time (sxmc_chain_run).  The ONLY code a
                          * caller may answer with that fallback; timeouts and aborts are SXMC_ESTATE */

/* numeric values of pdfz::Systematic::Type (src/pdfz.h:111-116) */
#define SXMC_SYST_SHIFT 0
#define SXMC_SYST_SCALE 1
#define SXMC_SYST_RESOLUTION_SCALE 2
#define SXMC_SYST_CTSCALE 3

#define SXMC_MAX_NFIELDS 10   /* pdfz::MAX_NFIELDS, src/pdfz.cpp:17 */
#define SXMC_MAX_SYST 8       /* systematics per PDF */
#define SXMC_MAX_NPARS 4      /* coefficients per systematic */
#define SXMC_MAX_PARAMETERS 256

typedef struct sxmc_hist sxmc_hist;     /* one pdfz::EvalHist (src/pdfz.h:402-574) */
typedef struct sxmc_chain sxmc_chain;   /* one MCMC object + its chain state (src/mcmc.h:36-138) */
typedef struct sxmc_comm sxmc_comm;     /* NCCL communicator for sample-sharded fits */

/* ---- library ---------------------------------------------------------------------- */
const char* sxmc_last_error(void);
const char* sxmc_version(void);
/* cudaSetDevice + capability check (compute capability 10.x required). */
int sxmc_init(int device);
int sxmc_device_synchronize(void);

/* ---- device buffers: the device half of hemi::Array<T>
 *      (contrib/hemi/hemi/array.h:21-267; allocate :211-226, copies :137-149, :228-260) --- */
int sxmc_buf_alloc(size_t bytes, void** dev_out);
int sxmc_buf_free(void* dev);
int sxmc_buf_upload(void* dev, const void* host, size_t bytes);     /* copyHostToDevice */
int sxmc_buf_download(void* host, const void* dev, size_t bytes);   /* copyDeviceToHost */
int sxmc_buf_zero(void* dev, size_t bytes);
/* pinned host staging (hemi::Array(n, usePinned=true), array.h:25-37) */
int sxmc_host_alloc(size_t bytes, void** host_out);
int sxmc_host_free(void* host);

/* ---- pdfz::EvalHist ---------------------------------------------------------------- */
/* pdfz::Eval::Eval + EvalHist::EvalHist (src/pdfz.cpp:57-97, :179-236).  `samples` is the
 * reference's host AoS array (nfloats = nsamples*nfields floats, row-major); it is copied
 * to the device and owned by the handle.  The n* arguments carry the vector sizes the
 * reference validates.  Errors (SXMC_EINVAL) carry the reference's messages. */
int sxmc_hist_create(const float* samples, size_t nfloats, int nfields, int nobservables,
                     const double* lower, int nlower, const double* upper, int nupper,
                     const int* nbins, int nnbins, unsigned dataset, sxmc_hist** out);
/* EvalHist::~EvalHist (src/pdfz.cpp:239-242) */
int sxmc_hist_destroy(sxmc_hist* h);
/* Eval::AddSystematic (src/pdfz.cpp:126-174).  `pars` are GLOBAL coefficient indices into
 * the bound parameter buffer (src/config.cpp:118-120), copied at call time. */
int sxmc_hist_add_systematic(sxmc_hist* h, int type, int obs, int extra_field, int npars,
                             const short* pars);
/* EvalHist::SetEvalPoints (src/pdfz.cpp:245-302): host rows of nobservables+1 floats. */
int sxmc_hist_set_eval_points(sxmc_hist* h, const float* points, size_t nfloats);
/* Eval::SetPDFValueBuffer / SetNormalizationBuffer / SetParameterBuffer
 * (src/pdfz.cpp:105-124): borrowed DEVICE pointers, re-pointable at any time. */
int sxmc_hist_set_pdf_buffer(sxmc_hist* h, float* dev, long offset, long stride);
int sxmc_hist_set_norm_buffer(sxmc_hist* h, unsigned* dev, long offset);
int sxmc_hist_set_param_buffer(sxmc_hist* h, const double* dev, long offset, long stride);
/* EvalHist::EvalAsync (src/pdfz.cpp:441-488): zero + transform + bin (+ eval_pdf) enqueued
 * on the handle's stream;  EvalHist::EvalFinished (:491-495): stream synchronise. */
int sxmc_hist_eval_async(sxmc_hist* h, int do_eval_pdf);
int sxmc_hist_eval_finished(sxmc_hist* h);
/* The `for all pdfs: EvalAsync(); for all pdfs: EvalFinished()` idiom of
 * src/mcmc.cpp:264-271 as ONE batched launch over all handles (first handle's stream). */
int sxmc_hists_eval_async(sxmc_hist* const* hs, int n, int do_eval_pdf);
/* Host-buffer convenience used for end-to-end timing: uploads `params` (nparams doubles),
 * evaluates all handles in one batched launch with their norm slots in one device array and
 * their pdf values signal-major (src/mcmc.cpp:236-237 layout), copies both back.
 * pdf_out may be NULL (=> EvalAsync(false)). */
int sxmc_hists_eval_host(sxmc_hist* const* hs, int n, const double* params, int nparams,
                         long param_offset, unsigned* norms_out, float* pdf_out);
/* introspection (the reference's protected members, src/pdfz.h:558-573) */
int sxmc_hist_total_nbins(const sxmc_hist* h);
double sxmc_hist_bin_volume(const sxmc_hist* h);
size_t sxmc_hist_nsamples(const sxmc_hist* h);
size_t sxmc_hist_neval_points(const sxmc_hist* h);
int sxmc_hist_get_bins(sxmc_hist* h, unsigned* host_out);          /* bins->hostPtr() */
int sxmc_hist_get_read_bins(sxmc_hist* h, int* host_out);          /* read_bins->hostPtr() */
/* EvalHist::GetSamples (src/pdfz.h:542-556): original order, original AoS layout. */
int sxmc_hist_get_samples(sxmc_hist* h, float* host_out);
/* layout control: samples are stored as bin-sorted, lane-interleaved SoA columns, built
 * lazily on the first evaluation (the reference's first-call Optimize(), src/pdfz.cpp:442-444).
 * sort=0 keeps the caller's sample order (used to measure what sorting buys). */
int sxmc_hist_set_layout(sxmc_hist* h, int sort);
int sxmc_hist_prepare(sxmc_hist* h);
/* tuning knobs of the binning launch (EvalHist::OptimizeBin analogue, src/pdfz.cpp:630-727).  The call
 * sets the DEFAULTS: every handle (hist, chain, batch) copies them when it is created and keeps them,
 * so two fits in one process can be tuned independently (set, create, set differently, create).
 * smem_hist_max_bins: histograms up to this many bins are privatised in shared memory;
 * blocks_per_sm <= 0 keeps the default;  kernel_select: 0 = automatic, bit 0 = force the
 * generic (interpreting) kernel, bit 1 = force run-time systematic dispatch in the fast
 * kernel (no compile-time program), bit 2 = the chain batch never takes its monotone endpoint
 * path, bits 3-4 = form of the peer-memory all-reduce (0 automatic: two-shot above two ranks,
 * 1 one-shot, 2 two-shot), bit 5 = chains set up from now on look events up one by one even when
 * every signal reads the same dataset (the multi-dataset form), bit 6 = chain steps are launched
 * without programmatic dependent launch, bit 7 = no run-time compilation of systematic programs (below),
 * bits 8-9 = how the fast binning kernel deals its chunks (0 production: a CTA's range dealt to its
 * warps round robin, global-memory histograms with warps independent of each other; 1 one contiguous
 * range per warp inside per-CTA segments; 2 global histograms, one contiguous range per warp of the
 * grid; 3 global histograms by per-CTA segments), bits 10-11 = pipeline chunks of the fast binning
 * kernel (0 automatic: four rows per chunk and 12 warps for launches that stream >= 1 GiB into
 * global-memory histograms of <= 65536 bins, two rows and 16 warps otherwise; 1 always two rows;
 * 2 four rows wherever two such stages fit) -- used by the parity tests and measurements to cover
 * every path. */
int sxmc_set_bin_tuning(int smem_hist_max_bins, int blocks_per_sm, int kernel_select);
/* device time (ms) of the last batched binning launch / last nll launch, from CUDA events
 * recorded on the launching stream when timing is enabled. */
int sxmc_enable_kernel_timing(int on);
int sxmc_last_kernel_ms(float* bin_ms, float* nll_ms);            /* of whichever chain was timed last */
/* Profiling hook of the fast binning kernel (no reference counterpart): while enabled every
 * launch writes [grid <= 256][32] %globaltimer stamps (ns) per CTA -- 0 entry, 1 pipeline
 * initialised, 2 first data arrived, 3 = number of segments, 4+3s+{0,1,2} parameters resolved /
 * warp 0 out of the chunk loop / histogram merged for segment s < 4, 11 = %smid + 1, 16+w warp w out of the
 * chunk loop of the last segment, 31 exit.  out256x32 (host, may be NULL) receives the stamps
 * of the last launch.  tools/prof_bin.py prints the prologue / steady state / epilogue split. */
int sxmc_bin_debug_stamps(int enable, unsigned long long* out256x32);
/* The fast binning kernel takes a fit's systematic program (which type acts on which column, in which
 * order; apply_systematic, src/pdfz.cpp:306-331) as a template parameter.  The build instantiates the
 * BASELINE workloads' programs; any other program of up to four single-coefficient systematics is
 * compiled for the fit with NVRTC when it is first planned (~0.4 s, cached per process; SXMC_NO_JIT=1
 * or kernel_select bit 7 keeps the run-time dispatching kernel instead).  This entry point only compiles (no GPU needed) and reports the cubin
 * size: nslots = columns the kernel loads, codes = 8 bits per systematic, type | slot_obs << 2 |
 * slot_extra << 5. */
int sxmc_jit_compile_check(int nobs, int nslots, int smem_hist, int nops, unsigned codes, size_t* cubin_bytes_out);
/* The same for the lookup + NLL + Metropolis kernel: [block < 1200][8] stamps -- 0 entry, 1 state
 * prefetched and next histogram buffer cleared, 2 per-signal constants ready, 3 lookup loop done,
 * 4 block sum done, 5 ticket taken -- and in row 1200 those of the block that finished the step plus
 * 6 event total known, 7 step finished (row 1201, word 0: its block index).
 * tools/prof_nll.py prints the phase split. */
int sxmc_nll_debug_stamps(int enable, unsigned long long* out1202x8);

/* ---- nll_kernels: standalone NLL of a materialised lut (MCMC::nll, src/mcmc.cpp:390-415:
 *      nll_event_chunks -> nll_event_reduce -> nll_total).  All pointers are DEVICE
 *      pointers except the two outputs. ------------------------------------------------ */
int sxmc_nll_eval(const float* lut_dev, size_t nevents, size_t nsignals,
                  const double* pars_dev, size_t nparameters, size_t nsources,
                  const double* means_dev, const double* sigmas_dev,
                  const double* nexpected_dev, const unsigned* n_mc_dev,
                  const short* source_id_dev, const unsigned* norms_dev,
                  double* nll_out_host, double* event_sum_out_host);
/* The same operation for callers that must not allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * A multi-block variant of the kernel behind sxmc_nll_eval whose scratch (per-block
 * partial sums + a ticket, 9.5 KB) is a file-scope device array every GPU that loads the library holds
 * until exit; the call issues one 256-byte cudaMemsetAsync and one kernel on `cuda_stream` (a
 * cudaStream_t; NULL = the default stream), calls no cudaMalloc / cudaFree / allocating routine and does
 * not synchronise.  All pointers are DEVICE pointers, the output too: out2_dev[0] = nll, out2_dev[1] =
 * the event sum, valid once the stream reaches that point.  Limits (beyond them: SXMC_EINVAL with a
 * message, nothing launched): nsignals <= 128, nparameters <= SXMC_MAX_PARAMETERS, nevents <= 2^31.
 * The scratch is shared by every call on a device and keeps the last call's contents (the ticket is
 * cleared at the start of each call): calls that OVERLAP on the GPU race -- serialise them on one
 * stream or with events.  Same arithmetic as sxmc_nll_eval per event; the event sum is reduced over more
 * blocks, so the two agree to rounding (<= 1e-12 relative), not bit for bit. */
int sxmc_nll_eval_noalloc(const float* lut_dev, size_t nevents, size_t nsignals,
                          const double* pars_dev, size_t nparameters, size_t nsources,
                          const double* means_dev, const double* sigmas_dev,
                          const double* nexpected_dev, const unsigned* n_mc_dev,
                          const short* source_id_dev, const unsigned* norms_dev,
                          double* out2_dev, void* cuda_stream);

/* ---- MCMC (src/mcmc.h:36-138) ------------------------------------------------------- */
typedef struct {
  int nsources;                 /* MCMC::nsources */
  int nsignals;                 /* MCMC::nsignals */
  int nparameters;              /* nsources + sum(systematics[i].npars), src/mcmc.cpp:47-54 */
  int nobservables;
  const double* parameter_means;   /* [nparameters] src/mcmc.cpp:60-80 */
  const double* parameter_sigma;   /* [nparameters] */
  const int* parameter_fixed;      /* [nparameters] */
  int systematics_fixed;           /* src/mcmc.cpp:67-72 */
  int nsystematics;                /* systematics.size() (:264 guard) */
  sxmc_hist* const* pdfs;          /* [nsignals] signals[i].histogram */
  const double* nexpected;         /* [nsignals] */
  const unsigned* n_mc;            /* [nsignals] */
  const short* source_id;          /* [nsignals] */
} sxmc_chain_config;

/* MCMC::MCMC (src/mcmc.cpp:28-130).  The handles are borrowed. */
int sxmc_chain_create(const sxmc_chain_config* cfg, sxmc_chain** out);
int sxmc_chain_destroy(sxmc_chain* c);
/* Philox4x32-10 seed (replaces init_device_rngs / gRandom->SetSeed, src/mcmc.cpp:118-128);
 * `stream_id` separates independent chains of an ensemble (experiment index). */
int sxmc_chain_set_seed(sxmc_chain* c, uint64_t seed, uint64_t stream_id);
/* Test hook: draw the chain's random numbers from host tapes (standard normals, U(0,1))
 * instead of Philox, in the reference's CPU call order (src/nll_kernels.cpp:45,65).
 * NULL/0 returns to Philox. */
int sxmc_chain_set_tapes(sxmc_chain* c, const double* normals, size_t n_normals,
                         const double* uniforms, size_t n_uniforms);
/* Optional sample-sharded mode: histograms are summed over the communicator after binning. */
int sxmc_chain_set_comm(sxmc_chain* c, sxmc_comm* comm);
/* MCMC::operator() (src/mcmc.cpp:143-387).  `data` = host rows of nobservables+1 floats.
 * The whole loop runs device-resident; sync_interval is accepted for signature parity
 * (the device chain buffer holds every step, so there is no periodic drain).
 * On return *nrows_out = rows left in the likelihood-space ntuple
 * (steps [2*burnin, nsteps) unless debug_mode). */
int sxmc_chain_run(sxmc_chain* c, const float* data, size_t data_nfloats, unsigned nsteps,
                   float burnin_fraction, int debug_mode, unsigned sync_interval,
                   long* nrows_out);
/* The same call in three parts, so that a caller (bench.py, an ensemble driver) can keep a
 * chain resident and advance it in segments:  begin = src/mcmc.cpp:143-258 (buffers, initial
 * widths, SetEvalPoints, initial evaluation and NLL, first proposal; `nsteps` is the total the
 * segments will add up to -- it fixes the burn-in points);  advance = `nsteps` iterations of
 * the loop at :261-378;  end = the final drain (:351-377) that makes rows/accepted readable. */
int sxmc_chain_begin(sxmc_chain* c, const float* data, size_t data_nfloats, unsigned nsteps,
                     float burnin_fraction, int debug_mode);
int sxmc_chain_advance(sxmc_chain* c, unsigned nsteps);
int sxmc_chain_end(sxmc_chain* c, long* nrows_out);
/* Measurement variant of advance: plain stream launches with a CUDA event around every kernel
 * and, when flush_bytes > 0, an L2 flush (memset of flush_bytes) before every step, outside
 * the timed intervals.  Outputs are sums over the nsteps steps, in ms. */
int sxmc_chain_advance_timed(sxmc_chain* c, unsigned nsteps, size_t flush_bytes,
                             float* step_ms_sum, float* bin_ms_sum, float* nll_ms_sum);
/* the part of bin_ms_sum of the last sxmc_chain_advance_timed spent between the end of the
 * binning kernel and the end of the histogram all-reduce (sample-sharded chains; else ~0) */
int sxmc_chain_last_comm_ms(sxmc_chain* c, float* ms_out);
/* which binning kernel the chain's steps launch (decided when the data are set): 0 per-sample interpreter
 * (multi-coefficient systematics, > 3 observables), 1 fast kernel with run-time dispatch of the systematics,
 * 2 fast kernel with a compile-time program of the build's table, 3 the same compiled for this fit by NVRTC */
int sxmc_chain_bin_kernel_kind(sxmc_chain* c, int* kind_out);
/* The pipeline shape the chain's binning launch was planned with (no reference counterpart; reported by
 * bench.py and checked by the parity tests): warps per CTA, 512-byte rows per pipeline chunk, stages per
 * warp, and whether the histograms are privatised in shared memory.  Zeros for the interpreter. */
int sxmc_chain_bin_plan(sxmc_chain* c, int* nwarps_out, int* rows_per_chunk_out, int* nstages_out, int* smem_hist_out);
/* device time (ms) of THIS chain's last evaluate-only pass (sxmc_chain_nll_at) while
 * sxmc_enable_kernel_timing is on; -1 before the first */
int sxmc_chain_last_kernel_ms(sxmc_chain* c, float* bin_ms, float* nll_ms);
/* the ntuple rows: nparameters+1 floats each (src/nll_kernels.cpp:78-84) */
int sxmc_chain_read(sxmc_chain* c, float* rows_out, long max_rows);
int sxmc_chain_accepted(sxmc_chain* c, long* accepted_out);
/* device time of the step loop of the last run (the interval the reference times,
 * src/mcmc.cpp:259-380), from CUDA events on the chain's stream. */
int sxmc_chain_loop_ms(sxmc_chain* c, float* ms_out);
/* number of kernel launches (graph kernel nodes included) issued by the last run */
int sxmc_chain_launches(sxmc_chain* c, long* launches_out);
/* per-step trace of the proposed-point NLL (nsteps doubles), enabled before run */
int sxmc_chain_enable_trace(sxmc_chain* c, int on);
int sxmc_chain_read_trace(sxmc_chain* c, double* nll_trace_out, double* proposed_out,
                          unsigned nsteps);
/* Re-evaluate PDFs (unless systematics_fixed... always when `rebin`) and the NLL at a host
 * parameter vector after data were set by a run or by sxmc_chain_set_data: the replay hook
 * for per-step NLL parity. */
int sxmc_chain_set_data(sxmc_chain* c, const float* data, size_t data_nfloats);
int sxmc_chain_nll_at(sxmc_chain* c, const double* pars, int rebin, double* nll_out,
                      double* event_sum_out);
/* norms / lut of the last evaluation (host copies) */
int sxmc_chain_get_norms(sxmc_chain* c, unsigned* norms_out);
int sxmc_chain_get_lut(sxmc_chain* c, float* lut_out);
int sxmc_chain_get_bins(sxmc_chain* c, unsigned* bins_out);   /* [nsignals][total_nbins] */
/* chain execution mode: 0 = automatic (default): the monotone edge-search kernel for 1-D fits
 * whose systematics are all shift / scale / ctscale, else the persistent cooperative chain kernel
 * when a step streams < 64 MB (the fit is latency-bound), else a CUDA graph of the per-step
 * kernel sequence; 1 = plain stream launches (debugging / profiling); 2 = persistent kernel
 * or fail; 3 = CUDA graph; 4 = monotone kernel or fail.  Sample-sharded chains
 * (sxmc_chain_set_comm) run 1 or 3 only. */
int sxmc_chain_set_mode(sxmc_chain* c, int mode);
/* what the last sxmc_chain_advance used: 1 graph / stream launches, 2 persistent, 4 monotone */
int sxmc_chain_last_path(sxmc_chain* c, int* path_out);
/* profiling hook of the persistent kernel: CTA 0 stamps %globaltimer at 8 phase boundaries of
 * the first 64 steps of each launch ([64][8] ns), followed at index 512 by [ncta][4] stamps of
 * every CTA in step 20; out64x8 (may be NULL, else 2048 entries) receives the last launch's. */
int sxmc_chain_phase_stamps(sxmc_chain* c, int enable, unsigned long long* out64x8);

/* ---- ensemble loop pieces (src/sxmc.cpp:62-134), SURVEY 8(f)-1,2 ------------------------ */
typedef struct sxmc_generator sxmc_generator;
/* make_fake_dataset (src/generator.cpp:10-48) in two parts.  create = what is the same for
 * every experiment: Signal::get_efficiency (src/signal.cpp:172-197) and the histogram
 * EvalHist::RandomSample samples from (src/pdfz.cpp:817-832, CreateHistogram :496-592), both at
 * `syst_vals` (the systematics' means), evaluated for all signals in one batched launch.
 * `uppers`/`lowers` (nlimits = 0 or >= nobservables) are the observables' limits of the
 * rejection test (:851-855).  The handles are borrowed; their bound buffers are not touched. */
int sxmc_generator_create(sxmc_hist* const* hs, int n, const double* nexpected, const unsigned* n_mc,
                          const double* syst_vals, int nsyst_vals, const float* uppers,
                          const float* lowers, int nlimits, sxmc_generator** out);
int sxmc_generator_destroy(sxmc_generator* g);
/* eff[j] = norm_j / n_mc_j (src/signal.cpp:196) */
int sxmc_generator_efficiency(sxmc_generator* g, double* eff_out);
/* sample = the per-experiment part (src/pdfz.cpp:834-919): per signal observed =
 * Poisson(nexpected*eff) or nint, then inversion of the bin CDF + uniform inside the bin
 * (ROOT TH1::GetRandom / TH2::GetRandom2 / TH3::GetRandom3 restated on Philox4x32-10; key =
 * (seed, stream_id = experiment index)).  events_out receives rows of nobservables+1 floats,
 * signal after signal (NULL: count only); *nfloats_out = floats generated. */
int sxmc_generator_sample(sxmc_generator* g, int poisson, uint64_t seed, uint64_t stream_id,
                          float* events_out, size_t max_floats, unsigned* observed_out,
                          size_t* nfloats_out);
/* sxmc::errors::Contour (src/error_estimators/contour.cpp:18-69) on LikelihoodSpace::get_contour
 * (src/likelihood.cpp:88-101) as device reductions over rows [nrows][nparameters+1] (DEVICE
 * pointer; last column = NLL): delta = 0.5*ChisquareQuantile(cl,1); per parameter the extents
 * of the rows with NLL-min < delta, and the point estimate (mid-range of the rows within 0.13
 * of the minimum, widened x5 until one qualifies).  The float->text->double round trip of the
 * reference's selection strings ("likelihood+" << -lmin << "<" << delta) is reproduced. */
int sxmc_rows_contour(const float* rows_dev, long nrows, int nparameters, float delta,
                      float* lower_out, float* upper_out, float* point_out, float* lmin_out,
                      long* ncontour_out);
/* the same on a chain's resident rows after sxmc_chain_run / sxmc_chain_end */
int sxmc_chain_contour(sxmc_chain* c, float delta, float* lower_out, float* upper_out,
                       float* point_out, float* lmin_out, long* ncontour_out);

/* ---- chain batch: the fits of an ensemble (src/sxmc.cpp:62-134: nexperiments x MCMC::operator()
 *      over the same signals) advanced in lockstep by ONE cooperative launch.  Up to 148 chains
 *      (one per SM) share the Monte-Carlo samples, which stay resident in shared memory; each
 *      chain has its own data events, random stream and state.  Small fits only (the samples of
 *      a CTA must fit in shared memory: workloads A, B, E); others fail with SXMC_ENOTSMALL and
 *      run one chain at a time through sxmc_chain_run.  `cfg` as for sxmc_chain_create. ------- */
typedef struct sxmc_batch sxmc_batch;
int sxmc_batch_create(const sxmc_chain_config* cfg, int max_chains, sxmc_batch** out);
int sxmc_batch_destroy(sxmc_batch* b);
int sxmc_batch_set_seed(sxmc_batch* b, int slot, uint64_t seed, uint64_t stream_id);
int sxmc_batch_set_tapes(sxmc_batch* b, int slot, const double* normals, size_t n_normals,
                         const double* uniforms, size_t n_uniforms);              /* test hook */
/* the chain's data events: host rows of nobservables+1 floats (SetEvalPoints, src/mcmc.cpp:235) */
int sxmc_batch_set_data(sxmc_batch* b, int slot, const float* data, size_t data_nfloats);
/* ... or a fake dataset drawn on the device and consumed there (no host round trip) */
int sxmc_batch_generate_data(sxmc_batch* b, int slot, sxmc_generator* g, int poisson, uint64_t seed,
                             uint64_t stream_id, unsigned* observed_out);
int sxmc_batch_enable_trace(sxmc_batch* b, int on);
/* MCMC::operator() (src/mcmc.cpp:143-387) for slots [0, nchains): initial evaluation, nsteps
 * steps with the two burn-in re-tunes, rows kept on the device. */
int sxmc_batch_run(sxmc_batch* b, int nchains, unsigned nsteps, float burnin_fraction, int debug_mode);
int sxmc_batch_nrows(sxmc_batch* b, int slot, long* nrows_out, long* accepted_out);
int sxmc_batch_read(sxmc_batch* b, int slot, float* rows_out, long max_rows);
int sxmc_batch_read_trace(sxmc_batch* b, int slot, double* nll_trace_out, double* proposed_out,
                          unsigned nsteps);
int sxmc_batch_contour(sxmc_batch* b, int slot, float delta, float* lower_out, float* upper_out,
                       float* point_out, float* lmin_out, long* ncontour_out);
/* device time of the last sxmc_batch_run launch (all chains, all steps) */
int sxmc_batch_loop_ms(sxmc_batch* b, float* ms_out);
/* which kernel the last run used: 1 = monotone edge search (1-D fits whose systematics are all
 * shift / scale / ctscale: bin edges are bisected over the sorted samples), 2 = sample by sample;
 * *max_edges_out = largest number of nominal bin edges in one CTA's sample range (path 1) */
int sxmc_batch_last_path(sxmc_batch* b, int* path_out, int* max_edges_out);
/* profiling hook: every CTA stamps %globaltimer at the phase boundaries of super-step 20
 * ([148][8] ns: start, bin done, barrier out, lookup done, tail done, barrier out) */
int sxmc_batch_phase_stamps(sxmc_batch* b, int enable, unsigned long long* out148x8);

/* ---- multi-GPU (no reference counterpart) ------------------------------- */
int sxmc_comm_unique_id(void* id128_out);                 /* ncclGetUniqueId, 128 bytes */
int sxmc_comm_create(const void* id128, int nranks, int rank, sxmc_comm** out);
int sxmc_comm_destroy(sxmc_comm* comm);
/* in-place uint32 sum over ranks of a device array (used by tests and by the chain) */
int sxmc_comm_allreduce_u32(sxmc_comm* comm, unsigned* dev, size_t count);

/* Sample sharding over NVLink peer memory instead of NCCL: each rank exports a window holding
 * its partial histograms (a 64-byte cudaIpcMemHandle_t), the ranks exchange the handles by any
 * means and connect; from then on the per-step exchange is ONE kernel that publishes a flag to
 * every peer, waits for theirs and sums the N partials with 128-bit peer loads
 * (p2p_kernels.cu).  All ranks must issue the same sequence of evaluations.  One box, <= 8 ranks. */
int sxmc_chain_p2p_export(sxmc_chain* c, void* handle64_out);
int sxmc_chain_p2p_connect(sxmc_chain* c, const void* handles, int nranks, int rank);
/* exchanges completed so far; aborted != 0 after a peer failed to show up within ~2 s */
int sxmc_chain_p2p_status(sxmc_chain* c, unsigned* exchanges_out, int* aborted_out);
/* profiling hook: nanoseconds (device globaltimer), summed over all exchanges so far, the
 * exchange kernel spent [0] publishing its flag and waiting for every peer's partial, [1]
 * summing (until the last block is done; fused form: the rank-sharded lookup), [2] closing the step
 * (second flag round); fused form only, block 0: [3] peers' counts gathered and per-signal constants ready, [4] bins evaluated, [5] block sum, [6] partial sums stored everywhere, fenced, ticket taken.  ns_out7: seven doubles. */
int sxmc_chain_p2p_clocks(sxmc_chain* c, double* ns_out7);

#ifdef __cplusplus
}
#endif
#endif /* SXMC_B200_H */
