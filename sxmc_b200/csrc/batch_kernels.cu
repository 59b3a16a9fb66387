// batch_kernels.cu -- MANY chains over the SAME Monte-Carlo samples in ONE cooperative launch:
// the ensemble loop of the reference (src/sxmc.cpp:62-134 runs nexperiments fits one after the
// other on the same signals) turned sideways.
//
// A single small fit is latency-bound: a step is two grid barriers and a serial tail around a few
// microseconds of work (chain_kernels.cu).  The experiments of an ensemble share the samples and
// differ only in data events, random stream and chain state, so up to one chain per SM advance in
// lockstep here; a "super-step" is one MCMC step of every chain:
//
//   bin      the CTA's share of the samples is resident in shared memory (loaded once per
//            launch with bulk async copies); for every chain it is transformed under THAT
//            chain's proposed systematics and binned into a shared-memory histogram, GROUP
//            chains per pass, then merged into the chain's global histogram.  The samples are
//            read from HBM once per launch, not once per step and chain.
//   -- grid barrier --
//   lookup   CTA e owns chain e: PDF lookup + event likelihood over the chain's own events,
//   + tail   block reduction, NLL total, Metropolis, row append, re-tune, next proposal
//            (finish_step, the same code as the one-chain kernels), histogram cleared.
//   -- grid barrier --
//
// The barriers and the tail are paid once per super-step instead of once per chain step.
#include "common.cuh"
#include "kernels.h"
#include "bin_device.cuh"
#include "nll_device.cuh"
#include "grid_sync.cuh"
#include "batch_device.cuh"

#include <cstring>

namespace sxmc {

// run_push4 with a short straight-line slow path: the run is flushed, three samples go
// straight to the histogram, the fourth starts the new run.  (A warp executes the slow path
// whenever ANY lane crosses a bin edge, so its length matters wherever bins are sparse.)
__device__ __forceinline__ void run_push4_direct(RunAcc& r, const int (&b)[4], unsigned* hist, int total_nbins) {
  if (b[0] == r.cur && b[1] == r.cur && b[2] == r.cur && b[3] == r.cur) {
    r.cnt += 4;
  } else {
    run_flush<true>(r, hist, total_nbins);
#pragma unroll
    for (int k = 0; k < 3; ++k)
      if (b[k] >= 0) { r.norm += 1; if (b[k] < total_nbins) atomicAdd(hist + b[k], 1u); }
    r.cur = b[3];
    r.cnt = 1;
  }
}

template <int NOBS, int NS, int NOPS, unsigned CODES>
__global__ void __launch_bounds__(FAST_THREADS, 1)
chain_batch_kernel(const __grid_constant__ FastParams P, const __grid_constant__ BatchParams Q) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long s_bar[FAST_WARPS * FAST_MAX_STAGES];
  __shared__ BatchShared S;
  __shared__ double s_par[BATCH_MAX_CHAINS][FAST_MAX_OPS];   // every chain's systematic coefficients
  __shared__ unsigned s_norm[BATCH_MAX_GROUP];
  unsigned* s_hist = reinterpret_cast<unsigned*>(smem_raw);

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const unsigned G = gridDim.x;
  const int ns = Q.ns, nsources = Q.nsources, nch = Q.nchains;
  const int tnb = P.geo.total_nbins;
  const Geo& geo = P.geo;
  unsigned* const bar = Q.sync + 0;
  unsigned* const abort_flag = Q.sync + 3;
  unsigned bar_count = 0;
  if (tid < BATCH_MAX_GROUP) s_norm[tid] = 0u;

  // ---- this CTA's resident samples: one job (job-aligned partition), chunks cut over the warps
  FastCtx cx;
  fast_ctx_init<NS>(cx, P, smem_raw, s_bar);
  unsigned long long t0, t1;
  int job_j = 0;
  fast_partition(P, t0, t1, job_j);
  const bool have_work = t0 < t1;
  const int R = P.R;
  const unsigned col_bytes = (unsigned)R * 512u, stage_bytes = NS * col_bytes;
  int nmine = 0;
  if (have_work) {
    const FastJob& job = P.job[job_j];
    const unsigned long long jbeg = P.prefix[job_j];
    const unsigned long long n = t1 - t0;
    const unsigned long long ca = (t0 - jbeg) + n * (unsigned)warp / FAST_WARPS;
    const unsigned long long cb = (t0 - jbeg) + n * (unsigned)(warp + 1) / FAST_WARPS;
    nmine = (int)(cb - ca);                     // <= P.S (checked on the host)
    if (lane == 0) {
      for (int i = 0; i < nmine; ++i) {
        const unsigned b = cx.my_bar0 + 8u * i;
        const unsigned dst = cx.my_stage0 + i * stage_bytes;
        mbar_expect_tx(b, stage_bytes);
#pragma unroll
        for (int k = 0; k < NS; ++k)
          bulk_g2s(dst + k * col_bytes, reinterpret_cast<const unsigned char*>(job.col[k]) + (ca + i) * col_bytes,
                   col_bytes, b);
      }
    }
  }

  // ---- the chain this CTA owns in the lookup/tail phase
  const int mine = blockIdx.x < (unsigned)nch ? (int)blockIdx.x : -1;
  batch_owner_init(S, Q, mine, tid);
  // the samples have landed (waited for once; the stages are never refilled)
  for (int i = 0; i < nmine; ++i) mbar_wait(cx.my_bar0 + 8u * i, 0u);
  __syncthreads();

  const int nsyst = have_work ? P.job[job_j].nsyst : 0;
  const unsigned total_super = Q.nsteps + (Q.do_init ? 1u : 0u);

#define BK_STAMP(slot)                                                              \
  do {                                                                              \
    if (Q.stamps && tid == 0 && s == 20u) Q.stamps[blockIdx.x * 8 + (slot)] = globaltimer_ns(); \
  } while (0)
  for (unsigned s = 0; s < total_super; ++s) {
    const int mode = (Q.do_init && s == 0) ? NLL_MODE_INIT : NLL_MODE_STEP;
    const bool bin_now = Q.rebin || mode == NLL_MODE_INIT;
    BK_STAMP(0);

    // ---- bin: every chain's histogram from the resident samples ------------------------------
    if (bin_now) {
      if (have_work) {
        const FastJob& job = P.job[job_j];
        for (int q = tid; q < nch * FAST_MAX_OPS; q += FAST_THREADS) {
          const int e = q / FAST_MAX_OPS, k = q % FAST_MAX_OPS;
          double par = 0.0;
          if (k < nsyst && job.npars[k]) par = __ldcg(&Q.st[e].proposed[nsources + job.pidx[k]]);
          s_par[e][k] = par;
        }
      }
      for (int g0 = 0; g0 < nch; g0 += Q.group) {
        const int hn = min(Q.group, nch - g0);
        for (int b = tid; b < hn * tnb; b += FAST_THREADS) s_hist[b] = 0u;
        __syncthreads();
        if (have_work) {
          const FastJob& job = P.job[job_j];
          for (int h = 0; h < hn; ++h) {
            int op_code[FAST_MAX_OPS];
            double op_c0[FAST_MAX_OPS], op_onep[FAST_MAX_OPS];
#pragma unroll
            for (int k = 0; k < FAST_MAX_OPS; ++k) {
              op_code[k] = 0; op_c0[k] = 0.0; op_onep[k] = 1.0;
              if (k < nsyst) {
                op_code[k] = job.code[k];
                op_c0[k] = __dadd_rn(0.0, __dmul_rn(s_par[g0 + h][k], 1.0));   // src/pdfz.cpp:310-314
                op_onep[k] = __dadd_rn(1.0, op_c0[k]);
              }
            }
            unsigned* hist = s_hist + h * tnb;
            RunAcc acc;
            acc.norm = 0; acc.cur = -1; acc.cnt = 0;
            // one row of 4 samples per lane, sample by sample (the reference's arithmetic)
            auto bin_row = [&](unsigned src) {
              double f[NS][4];
#pragma unroll
              for (int k = 0; k < NS; ++k) {
                const float4 v = lds128(src + k * col_bytes);
                f[k][0] = (double)v.x; f[k][1] = (double)v.y; f[k][2] = (double)v.z; f[k][3] = (double)v.w;
              }
              if constexpr (NOPS < 0) {
#pragma unroll
                for (int k = 0; k < FAST_MAX_OPS; ++k)
                  if (k < nsyst) apply_op<NS>(f, op_code[k], op_c0[k], op_onep[k]);
              } else {
                if constexpr (NOPS > 0) apply_static<NS, (CODES >> 0) & 255>(f, op_c0[0], op_onep[0]);
                if constexpr (NOPS > 1) apply_static<NS, (CODES >> 8) & 255>(f, op_c0[1], op_onep[1]);
                if constexpr (NOPS > 2) apply_static<NS, (CODES >> 16) & 255>(f, op_c0[2], op_onep[2]);
                if constexpr (NOPS > 3) apply_static<NS, (CODES >> 24) & 255>(f, op_c0[3], op_onep[3]);
              }
              int bq[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                double x[NOBS];
#pragma unroll
                for (int o = 0; o < NOBS; ++o) x[o] = f[o][k];
                bq[k] = flat_bin<NOBS>(x, geo);
              }
              run_push4_direct(acc, bq, hist, tnb);
            };
#pragma unroll 1
            for (int i = 0; i < nmine; ++i) {
              const unsigned src0 = cx.my_stage0 + i * stage_bytes + (unsigned)lane * 16u;
#pragma unroll 1
              for (int r = 0; r < R; ++r) bin_row(src0 + (unsigned)r * 512u);
            }
            run_flush<true>(acc, hist, tnb);
            const unsigned wnorm = __reduce_add_sync(0xffffffffu, acc.norm);
            if (lane == 0 && wnorm) atomicAdd(&s_norm[h], wnorm);
          }
        }
        __syncthreads();
        if (have_work) {
          for (int b = tid; b < hn * tnb; b += FAST_THREADS) {
            const unsigned cnt = s_hist[b];
            if (cnt) {
              const int h = b / tnb, k = b - h * tnb;
              atomicAdd(Q.hist + (size_t)(g0 + h) * Q.hist_words + (size_t)job_j * tnb + k, cnt);
            }
          }
          if (tid < hn) {
            const unsigned cn = s_norm[tid];
            if (cn) atomicAdd(Q.hist + (size_t)(g0 + tid) * Q.hist_words + (size_t)ns * tnb + job_j, cn);
            s_norm[tid] = 0u;
          }
        }
        __syncthreads();
      }
      BK_STAMP(1);
      cta_arrive(bar);
      ++bar_count;
      if (!cta_wait_ge(bar, bar_count * G, abort_flag, &S.ok)) break;
      BK_STAMP(2);
    }

    // ---- lookup + event likelihood + tail of the chain this CTA owns ---------------------------
    if (mine >= 0) batch_lookup_tail(S, Q, mine, mode, tnb, tid, Q.rebin && s + 1 < total_super);
    BK_STAMP(4);
    if (Q.rebin && s + 1 < total_super) {      // proposals published, histograms cleared
      cta_arrive(bar);
      ++bar_count;
      if (!cta_wait_ge(bar, bar_count * G, abort_flag, &S.ok)) break;
    }
    BK_STAMP(5);
  }
}

// precondition of the monotone path: the column holds no NaN and no -inf
__global__ void column_check_kernel(const float* __restrict__ col, unsigned long long n, unsigned* flag) {
  bool bad = false;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x)
    bad |= !(col[i] > __int_as_float(0xff800000));
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1u);
}

int column_has_nan_or_neginf(const float* col_dev, unsigned long long n, bool* out, cudaStream_t stream) {
  *out = false;
  if (n == 0) return SXMC_OK;
  unsigned* flag = nullptr;
  SXMC_CUDA(cudaMalloc(&flag, sizeof(unsigned)));
  cudaError_t e = cudaMemsetAsync(flag, 0, sizeof(unsigned), stream);
  unsigned long long blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (e == cudaSuccess) {
    column_check_kernel<<<(unsigned)blocks, 256, 0, stream>>>(col_dev, n, flag);
    e = cudaGetLastError();
  }
  unsigned h = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&h, flag, sizeof(unsigned), cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  cudaFree(flag);
  if (e != cudaSuccess) return cuda_fail(e, "column_has_nan_or_neginf", __FILE__, __LINE__);
  *out = h != 0;
  return SXMC_OK;
}

// ---------------------------------------------------------------------------------------------
typedef void (*BatchKernel)(const FastParams, const BatchParams);
struct BatchEntry {
  int nobs, ns, nops;
  unsigned codes;
  BatchKernel k;
};
#define FE(NOBS, NS, NOPS, CODES) { NOBS, NS, NOPS, (CODES), chain_batch_kernel<NOBS, NS, NOPS, (CODES)> }
static const BatchEntry g_batch_table[] = { SXMC_FAST_TABLE(FE) };
#undef FE

constexpr size_t BK_STATIC_RESERVE = 56 * 1024;   // static shared memory of the kernel above

// Shared-memory plan: GROUP histograms + resident stages.  Fails when the samples of a CTA do
// not fit (the fit is not "small": run its chains one at a time with the streaming kernels).
int plan_chain_batch(const BinLaunch& L0, int nchains, BatchPlan* plan) {
  if (!L0.fast || L0.njobs > FAST_MAX_JOBS || nchains < 1 || nchains > BATCH_MAX_CHAINS) {
    set_error("chain batch: fit shape not supported (fast binning path, <= 32 signals, <= 148 chains)");
    return SXMC_ENOTSMALL;
  }
  int dev = 0, sms = 0;
  SXMC_CUDA(cudaGetDevice(&dev));
  SXMC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  if (nchains > sms) { set_error("chain batch: more chains than SMs"); return SXMC_EINVAL; }
  for (int group = BATCH_MAX_GROUP; group >= 1; group >>= 1) {
    BinLaunch L = L0;
    L.nwarps = FAST_WARPS;        // the batch kernel runs full CTAs; its pipeline is re-planned below
    if (!bin_fast_plan(L.nslots, group * L.max_total_nbins, BK_STATIC_RESERVE, &L.rows_per_stage, &L.nstages,
                       &L.stage_off, &L.smem_bytes))
      continue;
    FastParams P;
    unsigned long long total = 0;
    build_fast_params(L, 0, L.njobs, sms, P, &total);
    if (!P.aligned) continue;
    // residency: every warp's chunk range must fit in its S stages
    bool ok = true;
    for (int j = 0; j < L.njobs && ok; ++j) {
      const unsigned long long nj = P.prefix[j + 1] - P.prefix[j];
      const unsigned cj = P.cta_first[j + 1] - P.cta_first[j];
      if (!nj) continue;
      const unsigned long long per_cta = (nj + cj - 1) / cj;
      const unsigned long long per_warp = (per_cta + FAST_WARPS - 1) / FAST_WARPS;
      if (per_warp > (unsigned long long)L.nstages) ok = false;
    }
    if (!ok) continue;
    plan->launch = L;
    plan->group = group;
    plan->grid = sms;
    return SXMC_OK;
  }
  set_error("chain batch: the samples do not fit in shared memory (not a small fit)");
  return SXMC_ENOTSMALL;
}

int launch_chain_batch(const BatchPlan& plan, const BatchParams& Q0, cudaStream_t stream) {
  const BinLaunch& L = plan.launch;
  BatchKernel fk = nullptr;
  for (const BatchEntry& e : g_batch_table)
    if (e.nobs == L.nobs && e.ns == L.nslots && e.nops == (L.jit ? -1 : L.prog_nops) &&      // (a program compiled
        (L.jit || L.prog_nops <= 0 || e.codes == L.prog_codes))                              // by NVRTC exists for bin_fast_kernel only)
      fk = e.k;
  if (!fk) { set_error("chain batch: no instantiation for this shape"); return SXMC_ENOTSMALL; }
  FastParams P;
  unsigned long long total = 0;
  build_fast_params(L, 0, L.njobs, plan.grid, P, &total);
  SXMC_CUDA(cudaFuncSetAttribute((const void*)fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes));
  int occ = 0;
  SXMC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)fk, FAST_THREADS, L.smem_bytes));
  if (occ < 1) { set_error("chain batch kernel does not fit on an SM"); return SXMC_ENOTSMALL; }
  BatchParams Q = Q0;
  Q.group = plan.group;
  void* args[2] = {(void*)&P, (void*)&Q};
  SXMC_CUDA(cudaLaunchCooperativeKernel((const void*)fk, dim3((unsigned)plan.grid), dim3(FAST_THREADS), args,
                                        L.smem_bytes, stream));
  return SXMC_OK;
}

}  // namespace sxmc
