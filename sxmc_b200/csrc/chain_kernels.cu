// chain_kernels.cu -- the persistent chain kernel: MANY MCMC steps in ONE cooperative launch.
//
// For small fits (workloads A, B) a step is ~12 MB of streaming and ~10^4 events: with one
// launch per kernel the step time is launch latency and dependent global round trips, not
// work.  Here one CTA per SM stays resident for the whole segment of the chain and the
// phases of a step (reference src/mcmc.cpp:261-348) are separated by grid barriers:
//
//   bin      every CTA transforms + bins its share of the samples (fast_bin_pass, the same
//            code as bin_fast_kernel) into the step's histogram buffer.  A warp whose chunks
//            all fit in its pipeline stages loads them ONCE: for A and B the Monte-Carlo
//            samples then live in shared memory across steps and the bin phase reads no
//            global memory at all.
//   -- grid barrier --
//   lookup   PDF lookup + event likelihood over all events (grid-stride), per-CTA partial sums;
//            the other histogram buffer is cleared for the next step.
//   -- arrive; CTA 0 waits --
//   finish   CTA 0 totals the partials in a fixed order and runs the step tail on its
//            shared-memory chain state (NLL total, Metropolis, row append, re-tune, proposal),
//            publishes the next proposed vector and releases the other CTAs.
//
// Launched with cudaLaunchCooperativeKernel (co-residency is checked by the runtime); every
// spin has a clock64 timeout that raises an abort flag instead of hanging the device.
#include "common.cuh"
#include "kernels.h"
#include "bin_device.cuh"
#include "nll_device.cuh"
#include "grid_sync.cuh"

#include <cstring>

namespace sxmc {

constexpr int PK_MAX_SIGNALS = 128;
// phase stamps of CTA 0 (debug/profiling): [step][8] ns, first PK_STAMP_STEPS steps of a launch
constexpr unsigned PK_STAMP_STEPS = 64;
#define PK_STAMP(slot)                                                                       \
  do {                                                                                       \
    if (Q.stamps && blockIdx.x == 0 && threadIdx.x == 0 && s < PK_STAMP_STEPS)               \
      Q.stamps[s * 8 + (slot)] = globaltimer_ns();                                           \
  } while (0)
// per-CTA stamps of step PK_CTA_STEP: stamps[512 + blockIdx * 4 + slot]
constexpr unsigned PK_CTA_STEP = 20;
#define PK_CTA_STAMP(slot)                                                                    \
  do {                                                                                        \
    if (Q.stamps && threadIdx.x == 0 && s == PK_CTA_STEP && blockIdx.x < 256)                 \
      Q.stamps[512 + blockIdx.x * 4 + (slot)] = globaltimer_ns();                             \
  } while (0)

template <int NOBS, int NS, bool SMEM, int NOPS, unsigned CODES>
__global__ void __launch_bounds__(FAST_THREADS, 1)
chain_persistent_kernel(const __grid_constant__ FastParams P, const __grid_constant__ PersistParams Q) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ SystRt s_syst[SXMC_MAX_SYST];
  __shared__ __align__(8) unsigned long long s_bar[FAST_WARPS * FAST_MAX_STAGES];
  __shared__ double s_c[PK_MAX_SIGNALS];
  __shared__ double s_bn[PK_MAX_SIGNALS];
  __shared__ double s_nexp[PK_MAX_SIGNALS];
  __shared__ double s_binvol[PK_MAX_SIGNALS];
  __shared__ const int* s_rb[PK_MAX_SIGNALS];
  __shared__ unsigned s_nmc[PK_MAX_SIGNALS];
  __shared__ short s_sid[PK_MAX_SIGNALS];
  __shared__ double s_warp[FAST_THREADS / 32];
  __shared__ double s_term[PK_MAX_SIGNALS + SXMC_MAX_PARAMETERS];
  __shared__ FinState F;
  __shared__ int s_ok;
  __shared__ double s_z[SXMC_MAX_PARAMETERS];   // this step's normals / uniform, drawn ahead of the tail
  __shared__ double s_u;
  __shared__ unsigned s_norm;
  if (threadIdx.x == 0) s_norm = 0u;
  unsigned* s_hist = reinterpret_cast<unsigned*>(smem_raw);

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const unsigned G = gridDim.x;
  const int ns = Q.ns, np = Q.np, nsources = Q.nsources;
  const int tnb = P.geo.total_nbins;
  unsigned* const bar1 = Q.sync + 0;
  unsigned* const bar2 = Q.sync + 1;
  unsigned* const go = Q.sync + 2;
  unsigned* const abort_flag = Q.sync + 3;

  FastCtx cx;
  fast_ctx_init<NS>(cx, P, smem_raw, s_bar);
  {  // does every chunk of this warp fit in its stages?  (same partition as fast_bin_pass)
    unsigned long long t0, t1, mine = 0;
    int j = 0;
    fast_partition(P, t0, t1, j);
    while (t0 < t1) {
      while (P.prefix[j + 1] <= t0) ++j;
      const unsigned long long jend = P.prefix[j + 1];
      const unsigned long long tend = t1 < jend ? t1 : jend;
      const unsigned long long n = tend - t0;
      mine += n * (unsigned)(warp + 1) / FAST_WARPS - n * (unsigned)warp / FAST_WARPS;
      t0 = tend;
    }
    cx.resident = Q.rebin && mine <= (unsigned long long)P.S;
  }

  prefetch_state(F, Q.st, np, tid, FAST_THREADS);
  if (tid < ns) {
    const NllSignal sg = Q.sigs[tid];
    s_nexp[tid] = sg.nexpected;
    s_binvol[tid] = sg.bin_volume;
    s_rb[tid] = sg.read_bins;
    s_nmc[tid] = sg.n_mc;
    s_sid[tid] = sg.source_id;
  }
  __syncthreads();

  const long long gtid = blockIdx.x * (long long)FAST_THREADS + tid;
  const long long gstride = (long long)G * FAST_THREADS;
  const long long etid = (long long)tid * G + blockIdx.x;   // events spread over all CTAs

  for (unsigned s = 0; s < Q.nsteps; ++s) {
    const unsigned istep = Q.step0 + s;
    const int b = Q.rebin ? (int)((istep + 1) & 1u) : 0;
    unsigned* const hb = Q.hist_buf[b];

    // ---- bin ------------------------------------------------------------------------------
    PK_STAMP(0);
    PK_CTA_STAMP(0);
    if (Q.rebin) {
      fast_bin_pass<NOBS, NS, SMEM, NOPS, CODES>(
          P, cx, s_hist, s_syst, &s_norm, hb, F.proposed + nsources,
          (Q.stamps && s == PK_CTA_STEP && blockIdx.x < 148) ? Q.stamps + 1200 + blockIdx.x * 4 : nullptr);
      PK_STAMP(1);
      PK_CTA_STAMP(1);
      cta_arrive(bar1);
      if (!cta_wait_ge(bar1, (s + 1u) * G, abort_flag, &s_ok)) break;
      PK_STAMP(2);
      PK_CTA_STAMP(2);
      // the buffer the NEXT step bins into: last read in the previous step's lookup phase
      unsigned* const other = Q.hist_buf[1 - b];
      for (unsigned long long w = gtid; w < Q.hist_words; w += gstride) other[w] = 0u;
    }

    // ---- lookup + event likelihood ------------------------------------------------------------
    // CTA 0's last warp draws the step's random numbers now (Philox is stateless: counter =
    // step index), off the serial tail
    const bool pre_rng = blockIdx.x == 0 && !F.tape_normals && !F.tape_uniforms;
    if (pre_rng && warp == FAST_WARPS - 1) {
      for (int k = lane; k < np; k += 32)
        s_z[k] = F.jump_width[k] > 0 ? fin_normal(F, F.step, (unsigned)k, 1u) : 0.0;
      if (lane == 0) {
        const uint4 r = fin_random(F, F.step, 0u, 2u);
        s_u = u01(r.x, r.y);
      }
    }
    if (tid < ns) {
      const unsigned norm = __ldcg(hb + (size_t)ns * tnb + tid);
      const float eff = (float)(1.0 * norm / s_nmc[tid]);                    // nll_kernels.cpp:105
      const double rate = __dmul_rn(F.proposed[s_sid[tid]], s_nexp[tid]);
      s_c[tid] = __dmul_rn(rate, (double)eff);
      s_bn[tid] = __dmul_rn((double)norm, s_binvol[tid]);                    // pdfz.cpp:417
      s_term[tid] = __dmul_rn(rate, (double)norm) / (double)s_nmc[tid];      // nll_kernels.cpp:167-171
    }
    __syncthreads();
    double sum = 0;
    for (long long i = etid; i < Q.nevents; i += gstride) {
      double acc = 0;
      for (int j = 0; j < ns; ++j) {
        const int rb = s_rb[j][i];
        float v = 0.0f;
        if (rb >= 0 && rb < tnb) {
          v = pdf_value(__ldcg(hb + (size_t)j * tnb + rb), s_bn[j]);          // pdfz.cpp:429-432
          if (isnan(v)) v = 0.0f;                                            // nll_kernels.cpp:107
        }
        acc = __dadd_rn(acc, __dmul_rn(s_c[j], (double)v));
      }
      if (acc > 0) sum += log(acc);
    }
    const double bsum = block_sum<FAST_THREADS>(sum, s_warp);
    if (tid == 0) __stcg(Q.partials + blockIdx.x, bsum);
    PK_STAMP(3);
    PK_CTA_STAMP(3);
    cta_arrive(bar2);

    // ---- finish (CTA 0) / wait for the next proposal (others) ---------------------------------
    double* const next = Q.bcast + (size_t)((s + 1u) & 1u) * SXMC_MAX_PARAMETERS;
    if (blockIdx.x == 0) {
      if (!cta_wait_ge(bar2, (s + 1u) * G, abort_flag, &s_ok)) break;
      PK_STAMP(4);
      double acc = 0;
      for (int k = tid; k < (int)G; k += FAST_THREADS) acc += __ldcg(Q.partials + k);
      const double total = block_sum<FAST_THREADS>(acc, s_warp);
      if (tid == 0) s_warp[0] = total;
      for (int i = tid; i < np; i += FAST_THREADS) {       // Gaussian constraints (nll_kernels.cpp:176-185)
        const double sg = F.sigmas[i];
        double t = 0;
        if (sg > 0) {
          const double x = __dadd_rn(F.proposed[i], -F.means[i]) / sg;
          t = __dmul_rn(__dmul_rn(0.5, x), x);
        }
        s_term[ns + i] = t;
      }
      __syncthreads();
      PK_STAMP(5);
      if (tid < 64)
        finish_step2(F, Q.st, np, ns, nsources, s_warp[0], NLL_MODE_STEP, s_term, tid & 31, tid >> 5,
                    pre_rng ? s_z : nullptr, pre_rng ? &s_u : nullptr);
      __syncthreads();
      PK_STAMP(6);
      for (int k = tid; k < np; k += FAST_THREADS) __stcg(next + k, F.proposed[k]);
      __syncthreads();
      if (tid == 0) {
        __threadfence();
        st_release_u32(go, s + 1u);
      }
      PK_STAMP(7);
    } else {
      if (!cta_wait_ge(go, s + 1u, abort_flag, &s_ok)) break;
      for (int k = tid; k < np; k += FAST_THREADS) F.proposed[k] = __ldcg(next + k);
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------
typedef void (*PersistKernel)(const FastParams, const PersistParams);
struct PersistEntry {
  int nobs, ns, nops;
  unsigned codes;
  PersistKernel k[2];
};
#define FE(NOBS, NS, NOPS, CODES)                                                     \
  { NOBS, NS, NOPS, (CODES),                                                          \
    { chain_persistent_kernel<NOBS, NS, false, NOPS, (CODES)>,                        \
      chain_persistent_kernel<NOBS, NS, true, NOPS, (CODES)> } }
static const PersistEntry g_persist_table[] = { SXMC_FAST_TABLE(FE) };
#undef FE

constexpr size_t PK_STATIC_RESERVE = 32 * 1024;   // static shared memory of the kernel above

static int sms_for_plan() {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms > 0 ? sms : 148;
}

int launch_chain_persistent(const BinLaunch& L0, const PersistParams& Q, cudaStream_t stream) {
  if (!L0.fast || L0.njobs > FAST_MAX_JOBS || Q.ns > PK_MAX_SIGNALS) {
    set_error("persistent chain kernel: fit shape not supported");
    return SXMC_ESTATE;
  }
  BinLaunch L = L0;
  L.nwarps = FAST_WARPS;          // the persistent kernel runs full CTAs; its pipeline is re-planned below
  // its own shared-memory plan: the chain state and the signal tables take static space
  L.smem_hist = L0.smem_hist &&
                bin_fast_plan(L.nslots, L.max_total_nbins, PK_STATIC_RESERVE, &L.rows_per_stage,
                              &L.nstages, &L.stage_off, &L.smem_bytes);
  if (!L.smem_hist &&
      !bin_fast_plan(L.nslots, 0, PK_STATIC_RESERVE, &L.rows_per_stage, &L.nstages, &L.stage_off,
                     &L.smem_bytes)) {
    set_error("persistent chain kernel: no shared-memory plan");
    return SXMC_ESTATE;
  }
  PersistKernel fk = nullptr;
  for (const PersistEntry& e : g_persist_table)
    if (e.nobs == L.nobs && e.ns == L.nslots && e.nops == (L.jit ? -1 : L.prog_nops) &&      // (a program compiled
        (L.jit || L.prog_nops <= 0 || e.codes == L.prog_codes))                              // by NVRTC exists for bin_fast_kernel only)
      fk = e.k[L.smem_hist ? 1 : 0];
  if (!fk) { set_error("persistent chain kernel: no instantiation for this shape"); return SXMC_ESTATE; }
  FastParams P;
  unsigned long long total = 0;
  build_fast_params(L, 0, L.njobs, sms_for_plan(), P, &total);
  SXMC_CUDA(cudaFuncSetAttribute((const void*)fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes));
  int occ = 0;
  SXMC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)fk, FAST_THREADS, L.smem_bytes));
  if (occ < 1) { set_error("persistent chain kernel does not fit on an SM"); return SXMC_ESTATE; }
  int dev = 0, sms = 0;
  SXMC_CUDA(cudaGetDevice(&dev));
  SXMC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  PersistParams Qc = Q;
  void* args[2] = {(void*)&P, (void*)&Qc};
  SXMC_CUDA(cudaLaunchCooperativeKernel((const void*)fk, dim3((unsigned)sms), dim3(FAST_THREADS), args,
                                        L.smem_bytes, stream));
  return SXMC_OK;
}

}  // namespace sxmc
