// bin_stream.cuh -- the transform + binning pass for launches that stream far more than the
// chip holds (workloads C, D: >= 64 chunks per warp), with the tail balanced dynamically.
//
// Why: with one static chunk range per CTA (bin_fast_kernel) the %globaltimer stamps of workloads
// C and D show every SM streaming at its own pace -- CTA lifetimes of 229 / 246 / 296 us
// (min / median / max) for the same 10.8 MB each -- so 15 % of the SM-time of a launch is spent
// waiting for the slowest CTA (profiles/r02_bin_phases.txt).  Here
//   * every CTA owns ONE job (signal) and a static range of it sized for ~70 % of its fair share,
//     binned into the CTA's shared-memory histogram as before;
//   * the rest of every job is cut into tiles of T chunks handed out through one global counter
//     to whichever WARP runs dry first; a tile's samples go to the job's global histogram with
//     L2 atomics (after per-thread run-length aggregation: a few hundred thousand per launch);
//   * nothing in steady state is CTA-wide: a warp's bulk-copy pipeline runs on from its static
//     range into its tiles without draining, the per-stage job id travels in shared memory next
//     to the stage, and the shared-memory histogram is merged slice by slice by the warps once
//     all 16 have left the static phase.
// Arithmetic, layout and run-length aggregation are those of bin_device.cuh; results are the
// same integers.
#pragma once
#include "bin_device.cuh"

namespace sxmc {

constexpr int STREAM_DYN = 0x100;     // stage metadata: job | STREAM_DYN for a tile of the dynamic part
constexpr int STREAM_END = -1;

__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

template <int NOBS, int NS, bool SMEM, int NOPS, unsigned CODES>
__global__ void __launch_bounds__(FAST_THREADS, 1)
bin_stream_kernel(const __grid_constant__ StreamParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long s_bar[FAST_WARPS * FAST_MAX_STAGES];
  __shared__ int s_meta[FAST_WARPS * FAST_MAX_STAGES];
  __shared__ double s_c0[FAST_MAX_JOBS][FAST_MAX_OPS];
  __shared__ double s_onep[FAST_MAX_JOBS][FAST_MAX_OPS];
  __shared__ unsigned s_done;

  const Geo& geo = P.geo;
  const int R = P.R, S = P.S;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned col_bytes = (unsigned)R * 512u, stage_bytes = NS * col_bytes;
  const unsigned my_stage0 = smem_u32(smem_raw + P.stage_off) + (unsigned)(warp * S) * stage_bytes;
  const unsigned my_bar0 = smem_u32(&s_bar[warp * FAST_MAX_STAGES]);
  int* const my_meta = &s_meta[warp * FAST_MAX_STAGES];
  unsigned* const s_hist = reinterpret_cast<unsigned*>(smem_raw);
  unsigned long long* const xd = P.xdbg ? P.xdbg + 32ull * blockIdx.x : nullptr;
  if (xd && threadIdx.x == 0) xd[0] = gtimer();

  // ---- this CTA's job and this warp's static chunk range of it ------------------------------
  int home = 0;
  while (home + 1 < P.njobs && P.cta_first[home + 1] <= blockIdx.x) ++home;
  const unsigned ncta = P.cta_first[home + 1] - P.cta_first[home], kcta = blockIdx.x - P.cta_first[home];
  const unsigned long long ns_home = P.nstatic[home];
  const unsigned ca0 = (unsigned)(ns_home * kcta / ncta), ca1 = (unsigned)(ns_home * (kcta + 1ull) / ncta);
  const unsigned wa = ca0 + (unsigned)((unsigned long long)(ca1 - ca0) * (unsigned)warp / FAST_WARPS);
  const unsigned wb = ca0 + (unsigned)((unsigned long long)(ca1 - ca0) * (unsigned)(warp + 1) / FAST_WARPS);

  // ---- issue side (lane 0): a cursor over [static range] ++ [tiles from the global queue] -----
  unsigned rng_c = wa, rng_end = wb;     // job-relative chunk cursor of the current range
  int rng_job = home, rng_dyn = 0, exhausted = 0;
  unsigned pending = 0;                  // the next tile index, fetched one tile ahead
  const unsigned ntiles = P.tprefix[P.njobs];
  if (lane == 0) {
    for (int s = 0; s < S; ++s) mbar_init(my_bar0 + 8u * s, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    pending = atomicAdd(P.queue, 1u);
  }
  auto issue_next = [&](unsigned slot) {          // lane 0 only
    const unsigned bar = my_bar0 + 8u * slot;
    while (!exhausted && rng_c >= rng_end) {      // next range: a tile of the dynamic part
      if (pending >= ntiles) { exhausted = 1; break; }
      // tmul is coprime with ntiles: a bijection that scatters consecutive queue positions over the
      // dynamic regions (tmul == 1: tiles in address order)
      const unsigned t = (unsigned)(((unsigned long long)pending * P.tmul) % ntiles);
      pending = atomicAdd(P.queue, 1u);
      int j = 0;
      while (P.tprefix[j + 1] <= t) ++j;
      rng_job = j;
      rng_dyn = STREAM_DYN;
      rng_c = P.nstatic[j] + (t - P.tprefix[j]) * (unsigned)P.T;
      rng_end = min(rng_c + (unsigned)P.T, P.nchunks[j]);
    }
    if (exhausted) {
      my_meta[slot] = STREAM_END;
      mbar_arrive(bar);
      return;
    }
    my_meta[slot] = rng_job | rng_dyn;
    const unsigned dst = my_stage0 + slot * stage_bytes;
    const unsigned long long off = (unsigned long long)rng_c * col_bytes;
    mbar_expect_tx(bar, stage_bytes);
#pragma unroll
    for (int k = 0; k < NS; ++k)
      bulk_g2s(dst + k * col_bytes, reinterpret_cast<const unsigned char*>(P.job[rng_job].col[k]) + off, col_bytes, bar);
    ++rng_c;
  };
  if (lane == 0)
    for (int s = 0; s < S; ++s) issue_next((unsigned)s);

  // ---- every job's systematic coefficients, once (p = 0 + par * pow(x, 0), src/pdfz.cpp:310-314)
  for (int q = threadIdx.x; q < P.njobs * FAST_MAX_OPS; q += FAST_THREADS) {
    const int j = q / FAST_MAX_OPS, k = q % FAST_MAX_OPS;
    const FastJob& job = P.job[j];
    double c0 = 0.0, onep = 1.0;
    if (k < job.nsyst) {
      const double par = job.npars[k] ? job.params[(long long)job.pidx[k] * job.param_stride] : 0.0;
      c0 = __dadd_rn(0.0, __dmul_rn(par, 1.0));
      onep = __dadd_rn(1.0, c0);
    }
    s_c0[j][k] = c0;
    s_onep[j][k] = onep;
  }
  if (SMEM)
    for (int b = threadIdx.x; b < geo.total_nbins; b += FAST_THREADS) s_hist[b] = 0;
  if (threadIdx.x == 0) s_done = 0u;
  __syncthreads();
  if (xd && threadIdx.x == 0) xd[1] = gtimer();

  // ---- consume side ----------------------------------------------------------------------------
  int cur = -2;                          // metadata of the range the accumulator belongs to
  int op_code[FAST_MAX_OPS];
  double op_c0[FAST_MAX_OPS], op_onep[FAST_MAX_OPS];
  int nsyst = 0;
  unsigned* gbins = nullptr;             // the current job's global histogram
  RunAcc acc;
  acc.norm = 0; acc.cur = -1; acc.cnt = 0;
  bool in_static = true;

  auto close_range = [&]() {             // flush the run and this range's in-domain count
    if (cur < 0) return;
    if (SMEM && !(cur & STREAM_DYN)) run_flush<true>(acc, s_hist, geo.total_nbins);
    else run_flush<false>(acc, gbins, geo.total_nbins);
    const unsigned wnorm = __reduce_add_sync(0xffffffffu, acc.norm);
    if (lane == 0 && wnorm) atomicAdd(P.job[cur & 0xff].norm, wnorm);
    acc.norm = 0; acc.cur = -1; acc.cnt = 0;
  };
  auto leave_static = [&]() {            // this warp's updates of the shared histogram are complete
    if (!in_static) return;
    in_static = false;
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
      atomicAdd(&s_done, 1u);
    }
    if (xd && lane == 0) xd[16 + warp] = gtimer();
  };

#pragma unroll 1
  for (unsigned it = 0;; ++it) {
    const unsigned slot = it % (unsigned)S;
    mbar_wait(my_bar0 + 8u * slot, (it / (unsigned)S) & 1u);
    if (xd && it == 0 && threadIdx.x == 0) xd[2] = gtimer();
    const int meta = my_meta[slot];
    if (meta != cur) {
      close_range();
      if (meta == STREAM_END) break;
      if (meta & STREAM_DYN) leave_static();
      cur = meta;
      const int j = meta & 0xff;
      nsyst = P.job[j].nsyst;
#pragma unroll
      for (int k = 0; k < FAST_MAX_OPS; ++k) {
        op_code[k] = P.job[j].code[k];
        op_c0[k] = s_c0[j][k];
        op_onep[k] = s_onep[j][k];
      }
      gbins = P.job[j].bins;
    }
    const unsigned src = my_stage0 + slot * stage_bytes + (unsigned)lane * 16u;
    const bool dyn = (cur & STREAM_DYN) != 0;
#pragma unroll 1
    for (int r = 0; r < R; ++r) {
      double f[NS][4];
#pragma unroll
      for (int k = 0; k < NS; ++k) {
        const float4 v = lds128(src + k * col_bytes + r * 512u);
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
      if (SMEM && !dyn) run_push4<true>(acc, bq, s_hist, geo.total_nbins);
      else run_push4<false>(acc, bq, gbins, geo.total_nbins);
    }
    __syncwarp();                       // every lane has read the stage before it is refilled
    if (lane == 0) issue_next(slot);
  }
  leave_static();                       // a warp that never saw a tile
  if (xd && lane == 0) xd[8] = gtimer();   // (last writer wins: a late warp's end of stream)

  // ---- merge the shared-memory histogram: warp w takes slice w once all warps left the static phase
  if (SMEM) {
    if (lane == 0) {
      unsigned spins = 0;
      while (*reinterpret_cast<volatile unsigned*>(&s_done) < (unsigned)FAST_WARPS)
        if (++spins > 64u) __nanosleep(64);
      __threadfence_block();
    }
    __syncwarp();
    unsigned* const job_bins = P.job[home].bins;
    const int nb = geo.total_nbins;
    const int lo = (int)((long long)nb * warp / FAST_WARPS), hi = (int)((long long)nb * (warp + 1) / FAST_WARPS);
    for (int b = lo + lane; b < hi; b += 32) {
      const unsigned cnt = s_hist[b];
      if (cnt) atomicAdd(job_bins + b, cnt);
    }
  }
  // ---- the last warp of the grid re-arms the tile queue for the next launch ---------------------
  if (lane == 0) {
    const unsigned total_warps = gridDim.x * FAST_WARPS;
    __threadfence();
    if (atomicAdd(P.queue + 1, 1u) == total_warps - 1u) {
      P.queue[0] = 0u;
      P.queue[1] = 0u;
      __threadfence();
    }
  }
  if (xd && lane == 0) xd[9] = gtimer();    // CTA exit (last writer wins)
}

}  // namespace sxmc
