// nll_sharded_kernels.cu -- the sample-sharded step's exchange FUSED into a rank-sharded lookup
// (no counterpart in the reference, whose fit loop is one process on one device).
//
// The plain sharded step runs three kernels per rank: binning of the rank's sample slice, the
// histogram all-reduce (p2p_kernels.cu: every rank ends up with all N x 3 MB summed), and the
// lookup + NLL + Metropolis kernel, replicated on every rank.  Here the last two are ONE launch
// and the lookup itself is dealt over the ranks:
//   * the grouped lookup's logical grid -- G blocks of 128 bins, exactly the grid the unsharded
//     chain launches (launch_nll) -- is cut into N contiguous ranges of blocks, one per rank;
//   * a physical block has 128 x N threads: thread (t, p) loads peer p's partial counts of the
//     block's bin t for every signal straight from p's window over NVLink (up to 16 loads in
//     flight per thread, so the block pays about one NVLink round trip), the N partials meet in
//     shared memory (integer atomics: order-free), and threads (t, 0) evaluate eval_pdf +
//     nll_event_chunks for their bin with the arithmetic, the order of additions and the block
//     reduction of lookup_nll_kernel<128>;
//   * the block's partial sum goes to slot [logical block] of EVERY rank's partial array (N
//     stores of 8 bytes instead of an all-gather of the histograms);
//   * the rank's last block tells every peer "my sums are in place", waits for theirs, totals
//     the G partials in the order the unsharded kernel's last block uses and finishes the step
//     (nll_total, jump_decider, pick_new_vector: finish_step2) -- on every rank alike.
// Every rank therefore takes the decision the unsharded chain takes, bit for bit, for any N
// (tests/test_sharding.py), and nobody ever holds the summed histograms: a step moves
// (N-1)/N x 3 MB over NVLink into each rank once, and the replicated lookup is gone.
//
// Two flag rounds per step, the same rows of the window the all-reduce kernel uses:
//   [p][r]      rank r's partial histogram of step s is complete (published after its binning kernel)
//   [p][8 + r]  rank r has read what it needs of step s and stored its partial sums everywhere
// Buffer reuse needs nothing more.  A rank clears partial buffer b^1 (for step s+1) during the
// lookup of step s: its peers read b^1 during step s-1 and said so ([8 + r] = s-1) before this rank
// left step s-1.  A peer's partial sums of step s+1 cannot arrive before this rank has read those
// of step s: the peer needs this rank's [p][r] = s+1 first, which this rank's NEXT launch publishes.
#include "common.cuh"
#include "kernels.h"
#include "nll_device.cuh"
#include "p2p_device.cuh"

namespace sxmc {

constexpr int SH_BINS = 128;             // bins per logical block == threads of lookup_nll_kernel<128>
constexpr int SH_MAX_SIGNALS = 128;

// block_sum<128> (nll_device.cuh) executed by the first four warps of a larger block: the same
// shuffle tree, the same order over the warp sums.  Result valid on thread 0.
__device__ __forceinline__ double block_sum128_first_warps(double v, double* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp < SH_BINS / 32) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  }
  __syncthreads();
  if (warp < SH_BINS / 32 && lane == 0) s_warp[warp] = v;
  __syncthreads();
  double t = 0;
  if (threadIdx.x == 0)
    for (int w = 0; w < SH_BINS / 32; ++w) t += s_warp[w];
  return t;
}

__global__ void __launch_bounds__(1024)
lookup_nll_sharded_kernel(const __grid_constant__ ShardNllParams Q) {
  extern __shared__ unsigned s_cnt[];                     // [ns][128] summed counts of the block's bins
  __shared__ double s_c[SH_MAX_SIGNALS];
  __shared__ double s_bn[SH_MAX_SIGNALS];
  __shared__ unsigned s_norm[SH_MAX_SIGNALS];
  __shared__ double s_warp[SH_BINS / 32];
  __shared__ double s_term[SH_MAX_SIGNALS + SXMC_MAX_PARAMETERS];
  __shared__ int s_ok, s_last;
  __shared__ FinState F;
  __shared__ double s_z[SXMC_MAX_PARAMETERS];
  __shared__ double s_u, s_bsum;

  const P2PParams& W = Q.W;
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int t = tid & (SH_BINS - 1), p = tid / SH_BINS;   // bin of the block, peer
  const int ns = Q.ns, np = Q.np, nb = Q.nbins;
  const int lb = Q.lb0 + blockIdx.x;                      // logical block of the unsharded grid
  const unsigned seq = *W.seq + 1u;
  unsigned long long* const clk = reinterpret_cast<unsigned long long*>(W.seq + 4);
  unsigned long long t0 = 0;

  // ---- prologue independent of this step's histograms (overlaps the binning kernel's tail) ----
  prefetch_state(F, Q.st, np, tid, nthreads);
  NllSignal my_sig;
  if (tid < ns) my_sig = Q.sigs[tid];
  for (unsigned long long w = blockIdx.x * (unsigned long long)nthreads + tid; w < Q.zero_words;
       w += (unsigned long long)gridDim.x * nthreads)
    Q.zero_ptr[w] = 0u;                                   // the buffer step s+1 bins into (header)
  for (int k = tid; k < ns * SH_BINS; k += nthreads) s_cnt[k] = 0u;
  if (tid < ns) s_norm[tid] = 0u;
  pdl_wait();
  pdl_trigger();

  // ---- round 1: my partial is complete (the binning kernel has finished); wait for everyone's ----
  if (blockIdx.x == 0 && tid < W.nranks) {
    if (tid == 0) t0 = globaltimer_ns();
    __threadfence_system();
    st_release_sys_u32(W.flags[tid] + W.rank, seq);
  }
  if (tid == 0) {
    s_ok = wait_peers(W, W.flags[W.rank], seq);
    if (blockIdx.x == 0) {
      const unsigned long long t1 = globaltimer_ns();
      clk[0] = t1;
      clk[1] += t1 - t0;
    }
  }
  __syncthreads();
  const bool ok = s_ok != 0;
  const size_t poff = (size_t)W.parity * W.pitch;
  const unsigned* part = W.part[p] + poff;

  const bool pre_rng = pre_rng_applies(F, Q.mode);
  if (pre_rng && blockIdx.x == 0 && tid >= nthreads - 32) pre_rng_draw(F, np, tid & 31, Q.st->pre_z, &Q.st->pre_u);

  // in-domain counts: every (signal, peer) pair one load
  if (ok)
    for (int k = tid; k < ns * W.nranks; k += nthreads) {
      const int j = k % ns, q = k / ns;
      atomicAdd(&s_norm[j], __ldcv(W.part[q] + poff + (size_t)ns * nb + j));
    }
  __syncthreads();
  if (tid < ns) {
    const int j = tid;
    const unsigned my_norm = s_norm[j];
    const float eff = (float)(1.0 * my_norm / my_sig.n_mc);                 // nll_kernels.cpp:105
    const double rate = __dmul_rn(F.proposed[my_sig.source_id], my_sig.nexpected);
    s_c[j] = __dmul_rn(rate, (double)eff);
    s_bn[j] = __dmul_rn((double)my_norm, my_sig.bin_volume);                // pdfz.cpp:417
    s_term[j] = __dmul_rn(rate, (double)my_norm) / (double)my_sig.n_mc;     // nll_kernels.cpp:167-171
  }
  for (int i = tid; i < np; i += nthreads) {                                // nll_kernels.cpp:176-185
    const double sg = F.sigmas[i];
    double x2 = 0;
    if (sg > 0) {
      const double x = __dadd_rn(F.proposed[i], -F.means[i]) / sg;
      x2 = __dmul_rn(__dmul_rn(0.5, x), x);
    }
    s_term[ns + i] = x2;
  }

  // ---- the block's bins: bin b of logical block lb, then b + G x 128, ... as the unsharded grid ----
  double sum = 0;
  const long long gstride = (long long)Q.logical_grid * SH_BINS;
  bool first = true;
  for (long long b0 = (long long)lb * SH_BINS; b0 < nb; b0 += gstride) {
    const long long b = b0 + t;
    if (!first) {
      __syncthreads();
      for (int k = tid; k < ns * SH_BINS; k += nthreads) s_cnt[k] = 0u;
      __syncthreads();
    }
    first = false;
    const unsigned m = (b < nb) ? Q.mult[b] : 0u;
    if (ok && m) {
      constexpr int LW = 16;    // loads in flight per thread (64 registers at 1024 threads)
      for (int j0 = 0; j0 < ns; j0 += LW) {
        unsigned h[LW];
#pragma unroll
        for (int q = 0; q < LW; ++q) h[q] = (j0 + q < ns) ? __ldcv(part + (size_t)(j0 + q) * nb + b) : 0u;
#pragma unroll
        for (int q = 0; q < LW; ++q)
          if (j0 + q < ns && h[q]) atomicAdd(&s_cnt[(j0 + q) * SH_BINS + t], h[q]);
      }
    }
    __syncthreads();
    if (p == 0 && m) {
      double s = 0;
      for (int j = 0; j < ns; ++j) {
        float v = pdf_value(s_cnt[j * SH_BINS + t], s_bn[j]);                // pdfz.cpp:429-432
        if (isnan(v)) v = 0.0f;                                             // nll_kernels.cpp:107
        s = __dadd_rn(s, __dmul_rn(s_c[j], (double)v));
      }
      if (s > 0) sum += (double)m * log(s);
    }
  }
  const double bsum = block_sum128_first_warps(sum, s_warp);
  if (tid == 0) s_bsum = bsum;
  __syncthreads();
  // the block's partial sum into slot lb of every rank's array
  if (tid < W.nranks)
    reinterpret_cast<double*>(const_cast<unsigned*>(W.part[tid]) + Q.partials_offset)[lb] = s_bsum;
  __syncthreads();
  if (tid == 0) {
    __threadfence_system();           // cumulative: the block's peer stores before its ticket
    s_last = atomicAdd(W.ticket, 1u) == gridDim.x - 1;
    if (s_last) __threadfence();
  }
  __syncthreads();
  if (!s_last) return;

  // ---- the rank's last block: round 2, then the step's tail on every rank alike -------------
  unsigned long long t2 = 0;
  if (tid == 0) {
    t2 = globaltimer_ns();
    clk[2] += t2 - *(volatile unsigned long long*)clk;
  }
  if (tid < W.nranks) {
    __threadfence_system();
    st_release_sys_u32(W.flags[tid] + P2P_MAX_RANKS + W.rank, seq);
  }
  if (tid == 0) {
    const int ok2 = wait_peers(W, W.flags[W.rank] + P2P_MAX_RANKS, seq);
    s_ok = ok && ok2;
    clk[3] += globaltimer_ns() - t2;
    *W.ticket = 0u;
    *W.seq = seq;
  }
  __syncthreads();
  if (!s_ok) return;                  // timed out: the abort flag is up, the host reports it
  const double* partials = reinterpret_cast<const double*>(W.part[W.rank] + Q.partials_offset);
  double acc = 0;
  if (tid < SH_BINS)
    for (int k = tid; k < Q.logical_grid; k += SH_BINS) acc += __ldcv(partials + k);
  const double total = block_sum128_first_warps(acc, s_warp);
  if (tid == 0) s_warp[0] = total;
  if (pre_rng) {
    for (int k = tid; k < np; k += nthreads) s_z[k] = __ldcg(Q.st->pre_z + k);
    if (tid == 0) s_u = __ldcg(&Q.st->pre_u);
  }
  __syncthreads();
  if (tid < 64)
    finish_step2(F, Q.st, np, ns, Q.nsources, s_warp[0], Q.mode, s_term, tid & 31, tid >> 5, pre_rng ? s_z : nullptr,
                 pre_rng ? &s_u : nullptr);
}

// The logical grid of the unsharded grouped lookup over many bins (launch_nll).
int nll_sharded_logical_grid(int nbins) {
  int grid = (nbins + SH_BINS - 1) / SH_BINS;
  if (grid > 148 * 8) grid = 148 * 8;
  return grid;
}

int launch_nll_sharded(ShardNllParams Q, bool pdl, cudaStream_t stream) {
  if (Q.ns > SH_MAX_SIGNALS || Q.W.nranks > P2P_MAX_RANKS || Q.W.nranks < 2) {
    set_error("launch_nll_sharded: 2..8 ranks, at most 128 signals");
    return SXMC_EINVAL;
  }
  Q.logical_grid = nll_sharded_logical_grid(Q.nbins);
  if (Q.logical_grid > P2P_PARTIALS_MAX) { set_error("launch_nll_sharded: logical grid exceeds the window's partial array"); return SXMC_ESTATE; }
  Q.lb0 = (int)((long long)Q.logical_grid * Q.W.rank / Q.W.nranks);
  const int lb1 = (int)((long long)Q.logical_grid * (Q.W.rank + 1) / Q.W.nranks);
  if (lb1 <= Q.lb0) { set_error("launch_nll_sharded: fewer logical blocks than ranks"); return SXMC_ESTATE; }
  const size_t smem = sizeof(unsigned) * (size_t)Q.ns * SH_BINS;
  if (smem > 48 * 1024)
    SXMC_CUDA(cudaFuncSetAttribute((const void*)lookup_nll_sharded_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SXMC_CUDA(launch_kernel_pdl(lookup_nll_sharded_kernel, dim3((unsigned)(lb1 - Q.lb0)), dim3((unsigned)(SH_BINS * Q.W.nranks)), smem,
                              stream, pdl, Q));
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

}  // namespace sxmc
