// nll_sharded_kernels.cu -- the sample-sharded step's exchange FUSED into a rank-sharded lookup
// (no counterpart in the reference, whose fit loop is one process on one device).
//
// The plain sharded step runs three kernels per rank: binning of the rank's sample slice, the
// histogram all-reduce (p2p_kernels.cu: every rank ends up with all N x 3 MB summed), and the
// lookup + NLL + Metropolis kernel, replicated on every rank.  Here the last two are ONE launch
// and the lookup itself is dealt over the ranks:
//   * the grouped lookup's logical grid -- G blocks of 128 bins, exactly the grid the unsharded
//     chain launches (launch_nll) -- is cut into N contiguous ranges of blocks, one per rank;
//   * a physical block (512 threads) gathers the N partial counts of its 128 bins for every
//     signal straight from the peers' windows over NVLink -- 128-bit loads, all of a thread's loads
//     (and the in-domain counts) issued before the first is consumed, so the block pays ONE NVLink
//     round trip -- and adds them up in shared memory (integer atomics: order-free);
//   * four threads per bin then evaluate eval_pdf's division and nll_event_chunks' product for the
//     bin's signals, and one of them adds the products in signal order: the arithmetic, the order
//     of additions and the block reduction of lookup_nll_kernel<128>;
//   * the block's sum goes to slot [logical block] of EVERY rank's partial array (N tagged stores
//     of 16 bytes instead of an all-gather of the histograms);
//   * the rank's last block collects the G sums as they arrive, totals them in the order the
//     unsharded kernel's last block uses and finishes the step (nll_total, jump_decider,
//     pick_new_vector: finish_step2) -- on every rank alike.
// Every rank therefore takes the decision the unsharded chain takes, bit for bit, for any N
// (tests/test_sharding.py), and nobody ever holds the summed histograms: a step moves
// (N-1)/N x 3 MB over NVLink into each rank once, and the replicated lookup is gone.
//
// Synchronisation per step: ONE flag round -- [p][r] = s in the window's flag row the all-reduce
// kernel uses too: rank r's partial histogram of step s is complete, published after its binning
// kernel -- and the partial sums themselves, each tagged with s.  A rank that holds all G sums of
// step s knows every rank has finished reading the partial histograms of step s.  Buffer reuse
// needs nothing more.  A rank clears partial buffer b^1 (for step s+1) during the lookup of step s:
// its peers read b^1 during step s-1 and this rank had their sums of s-1 before it left step s-1.  A
// peer's sums of step s+1 cannot overwrite those of step s unread: the peer needs this rank's
// [p][r] = s+1 first, which this rank's NEXT launch publishes.
#include <algorithm>

#include "common.cuh"
#include "kernels.h"
#include "nll_device.cuh"
#include "p2p_device.cuh"

namespace sxmc {

constexpr int LB_BINS = 128;             // bins per logical block == threads of lookup_nll_kernel<128>
constexpr int WARPS_PER_LB = LB_BINS / 32;
// A physical block stands for one logical block, four threads per bin.  (Measured and dropped:
// physical blocks of 32 bins -- one warp of the logical block each, four times as many blocks and
// SMs reading over NVLink, warp sums exchanged: the gather took 15 us instead of 11 at 8 GPUs;
// profiles/r02_bench_D_sharded_8gpu_fused_v2_32bin_blocks.json.)
constexpr int PB_BINS = LB_BINS;
constexpr int PB_THREADS = 4 * PB_BINS;
constexpr int SH_GATHER = 16;            // 128-bit peer loads in flight per thread
constexpr int SH_MAX_SIGNALS = 128;

// one tagged partial sum: {low half, step, high half, step} -- see the store below
__device__ __forceinline__ void st_tagged(unsigned* slot, double v, unsigned seq) {
  const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(slot), "r"((unsigned)bits), "r"(seq) : "memory");
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(slot + 2), "r"((unsigned)(bits >> 32)), "r"(seq) : "memory");
}
// spin until both halves carry `seq`; false on abort / timeout
__device__ __forceinline__ bool ld_tagged(const unsigned* slot, unsigned seq, const P2PParams& W,
                                          unsigned long long t_start, double& v) {
  unsigned lo, hi, f0, f1, n = 0;
  for (;;) {
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(lo), "=r"(f0) : "l"(slot) : "memory");
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(hi), "=r"(f1) : "l"(slot + 2) : "memory");
    if (f0 == seq && f1 == seq) break;
    if ((++n & 63u) == 0) {
      if (*(volatile unsigned*)W.abort_flag) return false;
      if (globaltimer_ns() - t_start > W.peer_timeout_ns) { atomicExch(W.abort_flag, 1u); return false; }
    }
  }
  v = __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
  return true;
}

// block_sum<128> (nll_device.cuh) executed by the first four warps of a block: the same shuffle
// tree, the same order over the warp sums.  Result valid on thread 0.
__device__ __forceinline__ double block_sum128_first_warps(double v, double* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp < WARPS_PER_LB) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  }
  __syncthreads();
  if (warp < WARPS_PER_LB && lane == 0) s_warp[warp] = v;
  __syncthreads();
  double t = 0;
  if (threadIdx.x == 0)
    for (int w = 0; w < WARPS_PER_LB; ++w) t += s_warp[w];
  return t;
}

__global__ void __launch_bounds__(PB_THREADS)
lookup_nll_sharded_kernel(const __grid_constant__ ShardNllParams Q) {
  extern __shared__ double s_dyn[];                       // [ns][128] products, then [ns][128] counts
  __shared__ double s_c[SH_MAX_SIGNALS];
  __shared__ double s_bn[SH_MAX_SIGNALS];
  __shared__ unsigned s_norm[SH_MAX_SIGNALS];
  __shared__ unsigned s_m[PB_BINS];
  __shared__ double s_warp[WARPS_PER_LB];
  __shared__ double s_term[SH_MAX_SIGNALS + SXMC_MAX_PARAMETERS];
  __shared__ int s_ok, s_last;
  __shared__ FinState F;
  __shared__ double s_z[SXMC_MAX_PARAMETERS];
  __shared__ double s_u, s_bsum;

  const P2PParams& W = Q.W;
  const int tid = threadIdx.x;
  const int t = tid & (PB_BINS - 1), kq = tid / PB_BINS;  // evaluation: bin of the block, quarter of the signals
  const int ns = Q.ns, np = Q.np, nb = Q.nbins, nr = W.nranks;
  double* const s_prod = s_dyn;                                            // [ns][128]
  unsigned* const s_cnt = reinterpret_cast<unsigned*>(s_dyn + (size_t)ns * PB_BINS);   // [ns][128]
  const int lb = Q.lb0 + blockIdx.x;                      // logical block of the unsharded grid
  const unsigned seq = *W.seq + 1u;
  unsigned long long* const clk = reinterpret_cast<unsigned long long*>(W.seq + 4);
  unsigned long long t0 = 0, tph = 0;      // phase clocks of block 0 (clk[4..7], sxmc_chain_p2p_clocks)
#define SH_PHASE(k)                                                          \
  do {                                                                       \
    if (blockIdx.x == 0 && tid == 0) {                                       \
      const unsigned long long t_ = globaltimer_ns();                        \
      clk[k] += t_ - tph;                                                    \
      tph = t_;                                                              \
    }                                                                        \
  } while (0)

  // ---- prologue independent of this step's histograms (overlaps the binning kernel's tail) ----
  prefetch_state(F, Q.st, np, tid, PB_THREADS);
  NllSignal my_sig;
  if (tid < ns) my_sig = Q.sigs[tid];
  const long long gstride = (long long)Q.logical_grid * LB_BINS;
  long long b0 = (long long)lb * LB_BINS;
  if (tid < PB_BINS) s_m[tid] = (b0 + tid < nb) ? Q.mult[b0 + tid] : 0u;   // events per bin: static data
  for (unsigned long long w = blockIdx.x * (unsigned long long)PB_THREADS + tid; w < Q.zero_words;
       w += (unsigned long long)gridDim.x * PB_THREADS)
    Q.zero_ptr[w] = 0u;                                   // the buffer step s+1 bins into (header)
  for (int k = tid; k < ns * PB_BINS; k += PB_THREADS) s_cnt[k] = 0u;
  if (tid < ns) s_norm[tid] = 0u;
  pdl_wait();
  pdl_trigger();

  // ---- my partial is complete (the binning kernel has finished): say so, wait for everyone's ----
  if (blockIdx.x == 0 && tid < nr) {
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
      tph = t1;
    }
  }
  __syncthreads();
  const bool ok = s_ok != 0;
  const size_t poff = (size_t)W.parity * W.pitch;

  const bool pre_rng = pre_rng_applies(F, Q.mode);
  if (pre_rng && blockIdx.x == 0 && tid >= PB_THREADS - 32) pre_rng_draw(F, np, tid & 31, Q.st->pre_z, &Q.st->pre_u);

  double sum = 0;
  bool first = true;
  for (; b0 < nb; b0 += gstride) {
    if (!first) {                           // (more than 148 x 8 x 128 bins: several bins per thread)
      __syncthreads();
      for (int k = tid; k < ns * PB_BINS; k += PB_THREADS) s_cnt[k] = 0u;
      if (tid < PB_BINS) s_m[tid] = (b0 + tid < nb) ? Q.mult[b0 + tid] : 0u;
      __syncthreads();
    }
    // ---- gather: the peers' counts of the block's 32 bins, every signal, ONE NVLink round trip:
    // the in-domain counts (first pass) and up to SH_GATHER 128-bit loads per thread are issued
    // before any of them is consumed.  Item i = (quad of bins, peer, signal), peers fastest after
    // the quad, so neighbouring groups of lanes read from different GPUs.
    constexpr int NRM = SH_MAX_SIGNALS * P2P_MAX_RANKS / PB_THREADS;     // (signal, peer) pairs per thread
    unsigned nrm[NRM];
    const bool want_norm = first && ok;
    if (want_norm)
#pragma unroll
      for (int u = 0; u < NRM; ++u) {
        const int k = tid + u * PB_THREADS;
        nrm[u] = (k < ns * nr) ? __ldcv(W.part[k / ns] + poff + (size_t)ns * nb + k % ns) : 0u;
      }
    constexpr int QUADS = PB_BINS / 4;
    const int items = QUADS * nr * ns;
    if (ok && (nb & 3) == 0) {
      for (int i0 = tid; i0 < items; i0 += PB_THREADS * SH_GATHER) {
        uint4 x[SH_GATHER];
#pragma unroll
        for (int u = 0; u < SH_GATHER; ++u) {
          const int i = i0 + u * PB_THREADS;
          x[u] = make_uint4(0u, 0u, 0u, 0u);
          if (i < items) {
            const int q = i % QUADS, pj = i / QUADS, pp = pj % nr, j = pj / nr;
            const long long b = b0 + 4 * q;
            if (b < nb) x[u] = __ldcv(reinterpret_cast<const uint4*>(W.part[pp] + poff + (size_t)j * nb + b));
          }
        }
#pragma unroll
        for (int u = 0; u < SH_GATHER; ++u) {
          const int i = i0 + u * PB_THREADS;
          if (i < items) {
            const int q = i % QUADS, j = (i / QUADS) / nr;
            unsigned* dst = s_cnt + j * PB_BINS + 4 * q;
            if (x[u].x) atomicAdd(dst + 0, x[u].x);
            if (x[u].y) atomicAdd(dst + 1, x[u].y);
            if (x[u].z) atomicAdd(dst + 2, x[u].z);
            if (x[u].w) atomicAdd(dst + 3, x[u].w);
          }
        }
      }
    } else if (ok) {                        // bins per PDF not a multiple of four: 32-bit loads
      for (int i = tid; i < PB_BINS * nr * ns; i += PB_THREADS) {
        const int tb = i % PB_BINS, pj = i / PB_BINS, pp = pj % nr, j = pj / nr;
        if (b0 + tb < nb && s_m[tb]) {
          const unsigned h = __ldcv(W.part[pp] + poff + (size_t)j * nb + b0 + tb);
          if (h) atomicAdd(s_cnt + j * PB_BINS + tb, h);
        }
      }
    }
    if (want_norm)
#pragma unroll
      for (int u = 0; u < NRM; ++u) {
        const int k = tid + u * PB_THREADS;
        if (k < ns * nr) atomicAdd(&s_norm[k % ns], nrm[u]);
      }
    __syncthreads();
    if (first) {
      for (int j = tid; j < ns; j += PB_THREADS) {
        const NllSignal sg = (j == tid) ? my_sig : Q.sigs[j];
        const unsigned my_norm = s_norm[j];
        const float eff = (float)(1.0 * my_norm / sg.n_mc);                     // nll_kernels.cpp:105
        const double rate = __dmul_rn(F.proposed[sg.source_id], sg.nexpected);
        s_c[j] = __dmul_rn(rate, (double)eff);
        s_bn[j] = __dmul_rn((double)my_norm, sg.bin_volume);                    // pdfz.cpp:417
        s_term[j] = __dmul_rn(rate, (double)my_norm) / (double)sg.n_mc;         // nll_kernels.cpp:167-171
      }
      for (int i = tid; i < np; i += PB_THREADS) {                              // nll_kernels.cpp:176-185
        const double sg = F.sigmas[i];
        double x2 = 0;
        if (sg > 0) {
          const double x = __dadd_rn(F.proposed[i], -F.means[i]) / sg;
          x2 = __dmul_rn(__dmul_rn(0.5, x), x);
        }
        s_term[ns + i] = x2;
      }
      __syncthreads();
    }
    SH_PHASE(first ? 4 : 5);
    // ---- evaluate: the division of eval_pdf and the product of nll_event_chunks for every (signal,
    // bin) by four threads per bin, then thread (t, 0) adds its bin's products in signal order --
    // each product is rounded on its own, so the sum is the unsharded kernel's, bit for bit
    const unsigned m = s_m[t];
    if (m)
      for (int j = kq; j < ns; j += PB_THREADS / PB_BINS) {
        float v = pdf_value(s_cnt[j * PB_BINS + t], s_bn[j]);                    // pdfz.cpp:429-432
        if (isnan(v)) v = 0.0f;                                                 // nll_kernels.cpp:107
        s_prod[j * PB_BINS + t] = __dmul_rn(s_c[j], (double)v);
      }
    __syncthreads();
    if (kq == 0 && m) {
      double sv = 0;
      for (int j = 0; j < ns; ++j) sv = __dadd_rn(sv, s_prod[j * PB_BINS + t]);
      if (sv > 0) sum += (double)m * log(sv);
    }
    first = false;
  }
  SH_PHASE(5);
  const double bsum = block_sum128_first_warps(sum, s_warp);      // threads (t, 0) are the first four warps
  if (tid == 0) s_bsum = bsum;
  __syncthreads();
  SH_PHASE(6);
  // The block's sum into slot lb of every rank's array, as two 8-byte words that each carry half of
  // the double AND the step number (the "LL" idiom of the collective libraries: an aligned 8-byte
  // store arrives whole, so the reader needs no fence and no separate flag -- a word whose tag is
  // the step's holds that step's data).
  if (tid < nr && ok) st_tagged(const_cast<unsigned*>(W.part[tid]) + Q.partials_offset + 4ull * lb, s_bsum, seq);
  if (tid == 0) {                     // elects the block that finishes the step; the fences order block 0's
    __threadfence();                  // pre-drawn random numbers (pre_z, pre_u) before the finisher's loads
    s_last = atomicAdd(W.ticket, 1u) == gridDim.x - 1;
    if (s_last) __threadfence();
  }
  SH_PHASE(7);
  __syncthreads();
  if (!s_last) return;

  // ---- the rank's last block: collect every logical block's sum (own and peers'), total them in
  // the unsharded kernel's order, finish the step -- on every rank alike.  Holding all sums of step s also means every rank is done reading this
  // step's partial histograms (a block stores its sum after its gather): the header's argument.
  unsigned long long t2 = 0;
  if (tid == 0) {
    t2 = globaltimer_ns();
    clk[2] += t2 - *(volatile unsigned long long*)clk;
    s_ok = ok ? 1 : 0;
  }
  __syncthreads();
  // every thread fetches a batch of slots at once (most have arrived), spins only on the late ones
  const unsigned* slots = W.part[W.rank] + Q.partials_offset;
  double* const s_val = s_dyn;                            // [G] (the launch sized the dynamic array for it)
  const int nslots = Q.logical_grid;
  if (ok) {
    const unsigned long long tw = globaltimer_ns();
    constexpr int BATCH = 8;
    bool good = true;
    for (int k0 = tid; k0 < nslots && good; k0 += PB_THREADS * BATCH) {
      uint4 x[BATCH];
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        const int k = k0 + u * PB_THREADS;
        x[u] = make_uint4(0u, 0u, 0u, 0u);
        if (k < nslots)
          asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(x[u].x), "=r"(x[u].y), "=r"(x[u].z), "=r"(x[u].w) : "l"(slots + 4ull * k) : "memory");
      }
#pragma unroll
      for (int u = 0; u < BATCH; ++u) {
        const int k = k0 + u * PB_THREADS;
        if (k < nslots && good) {
          double v;
          if (x[u].y == seq && x[u].w == seq) v = __longlong_as_double((long long)(((unsigned long long)x[u].z << 32) | x[u].x));
          else good = ld_tagged(slots + 4ull * k, seq, W, tw, v);
          s_val[k] = v;
        }
      }
    }
    if (!good) s_ok = 0;
  }
  __syncthreads();
  double acc = 0;
  if (s_ok && tid < LB_BINS)
    for (int k = tid; k < Q.logical_grid; k += LB_BINS) acc += s_val[k];
  __syncthreads();
  if (tid == 0) {
    clk[3] += globaltimer_ns() - t2;
    *W.ticket = 0u;
    *W.seq = seq;
  }
  if (!s_ok) return;                  // timed out: the abort flag is up, the host reports it
  const double total = block_sum128_first_warps(acc, s_warp);
  if (tid == 0) s_warp[0] = total;
  if (pre_rng) {
    for (int k = tid; k < np; k += PB_THREADS) s_z[k] = __ldcg(Q.st->pre_z + k);
    if (tid == 0) s_u = __ldcg(&Q.st->pre_u);
  }
  __syncthreads();
  if (tid < 64)
    finish_step2(F, Q.st, np, ns, Q.nsources, s_warp[0], Q.mode, s_term, tid & 31, tid >> 5, pre_rng ? s_z : nullptr,
                 pre_rng ? &s_u : nullptr);
#undef SH_PHASE
}

// The logical grid of the unsharded grouped lookup over many bins (launch_nll).
int nll_sharded_logical_grid(int nbins) {
  int grid = (nbins + LB_BINS - 1) / LB_BINS;
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
  // this rank's contiguous range of logical blocks
  Q.lb0 = (int)((long long)Q.logical_grid * Q.W.rank / Q.W.nranks);
  const int lb1 = (int)((long long)Q.logical_grid * (Q.W.rank + 1) / Q.W.nranks);
  if (lb1 <= Q.lb0) { set_error("launch_nll_sharded: fewer logical blocks than ranks"); return SXMC_ESTATE; }
  size_t smem = (sizeof(double) + sizeof(unsigned)) * (size_t)Q.ns * PB_BINS;
  smem = std::max(smem, sizeof(double) * (size_t)Q.logical_grid);                    // the finisher's slot values
  // (static + dynamic shared memory pass 48 KB from 17 signals on: always opt in)
  SXMC_CUDA(cudaFuncSetAttribute((const void*)lookup_nll_sharded_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SXMC_CUDA(launch_kernel_pdl(lookup_nll_sharded_kernel, dim3((unsigned)(lb1 - Q.lb0)), dim3(PB_THREADS), smem, stream, pdl, Q));
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

}  // namespace sxmc
