// p2p_kernels.cu -- the histogram all-reduce of a sample-sharded step over NVLink peer memory
// (SURVEY 8(e)-2), as ONE kernel per step in place of ncclAllReduce.
//
// Every rank bins its slice of the samples into a partial histogram that lives in a window
// other GPUs can read (CUDA IPC; NVLink 5 / NVSwitch gives every peer full bandwidth).  The
// kernel, stream-ordered after the rank's binning kernel,
//   1. publishes "my partial of step s is complete" to every peer (release store, system scope),
//   2. waits until every peer has published the same (acquire loads, bounded spin),
//   3. sums the N partials with 128-bit peer loads into the rank's private, complete histogram.
// One synchronisation, (N-1) x histogram bytes of NVLink reads per rank (D: 3 MB per peer), no
// intermediate copies.  That is the form for two ranks; for more, reading every peer's whole
// partial costs (N-1) x 3 MB per rank (8 ranks: 68 us per step measured), so the kernel
// switches to reduce-scatter + all-gather inside the same launch: each rank sums 1/N of the
// words and stores that slice into every rank's sum buffer (which therefore lives in the
// window too), then a second flag round says "slice in place".  Integer sums, so every rank
// holds bit-identical histograms in either form.
//
// Buffer reuse is safe without further handshakes: a peer overwrites the partial of parity b
// only two steps later, and it cannot pass its own step s+1 kernel before this rank has
// published step s+1 -- which this rank does only after it finished reading step s.
#include "common.cuh"
#include "kernels.h"
#include "grid_sync.cuh"
#include "p2p_device.cuh"

namespace sxmc {

__global__ void __launch_bounds__(256)
hist_allreduce_p2p_kernel(const __grid_constant__ P2PParams W) {
  __shared__ int s_ok, s_last;
  const unsigned seq = *W.seq + 1u;              // the step being exchanged (same on every block)
  unsigned long long* const clk = reinterpret_cast<unsigned long long*>(W.seq + 4);   // [0] scratch, [1..3] sums (ns)
  unsigned long long t0 = 0;
  // the binning kernel finished before this launch started (stream order), so its partial is
  // complete in this rank's memory: tell every peer, one thread per peer so the system-scope
  // release stores travel in parallel
  if (blockIdx.x == 0 && threadIdx.x < W.nranks) {
    if (threadIdx.x == 0) t0 = globaltimer_ns();
    __threadfence_system();
    st_release_sys_u32(W.flags[threadIdx.x] + W.rank, seq);
  }
  if (threadIdx.x == 0) {
    s_ok = wait_peers(W, W.flags[W.rank], seq);
    if (blockIdx.x == 0) {
      const unsigned long long t1 = globaltimer_ns();
      clk[0] = t1;
      clk[1] += t1 - t0;
    }
  }
  __syncthreads();
  const unsigned long long gtid = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  const unsigned long long gstride = (unsigned long long)gridDim.x * blockDim.x;
  const size_t poff = (size_t)W.parity * W.pitch;
  if (s_ok && !W.twoshot) {
    const unsigned long long nvec = W.words / 4;
    for (unsigned long long v = gtid; v < nvec; v += gstride) {
      uint4 acc = make_uint4(0u, 0u, 0u, 0u);
      for (int p = 0; p < W.nranks; ++p) {
        const uint4 x = __ldcv(reinterpret_cast<const uint4*>(W.part[p] + poff) + v);
        acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
      }
      reinterpret_cast<uint4*>(W.sum)[v] = acc;
    }
    for (unsigned long long w = nvec * 4 + gtid; w < W.words; w += gstride) {
      unsigned acc = 0;
      for (int p = 0; p < W.nranks; ++p) acc += __ldcv(W.part[p] + poff + w);
      W.sum[w] = acc;
    }
  } else if (s_ok) {
    // reduce-scatter + all-gather: this rank sums its 1/N slice of the (128-byte padded) window
    // and stores the result into every rank's sum buffer: 2 (N-1)/N x histogram bytes of NVLink
    // traffic per rank instead of (N-1) x.
    const unsigned long long nvec = W.pitch / 4;
    const unsigned long long v0 = nvec * (unsigned)W.rank / (unsigned)W.nranks;
    const unsigned long long v1 = nvec * (unsigned)(W.rank + 1) / (unsigned)W.nranks;
    for (unsigned long long v = v0 + gtid; v < v1; v += gstride) {
      uint4 acc = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
      for (int p = 0; p < P2P_MAX_RANKS; ++p)
        if (p < W.nranks) {
          const uint4 x = __ldcv(reinterpret_cast<const uint4*>(W.part[p] + poff) + v);
          acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
        }
#pragma unroll
      for (int p = 0; p < P2P_MAX_RANKS; ++p)
        if (p < W.nranks)
          reinterpret_cast<uint4*>(const_cast<unsigned*>(W.part[p]) + W.sum_offset)[v] = acc;
    }
  }
  // the last block to finish closes the step: (two-shot) tell every peer this rank's slice is
  // in place and wait for theirs, then advance the sequence number for the next launch
  __syncthreads();
  if (threadIdx.x == 0) {
    // barrier, then ONE cumulative fence per block (the grid-sync idiom): the block's peer
    // stores are ordered before its ticket, hence before the flags the last block publishes.
    // (A system fence in every storing thread cost ~10 us per exchange.)
    if (W.twoshot) __threadfence_system(); else __threadfence();
    s_last = atomicAdd(W.ticket, 1u) == gridDim.x - 1;
    if (s_last) __threadfence();
  }
  __syncthreads();
  if (!s_last) return;
  unsigned long long t2 = 0;
  if (threadIdx.x == 0) {
    t2 = globaltimer_ns();
    clk[2] += t2 - *(volatile unsigned long long*)clk;
  }
  if (W.twoshot && s_ok) {
    if (threadIdx.x < W.nranks) {
      __threadfence_system();
      st_release_sys_u32(W.flags[threadIdx.x] + P2P_MAX_RANKS + W.rank, seq);
    }
    if (threadIdx.x == 0) wait_peers(W, W.flags[W.rank] + P2P_MAX_RANKS, seq);
  }
  if (threadIdx.x == 0) {
    clk[3] += globaltimer_ns() - t2;
    *W.ticket = 0u;
    *W.seq = seq;
  }
}

int launch_hist_allreduce_p2p(const P2PParams& W, cudaStream_t stream) {
  unsigned long long blocks = (W.words / 4 + 255) / 256;
  if (blocks > 148 * 2) blocks = 148 * 2;      // every block spins on the flags: keep them co-resident
  if (blocks < 1) blocks = 1;
  hist_allreduce_p2p_kernel<<<(unsigned)blocks, 256, 0, stream>>>(W);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

}  // namespace sxmc
