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
// intermediate copies: at these sizes the exchange is latency-bound, and one flag round trip
// beats the two or more a ring or tree needs.  Integer sums in a fixed rank order, so every
// rank holds bit-identical histograms.
//
// Buffer reuse is safe without further handshakes: a peer overwrites the partial of parity b
// only two steps later, and it cannot pass its own step s+1 kernel before this rank has
// published step s+1 -- which this rank does only after it finished reading step s.
#include "common.cuh"
#include "kernels.h"
#include "grid_sync.cuh"

namespace sxmc {

__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(256)
hist_allreduce_p2p_kernel(const __grid_constant__ P2PParams W) {
  __shared__ int s_ok;
  const unsigned seq = *W.seq + 1u;              // the step being exchanged (same on every block)
  if (threadIdx.x == 0) {
    int ok = 1;
    if (blockIdx.x == 0) {
      __threadfence_system();                    // the binning kernel's writes are ordered before the flags
      for (int p = 0; p < W.nranks; ++p) st_release_sys_u32(W.flags[p] + W.rank, seq);
    }
    const long long t0 = clock64();
    for (int p = 0; p < W.nranks && ok; ++p) {
      unsigned n = 0;
      while ((int)(ld_acquire_sys_u32(W.flags[W.rank] + p) - seq) < 0) {
        if ((++n & 63u) == 0) {
          if (*(volatile unsigned*)W.abort_flag) { ok = 0; break; }
          if (clock64() - t0 > PK_TIMEOUT_CYCLES) { atomicExch(W.abort_flag, 1u); ok = 0; break; }
        }
      }
    }
    s_ok = ok;
  }
  __syncthreads();
  if (s_ok) {
    const unsigned long long nvec = W.words / 4;
    const unsigned long long gtid = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
    const unsigned long long gstride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long v = gtid; v < nvec; v += gstride) {
      uint4 acc = make_uint4(0u, 0u, 0u, 0u);
      for (int p = 0; p < W.nranks; ++p) {
        const uint4 x = __ldcv(reinterpret_cast<const uint4*>(W.part[p] + (size_t)W.parity * W.pitch) + v);
        acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
      }
      reinterpret_cast<uint4*>(W.sum)[v] = acc;
    }
    for (unsigned long long w = nvec * 4 + gtid; w < W.words; w += gstride) {
      unsigned acc = 0;
      for (int p = 0; p < W.nranks; ++p) acc += __ldcv(W.part[p] + (size_t)W.parity * W.pitch + w);
      W.sum[w] = acc;
    }
  }
  // the last block to finish advances the sequence number for the next step's launch
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(W.ticket, 1u) == gridDim.x - 1) {
      *W.ticket = 0u;
      *W.seq = seq;
    }
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
