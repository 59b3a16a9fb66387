// p2p_device.cuh -- flag primitives of the peer-memory exchanges (p2p_kernels.cu, nll_sharded_kernels.cu):
// system-scope release / acquire accesses to flag words other GPUs map through CUDA IPC, and the
// bounded wait for every peer of a window.
#pragma once
#include "common.cuh"
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

// Spin until every peer's flag in this rank's own flag row `base` has reached `seq`.
__device__ __forceinline__ int wait_peers(const P2PParams& W, const unsigned* base, unsigned seq) {
  // The peers are other host processes: their skew (a GIL stall, a lazy module load, a graph
  // instantiation) is bounded in wall-clock time, far above the grid barriers' cycle budget.
  const unsigned long long t0 = globaltimer_ns();
  for (int p = 0; p < W.nranks; ++p) {
    unsigned n = 0;
    while ((int)(ld_acquire_sys_u32(base + p) - seq) < 0) {
      if ((++n & 63u) == 0) {
        if (*(volatile unsigned*)W.abort_flag) return 0;
        if (globaltimer_ns() - t0 > W.peer_timeout_ns) { atomicExch(W.abort_flag, 1u); return 0; }
      }
    }
  }
  return 1;
}

}  // namespace sxmc
