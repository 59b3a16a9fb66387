// grid_sync.cuh -- grid-wide synchronisation of the persistent (cooperative) kernels: release/
// acquire counters in global memory, one spinning thread per CTA, every spin bounded by a
// clock64 timeout that raises an abort flag instead of hanging the device.
#pragma once
#include "common.cuh"

namespace sxmc {

constexpr long long PK_TIMEOUT_CYCLES = 4000000000ll;   // ~2 s at 1.9 GHz

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// thread 0 spins until *p >= target (wrap-safe); false on abort/timeout.  Ends with a CTA barrier.
__device__ __forceinline__ bool cta_wait_ge(const unsigned* p, unsigned target, unsigned* abort_flag,
                                            int* s_ok) {
  if (threadIdx.x == 0) {
    int ok = 1;
    const long long t0 = clock64();
    unsigned n = 0;
    while ((int)(ld_acquire_u32(p) - target) < 0) {
      if ((++n & 255u) == 0) {
        if (ld_acquire_u32(abort_flag)) { ok = 0; break; }
        if (clock64() - t0 > PK_TIMEOUT_CYCLES) { atomicExch(abort_flag, 1u); ok = 0; break; }
      }
    }
    *s_ok = ok;
  }
  __syncthreads();
  return *s_ok != 0;
}

// all threads' prior global writes become visible before the counter moves
__device__ __forceinline__ void cta_arrive(unsigned* counter) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
  }
}

}  // namespace sxmc
