// noalloc_kernels.cu -- sxmc_nll_eval_noalloc: MCMC::nll over a caller-held lut (the operation of
// sxmc_nll_eval / nll_lut_kernel, reference src/mcmc.cpp:390-415) for callers that must not
// allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// The kernel is a multi-block variant of nll_lut_kernel: every block sums log(sum_j N_j P_j) over
// its events, writes a partial sum, takes a ticket; the last block totals the partials in block
// order and adds nll_total's normalisation and constraint terms.  Partials and ticket are SCRATCH:
// they live in a file-scope device array of fixed size, not in memory allocated per call.
//   * limit: the scratch holds NOALLOC_MAX_BLOCKS partial sums + one ticket (9.5 KB, held by every
//     GPU that loads the library until the process exits); the grid is capped to that, signals to
//     NOALLOC_MAX_SIGNALS (shared-memory coefficients), parameters to SXMC_MAX_PARAMETERS, events to
//     NOALLOC_MAX_EVENTS -- larger calls are rejected with SXMC_EINVAL;
//   * every call shares the array and it keeps the last call's contents: the ticket is cleared by a
//     stream-ordered memset at the start of every call (an aborted earlier call cannot poison it),
//     the partials are overwritten before they are read;
//   * therefore calls that OVERLAP on the GPU (two streams, two host threads) race on the scratch:
//     callers serialise them (one stream, or events between streams).
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the
// kernel gets comes from cudaGetSymbolAddress, taken once per device with that device current.
#include <mutex>

#include "common.cuh"
#include "kernels.h"
#include "capi_internal.h"
#include "nll_device.cuh"

namespace sxmc {

constexpr int NOALLOC_THREADS = 256;
constexpr int NOALLOC_MAX_BLOCKS = 148 * 8;
constexpr int NOALLOC_MAX_SIGNALS = 128;
constexpr unsigned long long NOALLOC_MAX_EVENTS = 1ull << 31;
constexpr int NOALLOC_MAX_DEVICES = 64;
constexpr size_t NOALLOC_TICKET_BYTES = 256;                       // the ticket has a 256-byte line to itself
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_TICKET_BYTES + sizeof(double) * NOALLOC_MAX_BLOCKS;

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

__global__ void __launch_bounds__(NOALLOC_THREADS)
nll_lut_noalloc_kernel(const float* __restrict__ lut, long long ne, int ns, const double* __restrict__ pars, int npars,
                       int nsources, const double* __restrict__ means, const double* __restrict__ sigmas,
                       const double* __restrict__ nexpected, const unsigned* __restrict__ n_mc,
                       const short* __restrict__ source_id, const unsigned* __restrict__ norms, double* out2,
                       unsigned char* scratch) {
  __shared__ double s_c[NOALLOC_MAX_SIGNALS];
  __shared__ double s_warp[NOALLOC_THREADS / 32];
  __shared__ int s_last;
  unsigned* const ticket = reinterpret_cast<unsigned*>(scratch);
  double* const partials = reinterpret_cast<double*>(scratch + NOALLOC_TICKET_BYTES);
  const int tid = threadIdx.x;
  for (int j = tid; j < ns; j += NOALLOC_THREADS) {
    const float eff = (float)(1.0 * norms[j] / n_mc[j]);                                  // nll_kernels.cpp:105
    s_c[j] = __dmul_rn(__dmul_rn(pars[source_id[j]], nexpected[j]), (double)eff);
  }
  __syncthreads();
  double sum = 0;
  for (long long i = blockIdx.x * (long long)NOALLOC_THREADS + tid; i < ne; i += (long long)gridDim.x * NOALLOC_THREADS) {
    double s = 0;
    for (int j = 0; j < ns; ++j) {
      const float v = lut[(long long)j * ne + i];
      s = __dadd_rn(s, __dmul_rn(s_c[j], (double)(isnan(v) ? 0.0f : v)));                 // nll_kernels.cpp:107
    }
    if (s > 0) sum += log(s);                                                             // :109
  }
  const double bsum = block_sum<NOALLOC_THREADS>(sum, s_warp);
  if (tid == 0) {
    partials[blockIdx.x] = bsum;
    __threadfence();
    s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double acc = 0;
  for (int b = tid; b < (int)gridDim.x; b += NOALLOC_THREADS) acc += __ldcg(partials + b);
  const double ev = block_sum<NOALLOC_THREADS>(acc, s_warp);
  if (tid != 0) return;
  *ticket = 0u;
  // nll_total (src/nll_kernels.cpp:150-188), as nll_lut_kernel does it
  double t = -ev;
  double nll = t;
  if (isnan(t)) {
    nll = 1e18;
  } else {
    for (int i = 0; i < ns; ++i)
      t = __dadd_rn(t, __dmul_rn(__dmul_rn(pars[source_id[i]], nexpected[i]), (double)norms[i]) / (double)n_mc[i]);
    nll = t;
    for (int i = 0; i < npars; ++i) {
      if (i < nsources && pars[i] < 0) { nll = 1e18; break; }
      if (sigmas[i] > 0) {
        const double x = __dadd_rn(pars[i], -means[i]) / sigmas[i];
        t = __dadd_rn(t, __dmul_rn(__dmul_rn(0.5, x), x));
        nll = t;
      }
    }
  }
  out2[0] = nll;
  out2[1] = ev;
}

// the scratch array's DEVICE address on the current device (taken once per device)
static int noalloc_scratch_ptr(unsigned char** out) {
  static std::mutex mu;
  static unsigned char* ptrs[NOALLOC_MAX_DEVICES] = {nullptr};
  int dev = 0;
  SXMC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) SXMC_INVAL("sxmc_nll_eval_noalloc: device ordinal out of range (< 64)");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    unsigned char* p = nullptr;
    SXMC_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    if (!p) { set_error("sxmc_nll_eval_noalloc: cudaGetSymbolAddress returned no address"); return SXMC_ECUDA; }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return SXMC_OK;
}

}  // namespace sxmc

using namespace sxmc;

extern "C" int sxmc_nll_eval_noalloc(const float* lut_dev, size_t nevents, size_t nsignals, const double* pars_dev,
                                     size_t nparameters, size_t nsources, const double* means_dev,
                                     const double* sigmas_dev, const double* nexpected_dev, const unsigned* n_mc_dev,
                                     const short* source_id_dev, const unsigned* norms_dev, double* out2_dev,
                                     void* cuda_stream) {
  if (!out2_dev || !pars_dev || !means_dev || !sigmas_dev || !nexpected_dev || !n_mc_dev || !source_id_dev || !norms_dev ||
      (!lut_dev && nevents && nsignals))
    SXMC_INVAL("sxmc_nll_eval_noalloc: NULL argument");
  if (nsignals > (size_t)NOALLOC_MAX_SIGNALS) SXMC_INVAL("sxmc_nll_eval_noalloc: more than 128 signals (fixed scratch; use sxmc_nll_eval)");
  if (nparameters > (size_t)SXMC_MAX_PARAMETERS || nsources > nparameters)
    SXMC_INVAL("sxmc_nll_eval_noalloc: nparameters beyond SXMC_MAX_PARAMETERS, or nsources > nparameters");
  if (nevents > NOALLOC_MAX_EVENTS) SXMC_INVAL("sxmc_nll_eval_noalloc: more than 2^31 events (fixed scratch; use sxmc_nll_eval)");
  unsigned char* scratch = nullptr;
  SXMC_TRY(noalloc_scratch_ptr(&scratch));
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  unsigned long long grid = (nevents + NOALLOC_THREADS - 1) / NOALLOC_THREADS;
  if (grid < 1) grid = 1;
  if (grid > (unsigned long long)NOALLOC_MAX_BLOCKS) grid = NOALLOC_MAX_BLOCKS;
  // the array keeps the last call's contents: the ticket must start at zero
  SXMC_CUDA(cudaMemsetAsync(scratch, 0, NOALLOC_TICKET_BYTES, stream));
  nll_lut_noalloc_kernel<<<(unsigned)grid, NOALLOC_THREADS, 0, stream>>>(
      lut_dev, (long long)nevents, (int)nsignals, pars_dev, (int)nparameters, (int)nsources, means_dev, sigmas_dev,
      nexpected_dev, n_mc_dev, source_id_dev, norms_dev, out2_dev, scratch);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}
