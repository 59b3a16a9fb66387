// bin_kernels.cu -- systematic transform + N-D binning (replaces zero_hist + bin_samples,
// reference src/pdfz.cpp:334-408), SetEvalPoints (:245-302) and eval_pdf (:411-436) as
// hand-written sm_100a CUDA.
//
// Bit-exactness contract: every double operation that feeds a bin index is done with the
// explicit round-to-nearest intrinsics (__dadd_rn/__dmul_rn: never contracted into FMA), in
// the order the reference's CPU build performs them (x86-64 g++ -O2 emits no FMA).  The TU
// is also compiled with -fmad=false.
#include "common.cuh"
#include "kernels.h"
#include "bin_device.cuh"

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cstring>
#include <vector>

namespace sxmc {

// =========================================================================================
// generic kernel: any nobs/nfields <= 10, any systematics; per-sample interpreter
// =========================================================================================
template <bool SMEM>
__global__ void __launch_bounds__(BIN_THREADS)
bin_generic_kernel(const BinJob* __restrict__ jobs, const unsigned long long* __restrict__ btile_prefix,
                   int njobs, int smem_bins) {
  extern __shared__ unsigned s_hist[];
  __shared__ SystRt s_syst[SXMC_MAX_SYST];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const unsigned long long total = btile_prefix[njobs];
  unsigned long long t0 = total * blockIdx.x / gridDim.x;
  const unsigned long long t1 = total * (blockIdx.x + 1ull) / gridDim.x;
  if (t0 >= t1) return;
  int j = 0;
  while (btile_prefix[j + 1] <= t0) ++j;

  while (t0 < t1) {
    const BinJob& job = jobs[j];
    const unsigned long long jbeg = btile_prefix[j], jend = btile_prefix[j + 1];
    const unsigned long long tend = t1 < jend ? t1 : jend;
    const int nsyst = job.nsyst, nobs = job.nobs, nfields = job.nfields;
    const int total_nbins = job.total_nbins;
    const bool use_smem = SMEM && total_nbins <= smem_bins;
    if (threadIdx.x < nsyst) resolve_syst(job, threadIdx.x, s_syst[threadIdx.x], false);
    if (use_smem)
      for (int b = threadIdx.x; b < total_nbins; b += BIN_THREADS) s_hist[b] = 0;
    __syncthreads();
    unsigned* hist = use_smem ? s_hist : job.bins;

    RunAcc acc;
    acc.norm = 0;
    for (unsigned long long bt = t0; bt < tend; ++bt) {
      const unsigned long long wt = (bt - jbeg) * BIN_WARPS + warp;
      if (wt >= job.nwtiles) continue;
      acc.cur = -1;
      acc.cnt = 0;
      for (int i = 0; i < LANE_RUN; ++i) {
        const unsigned long long off =
            ((wt * WT_ROWS + (unsigned)(i >> 2)) * 32ull + lane) * 4ull + (i & 3);
        double f[SXMC_MAX_NFIELDS];
        for (int c = 0; c < nfields; ++c) f[c] = (double)__ldg(job.cols[c] + off);
        for (int k = 0; k < nsyst; ++k) {
          const SystRt& s = s_syst[k];
          f[s.obs] = syst_apply_one(s, f[s.obs], f[s.extra]);
        }
        bool in = true;
        int b = 0;
        for (int o = 0; o < nobs; ++o) {
          const double x = f[o];
          if (!(x >= job.lower[o]) || !(x < job.upper[o])) { in = false; break; }
          b += __double2int_rz(__dmul_rn(__dadd_rn(x, -job.lower[o]), job.scale[o])) * job.stride[o];
        }
        run_push<SMEM>(acc, in ? b : -1, hist, total_nbins);
      }
      run_flush<SMEM>(acc, hist, total_nbins);
    }
    const unsigned wnorm = __reduce_add_sync(0xffffffffu, acc.norm);
    if (lane == 0 && wnorm) atomicAdd(job.norm, wnorm);
    if (use_smem) {
      __syncthreads();
      for (int b = threadIdx.x; b < total_nbins; b += BIN_THREADS) {
        const unsigned c = s_hist[b];
        if (c) atomicAdd(job.bins + b, c);
      }
    }
    __syncthreads();
    t0 = tend;
    ++j;
  }
}

// =========================================================================================
// launch dispatch
// =========================================================================================
typedef void (*FastKernel)(const FastParams);

struct FastEntry {
  int nobs, ns, nops;
  unsigned codes;
  FastKernel k[2];   // [global-memory histogram, shared-memory histogram]
};
#define FE(NOBS, NS, NOPS, CODES)                                                   \
  { NOBS, NS, NOPS, (CODES),                                                        \
    { bin_fast_kernel<NOBS, NS, false, NOPS, (CODES)>, bin_fast_kernel<NOBS, NS, true, NOPS, (CODES)> } }
static const FastEntry g_fast_table[] = { SXMC_FAST_TABLE(FE) };
#undef FE

// nops < 0: the dispatching instantiation
static const FastEntry* pick_entry(int nobs, int ns, int nops, unsigned codes) {
  for (const FastEntry& e : g_fast_table)
    if (e.nobs == nobs && e.ns == ns && e.nops == nops && (nops <= 0 || e.codes == codes)) return &e;
  return nullptr;
}
static FastKernel pick_fast(int nobs, int ns, int nops, unsigned codes, bool smem) {
  const FastEntry* e = pick_entry(nobs, ns, nops, codes);
  return e ? e->k[smem ? 1 : 0] : nullptr;
}

bool bin_fast_supported(int nobs, int nslots) {
  return pick_fast(nobs, nslots, -1, 0u, false) != nullptr;
}
bool bin_fast_has_program(int nobs, int nslots, int nops, unsigned codes) {
  return nops >= 0 && pick_fast(nobs, nslots, nops, codes, false) != nullptr;
}

// Shared-memory plan of the fast kernel: [histogram][FAST_WARPS x S stages of NS x R x 512 B].
// Returns false when not even 2 single-row stages fit beside the histogram.
bool bin_fast_plan(int nslots, int hist_bins, size_t static_reserve, int* R_out, int* S_out,
                   unsigned* stage_off_out, size_t* smem_out) {
  const size_t limit = 227 * 1024 - static_reserve;      // opt-in maximum minus the static part
  const size_t hist = ((size_t)hist_bins * sizeof(unsigned) + 127) & ~(size_t)127;
  if (hist >= limit) return false;
  for (int R = 2; R >= 1; --R) {
    const size_t per_stage = (size_t)FAST_WARPS * nslots * R * 512;
    int S = (int)((limit - hist) / per_stage);
    if (S > FAST_MAX_STAGES) S = FAST_MAX_STAGES;
    if (S >= 3 || (R == 1 && S >= 2)) {
      *R_out = R; *S_out = S; *stage_off_out = (unsigned)hist; *smem_out = hist + per_stage * S;
      return true;
    }
  }
  return false;
}

// The same with chunks of FOUR rows for launches that stream a gigabyte or more: per chunk a warp
// spends ~130 instructions on pipeline bookkeeping (barrier wait, expect-tx, one bulk copy per column,
// cursor) whatever the chunk holds, and the kernel is short of issue slots, not of bytes in flight
// -- 2 stages of 4 rows keep as many bytes in flight as 3-4 stages of 2.  16 warps when two such
// stages fit beside the histogram, else 12 (D's four columns: 12 x 2 x 4 x 4 x 512 B = 192 KB;
// measured 299 -> 275 us per step on a power-capped box, where the ~13 % fewer instructions also mean
// higher clocks; plan_launch says where it is taken; profiles/r02_bin_chunk_rows_ab.txt).
bool bin_fast_plan_big(int nslots, int hist_bins, size_t static_reserve, int* nwarps_out, int* R_out, int* S_out,
                       unsigned* stage_off_out, size_t* smem_out) {
  const size_t limit = 227 * 1024 - static_reserve;
  const size_t hist = ((size_t)hist_bins * sizeof(unsigned) + 127) & ~(size_t)127;
  if (hist >= limit) return false;
  const int R = 4;
  for (int nw = FAST_WARPS; nw >= 12; nw -= 4) {
    const size_t per_stage = (size_t)nw * nslots * R * 512;
    int S = (int)((limit - hist) / per_stage);
    if (S > 3) S = 3;
    if (S >= 2) {
      *nwarps_out = nw; *R_out = R; *S_out = S; *stage_off_out = (unsigned)hist; *smem_out = hist + per_stage * S;
      return true;
    }
  }
  return false;
}

static int g_num_sms = 0;
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

void build_fast_params(const BinLaunch& L, int g0, int ng, int grid, FastParams& P,
                       unsigned long long* total_out) {
  memset(&P, 0, sizeof P);
  for (int i = 0; i < 3; ++i) {
    P.geo.lo[i] = L.lower[i]; P.geo.hi[i] = L.upper[i]; P.geo.sc[i] = L.scale[i]; P.geo.st[i] = L.stride[i];
  }
  P.geo.total_nbins = L.max_total_nbins;
  P.R = L.rows_per_stage; P.S = L.nstages; P.stage_off = L.stage_off; P.njobs = ng;
  P.nwarps = L.nwarps > 0 ? L.nwarps : FAST_WARPS;
  unsigned long long total = 0;
  P.prefix[0] = 0;
  for (int j = 0; j < ng; ++j) {
    const BinJob& src = L.jobs_host[g0 + j];
    FastJob& d = P.job[j];
    for (int c = 0; c < MAX_SLOTS; ++c) d.col[c] = src.cols[src.slot_field[c < src.nslots ? c : 0]];
    d.bins = src.bins; d.norm = src.norm; d.params = src.params;
    d.param_stride = (int)src.param_stride;
    d.nsyst = src.nsyst;
    for (int k = 0; k < src.nsyst && k < FAST_MAX_OPS; ++k) {
      d.code[k] = (unsigned short)(src.syst[k].type | (src.syst[k].obs_slot << 4) | (src.syst[k].extra_slot << 8));
      d.pidx[k] = src.syst[k].pidx[0];
      d.npars[k] = (unsigned char)(src.syst[k].npars > 0);
    }
    total += src.nwtiles * (unsigned long long)(WT_ROWS / L.rows_per_stage);
    P.prefix[j + 1] = total;
  }
  *total_out = total;
  // Small fits: give every job whole CTAs (no CTA spans two jobs).  Big fits keep the even
  // split, whose balance is perfect and whose second segment costs nothing in proportion.
  P.aligned = 0;
  P.cta_split = 0;
  int nonempty = 0;
  for (int j = 0; j < ng; ++j) nonempty += P.prefix[j + 1] > P.prefix[j];
  if (total > 0 && nonempty <= grid && total / ((unsigned long long)grid * P.nwarps) < 32) {
    int cta[FAST_MAX_JOBS];
    int used = 0;
    for (int j = 0; j < ng; ++j) {
      const unsigned long long nj = P.prefix[j + 1] - P.prefix[j];
      cta[j] = nj ? (int)std::max<unsigned long long>(1, nj * grid / total) : 0;
      used += cta[j];
    }
    while (used != grid) {      // hand out / take back CTAs where it changes the load least
      int best = -1;
      double score = 0;
      for (int j = 0; j < ng; ++j) {
        const unsigned long long nj = P.prefix[j + 1] - P.prefix[j];
        if (!nj || (used > grid && cta[j] <= 1)) continue;
        const double load = used < grid ? (double)nj / cta[j] : -(double)nj / (cta[j] - 1);
        if (best < 0 || load > score) { best = j; score = load; }
      }
      if (best < 0) break;
      cta[best] += used < grid ? 1 : -1;
      used += used < grid ? 1 : -1;
    }
    if (used == grid) {
      P.cta_first[0] = 0;
      for (int j = 0; j < ng; ++j) P.cta_first[j + 1] = (unsigned short)(P.cta_first[j] + cta[j]);
      P.aligned = 1;
    }
  }
}

static unsigned long long* g_bin_xdbg = nullptr;   // profiling hook, NULL in production
void set_bin_xdbg(unsigned long long* dev) { g_bin_xdbg = dev; }

int launch_bin(const BinLaunch& L, cudaStream_t stream) {
  if (L.total_units == 0) return SXMC_OK;
  if (L.fast) {
    FastKernel fk = pick_fast(L.nobs, L.nslots, L.prog_nops, L.prog_codes, L.smem_hist);
    void* jk = nullptr;          // the fit's own program, compiled at run time (jit.cu)
    if (!fk && L.jit) jk = jit_bin_fast_kernel(L.nobs, L.nslots, L.smem_hist, L.prog_nops, L.prog_codes);
    if (!fk && !jk) fk = pick_fast(L.nobs, L.nslots, -1, 0u, L.smem_hist);      // run-time dispatch
    if (!fk && !jk) { set_error("launch_bin: no fast kernel for this shape"); return SXMC_ESTATE; }
    if (fk) SXMC_CUDA(cudaFuncSetAttribute((const void*)fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes));
    // jobs travel in the kernel parameters (constant bank), FAST_MAX_JOBS per launch
    for (int g0 = 0; g0 < L.njobs; g0 += FAST_MAX_JOBS) {
      const int ng = L.njobs - g0 < FAST_MAX_JOBS ? L.njobs - g0 : FAST_MAX_JOBS;
      FastParams P;
      unsigned long long total = 0;
      build_fast_params(L, g0, ng, 0, P, &total);
      if (total == 0) continue;
      // one persistent CTA per SM; at least one chunk per warp
      unsigned long long grid = (unsigned long long)num_sms();
      const unsigned nwarps = (unsigned)(L.nwarps > 0 ? L.nwarps : FAST_WARPS);
      const unsigned long long max_grid = (total + nwarps - 1) / nwarps;
      if (grid > max_grid) grid = max_grid;
      build_fast_params(L, g0, ng, (int)grid, P, &total);
      P.xdbg = g_bin_xdbg;
      P.cta_split = L.cta_split;
      const bool pdl = L.pdl && L.njobs <= FAST_MAX_JOBS;
      if (jk) {
        const int rc = jit_launch_bin(jk, (unsigned)grid, L.smem_bytes, stream, pdl, P);
        if (rc) return rc;
      } else {
        SXMC_CUDA(launch_kernel_pdl(fk, dim3((unsigned)grid), dim3(nwarps * 32u), L.smem_bytes, stream, pdl, P));
      }
    }
    return SXMC_OK;
  }
  const size_t smem = L.smem_hist ? sizeof(unsigned) * (size_t)L.max_total_nbins : 0;
  const void* fn = L.smem_hist ? (const void*)bin_generic_kernel<true> : (const void*)bin_generic_kernel<false>;
  if (smem > 48 * 1024)
    SXMC_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  SXMC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, BIN_THREADS, smem));
  if (occ < 1) occ = 1;
  if (L.blocks_per_sm > 0 && L.blocks_per_sm < occ) occ = L.blocks_per_sm;
  unsigned long long grid = (unsigned long long)num_sms() * occ;
  if (grid > L.total_units) grid = L.total_units;
  if (L.smem_hist) {
    bin_generic_kernel<true><<<(unsigned)grid, BIN_THREADS, smem, stream>>>(
        L.jobs_dev, L.prefix_dev, L.njobs, L.max_total_nbins);
  } else {
    bin_generic_kernel<false><<<(unsigned)grid, BIN_THREADS, 0, stream>>>(
        L.jobs_dev, L.prefix_dev, L.njobs, 0);
  }
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

// zero_hist (src/pdfz.cpp:334-346) for every job of a batch in one launch: block j clears
// jobs[j].bins and jobs[j].norm
__global__ void zero_jobs_kernel(const BinJob* __restrict__ jobs) {
  const BinJob& job = jobs[blockIdx.x];
  for (int b = threadIdx.x; b < job.total_nbins; b += blockDim.x) job.bins[b] = 0u;
  if (threadIdx.x == 0) *job.norm = 0u;
}

int launch_zero_jobs(const BinJob* jobs_dev, int njobs, cudaStream_t stream) {
  if (njobs <= 0) return SXMC_OK;
  zero_jobs_kernel<<<njobs, 256, 0, stream>>>(jobs_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

// =========================================================================================
// layout preparation: sort key, gather into the tiled SoA layout
// =========================================================================================
__device__ __forceinline__ unsigned ordered_bits(float v) {
  unsigned u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// key = (nominal flat bin with every coordinate clamped into the histogram) << 32 |
//       order-preserving bits of the primary field (the observable the first systematic moves)
__global__ void make_keys_kernel(const float* __restrict__ aos, unsigned long long n, int nfields,
                                 int nobs, PrepGeo g, int primary, unsigned long long* keys,
                                 unsigned* idx) {
  const unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* row = aos + i * nfields;
  unsigned long long b = 0;
  for (int o = 0; o < nobs; ++o) {
    const double x = (double)row[o];
    int k = g.nbins[o] - 1;   // NaN and >= upper go last
    if (x < g.upper[o]) {
      k = __double2int_rz((x - g.lower[o]) * g.scale[o]);
      if (!(x >= g.lower[o])) k = 0;
      if (k > g.nbins[o] - 1) k = g.nbins[o] - 1;
      if (k < 0) k = 0;
    }
    b += (unsigned long long)k * (unsigned)g.stride[o];
  }
  keys[i] = (b << 32) | ordered_bits(row[primary]);
  idx[i] = (unsigned)i;
}

__global__ void iota_kernel(unsigned* idx, unsigned long long n) {
  const unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (i < n) idx[i] = (unsigned)i;
}

__global__ void gather_tiled_kernel(const float* __restrict__ aos, const unsigned* __restrict__ idx,
                                    unsigned long long n, unsigned long long npad, int nfields,
                                    int nobs, float* cols, unsigned long long col_pitch,
                                    unsigned* perm_out) {
  const unsigned long long s = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (s >= npad) return;
  const unsigned long long off = tiled_offset(s);
  if (s < n) {
    const unsigned src = idx[s];
    if (perm_out) perm_out[s] = src;
    for (int f = 0; f < nfields; ++f) cols[f * col_pitch + off] = aos[(unsigned long long)src * nfields + f];
  } else {
    // padding: +inf observables are outside every domain under every transform
    for (int f = 0; f < nfields; ++f)
      cols[f * col_pitch + off] = f < nobs ? __int_as_float(0x7f800000) : 0.0f;
  }
}

// inverse of gather_tiled: original-order AoS rows (EvalHist::GetSamples)
__global__ void scatter_aos_kernel(const float* __restrict__ cols, unsigned long long col_pitch,
                                   const unsigned* __restrict__ perm, unsigned long long n,
                                   int nfields, float* aos) {
  const unsigned long long s = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (s >= n) return;
  const unsigned long long off = tiled_offset(s);
  const unsigned dst = perm[s];
  for (int f = 0; f < nfields; ++f) aos[(unsigned long long)dst * nfields + f] = cols[f * col_pitch + off];
}

int prepare_layout(const float* aos_dev, unsigned long long n, int nfields, int nobs,
                   const PrepGeo& g, int primary, bool sort, float* cols_dev,
                   unsigned long long col_pitch, unsigned* perm_dev, cudaStream_t stream) {
  const unsigned long long npad = col_pitch;
  if (n >= (1ull << 32)) { set_error("prepare_layout: more than 2^32-1 samples in one PDF"); return SXMC_EINVAL; }
  unsigned* idx = nullptr;
  unsigned long long* keys = nullptr;
  unsigned* idx_alt = nullptr;
  unsigned long long* keys_alt = nullptr;
  void* tmp = nullptr;
  const unsigned nblk = (unsigned)((n + 255) / 256);
  int rc = SXMC_OK;
  if (n > 0) {
    SXMC_CUDA(cudaMalloc(&idx, sizeof(unsigned) * n));
    if (sort) {
      SXMC_CUDA(cudaMalloc(&keys, sizeof(unsigned long long) * n));
      SXMC_CUDA(cudaMalloc(&keys_alt, sizeof(unsigned long long) * n));
      SXMC_CUDA(cudaMalloc(&idx_alt, sizeof(unsigned) * n));
      make_keys_kernel<<<nblk, 256, 0, stream>>>(aos_dev, n, nfields, nobs, g, primary, keys, idx);
      cub::DoubleBuffer<unsigned long long> dk(keys, keys_alt);
      cub::DoubleBuffer<unsigned> dv(idx, idx_alt);
      size_t tmp_bytes = 0;
      int end_bit = 32;
      while (end_bit < 64 && (g.total_nbins >> (end_bit - 32)) != 0) ++end_bit;
      SXMC_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (long long)n, 0, end_bit, stream));
      SXMC_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
      SXMC_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, (long long)n, 0, end_bit, stream));
      if (dv.Current() != idx) { unsigned* t = idx; idx = idx_alt; idx_alt = t; }
    } else {
      iota_kernel<<<nblk, 256, 0, stream>>>(idx, n);
    }
  }
  if (npad > 0) {
    const unsigned nb2 = (unsigned)((npad + 255) / 256);
    gather_tiled_kernel<<<nb2, 256, 0, stream>>>(aos_dev, idx, n, npad, nfields, nobs, cols_dev,
                                                 col_pitch, perm_dev);
  }
  cudaError_t e = cudaStreamSynchronize(stream);
  if (e == cudaSuccess) e = cudaGetLastError();
  cudaFree(idx); cudaFree(keys); cudaFree(keys_alt); cudaFree(idx_alt); cudaFree(tmp);
  if (e != cudaSuccess) rc = cuda_fail(e, "prepare_layout", __FILE__, __LINE__);
  return rc;
}

int unprepare_to_aos(const float* cols_dev, unsigned long long col_pitch, const unsigned* perm_dev,
                     unsigned long long n, int nfields, float* aos_dev, cudaStream_t stream) {
  if (n == 0) return SXMC_OK;
  scatter_aos_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(cols_dev, col_pitch, perm_dev,
                                                                      n, nfields, aos_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

// =========================================================================================
// SetEvalPoints (src/pdfz.cpp:245-302) and eval_pdf (src/pdfz.cpp:411-436)
// =========================================================================================
__global__ void read_bins_kernel(const float* __restrict__ points, long long npoints, int nobs,
                                 PrepGeo g, unsigned dataset, int* __restrict__ read_bins) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= npoints) return;
  const float* row = points + i * (nobs + 1);
  bool in = true;
  int b = 0;
  for (int o = 0; o < nobs; ++o) {
    const double x = (double)row[o];
    if (!(x >= g.lower[o]) || !(x < g.upper[o])) { in = false; break; }
    b += __double2int_rz(__dmul_rn(__dadd_rn(x, -g.lower[o]), g.scale[o])) * g.stride[o];
  }
  // `points[idx] != this->dataset`: float vs unsigned compares in float (:290-293)
  if (row[nobs] != (float)dataset) b = -2;
  read_bins[i] = in ? b : -1;
}

int launch_read_bins(const float* points_dev, long long npoints, int nobs, const PrepGeo& g,
                     unsigned dataset, int* read_bins_dev, cudaStream_t stream) {
  if (npoints <= 0) return SXMC_OK;
  read_bins_kernel<<<(unsigned)((npoints + 255) / 256), 256, 0, stream>>>(points_dev, npoints, nobs, g,
                                                                          dataset, read_bins_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

__global__ void eval_pdf_kernel(const EvalPdfJob* __restrict__ jobs, int njobs) {
  const EvalPdfJob job = jobs[blockIdx.y];
  // bin_norm = *norm * bin_volume  (unsigned -> double, :417)
  const double bin_norm = __dmul_rn((double)*job.norm, job.bin_volume);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < job.npoints;
       i += (long long)gridDim.x * blockDim.x) {
    const int b = job.read_bins[i];
    float v;
    if (b == -2) v = 0.0f;
    else if (b < 0) v = __int_as_float(0x7fc00000);
    else if (b >= job.total_nbins) v = 0.0f;   // reference reads past the array (undefined)
    else v = pdf_value(job.bins[b], bin_norm);
    job.out[job.stride * i] = v;
  }
}

int launch_eval_pdf(const EvalPdfJob* jobs_dev, int njobs, long long max_points, cudaStream_t stream) {
  if (njobs <= 0 || max_points <= 0) return SXMC_OK;
  long long gx = (max_points + 255) / 256;
  if (gx > 1024) gx = 1024;
  dim3 grid((unsigned)gx, (unsigned)njobs);
  eval_pdf_kernel<<<grid, 256, 0, stream>>>(jobs_dev, njobs);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

}  // namespace sxmc
