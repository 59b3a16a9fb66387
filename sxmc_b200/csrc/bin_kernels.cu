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

#include <cub/device/device_radix_sort.cuh>

namespace sxmc {

// =========================================================================================
// shared helpers
// =========================================================================================

// Per-launch geometry (identical for every job of a fast-path launch), passed by value so it
// sits in the constant bank and costs no registers.
typedef FastGeo Geo;

struct SystRt {       // one systematic, resolved for the current parameter vector
  int type, obs, extra, npars;
  double c0;          // 0.0 + par0 * pow(x,0)   (src/pdfz.cpp:310-314 with i = 0)
  double c1, c2, c3;  // raw coefficients of x, x^2, x^3
  double onep;        // 1 + p, valid when npars == 1 (SCALE / CTSCALE, src/pdfz.cpp:320,323)
};

__device__ __forceinline__ void resolve_syst(const BinJob& job, int k, SystRt& s, bool slots) {
  const SystDesc& d = job.syst[k];
  s.type = d.type;
  s.obs = slots ? d.obs_slot : d.obs;
  s.extra = slots ? d.extra_slot : d.extra;
  s.npars = d.npars;
  double c[SXMC_MAX_NPARS] = {0, 0, 0, 0};
  for (int i = 0; i < d.npars && i < SXMC_MAX_NPARS; ++i)
    c[i] = job.params[(long long)d.pidx[i] * job.param_stride];
  s.c0 = __dadd_rn(0.0, __dmul_rn(c[0], 1.0));
  s.c1 = c[1]; s.c2 = c[2]; s.c3 = c[3];
  s.onep = __dadd_rn(1.0, s.c0);
}

// p = sum_i par_i * pow(x, i) accumulated left to right (src/pdfz.cpp:310-314).
// pow(x,0)=1 and pow(x,1)=x are exact; pow(x,2.0)==x*x for glibc; i=3 uses the
// device pow (<=1 ulp-level difference from glibc possible; documented in DESIGN.md).
__device__ __forceinline__ double syst_poly(const SystRt& s, double x) {
  double p = s.c0;
  if (s.npars > 1) {
    p = __dadd_rn(p, __dmul_rn(s.c1, x));
    if (s.npars > 2) {
      p = __dadd_rn(p, __dmul_rn(s.c2, __dmul_rn(x, x)));
      if (s.npars > 3) p = __dadd_rn(p, __dmul_rn(s.c3, pow(x, 3.0)));
    }
  }
  return p;
}

// src/pdfz.cpp:316-330
__device__ __forceinline__ double syst_apply_one(const SystRt& s, double x, double t) {
  const double p = syst_poly(s, x);
  switch (s.type) {
    case SXMC_SYST_SHIFT: return __dadd_rn(x, p);
    case SXMC_SYST_SCALE: return __dmul_rn(x, __dadd_rn(1.0, p));
    case SXMC_SYST_CTSCALE:
      return __dadd_rn(1.0, __dmul_rn(__dadd_rn(x, -1.0), __dadd_rn(1.0, p)));
    case SXMC_SYST_RESOLUTION_SCALE: return __dadd_rn(x, __dmul_rn(p, __dadd_rn(x, -t)));
  }
  return x;
}

// Run-length aggregation of histogram updates: a thread walks a sorted run, so `b` changes
// rarely; one atomic per run instead of one per sample.
struct RunAcc {
  int cur;
  unsigned cnt;
  unsigned norm;
};

template <bool SMEM>
__device__ __forceinline__ void run_flush(RunAcc& r, unsigned* hist, int total_nbins) {
  if (r.cur >= 0) {
    r.norm += r.cnt;                       // in-domain count (src/pdfz.cpp:401-403)
    if (r.cur < total_nbins) atomicAdd(hist + r.cur, r.cnt);
  }
}

template <bool SMEM>
__device__ __forceinline__ void run_push(RunAcc& r, int b, unsigned* hist, int total_nbins) {
  if (b != r.cur) {
    run_flush<SMEM>(r, hist, total_nbins);
    r.cur = b;
    r.cnt = 0;
  }
  r.cnt += 1;
}

__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}

// =========================================================================================
// fast kernel: NOBS <= 3 observables, NS <= MAX_SLOTS register-resident columns
// =========================================================================================

// All systematics of a fast launch have a single coefficient (npars <= 1): p is the same for
// every sample, resolved once per segment into SystRt::c0 / onep.  The (type, slot, slot)
// triple of each systematic is warp-uniform; it is dispatched with switch statements whose
// cases index the slot array with compile-time constants only, so the 4-sample working set
// stays in registers (a loop with `if (s.obs != o) continue` gets rewritten by the compiler
// into a dynamically indexed local-memory array).

template <int NS, int O, int E>
__device__ __forceinline__ void op_resolution(double (&f)[NS][4], const double c0) {
#pragma unroll
  for (int k = 0; k < 4; ++k)
    f[O][k] = __dadd_rn(f[O][k], __dmul_rn(c0, __dadd_rn(f[O][k], -f[E][k])));
}

template <int NS, int O>
__device__ __forceinline__ void op_on(double (&f)[NS][4], const int type, const int ext,
                                      const double c0, const double onep) {
  if (type == SXMC_SYST_SHIFT) {
#pragma unroll
    for (int k = 0; k < 4; ++k) f[O][k] = __dadd_rn(f[O][k], c0);
  } else if (type == SXMC_SYST_SCALE) {
#pragma unroll
    for (int k = 0; k < 4; ++k) f[O][k] = __dmul_rn(f[O][k], onep);
  } else if (type == SXMC_SYST_CTSCALE) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      f[O][k] = __dadd_rn(1.0, __dmul_rn(__dadd_rn(f[O][k], -1.0), onep));
  } else {  // RESOLUTION_SCALE
    switch (ext) {
      case 0: op_resolution<NS, O, 0>(f, c0); break;
      case 1: if constexpr (NS > 1) op_resolution<NS, O, 1>(f, c0); break;
      case 2: if constexpr (NS > 2) op_resolution<NS, O, 2>(f, c0); break;
      case 3: if constexpr (NS > 3) op_resolution<NS, O, 3>(f, c0); break;
      case 4: if constexpr (NS > 4) op_resolution<NS, O, 4>(f, c0); break;
      default: if constexpr (NS > 5) op_resolution<NS, O, 5>(f, c0); break;
    }
  }
}

// code = type | obs_slot << 4 | extra_slot << 8
template <int NS>
__device__ __forceinline__ void apply_op(double (&f)[NS][4], const int code, const double c0,
                                         const double onep) {
  const int type = code & 15, ext = code >> 8;
  switch ((code >> 4) & 15) {
    case 0: op_on<NS, 0>(f, type, ext, c0, onep); break;
    case 1: if constexpr (NS > 1) op_on<NS, 1>(f, type, ext, c0, onep); break;
    case 2: if constexpr (NS > 2) op_on<NS, 2>(f, type, ext, c0, onep); break;
    case 3: if constexpr (NS > 3) op_on<NS, 3>(f, type, ext, c0, onep); break;
    case 4: if constexpr (NS > 4) op_on<NS, 4>(f, type, ext, c0, onep); break;
    default: if constexpr (NS > 5) op_on<NS, 5>(f, type, ext, c0, onep); break;
  }
}

// Compile-time systematic programs.  A program is up to FAST_MAX_OPS op codes of 8 bits,
// code = type | slot_obs << 2 | slot_extra << 5, packed little-endian in CODES; NOPS < 0 selects
// the run-time dispatch above.  Kernels are instantiated for the programs of the BASELINE
// workloads (table in pick_fast); any other fit runs the dispatching instantiation.
template <int NS, int CODE>
__device__ __forceinline__ void apply_static(double (&f)[NS][4], const double c0, const double onep) {
  constexpr int type = CODE & 3, O = (CODE >> 2) & 7, E = (CODE >> 5) & 7;
  static_assert(O < NS && E < NS, "systematic program addresses a slot the kernel does not load");
  if constexpr (type == SXMC_SYST_RESOLUTION_SCALE) {
    op_resolution<NS, O, E>(f, c0);
  } else {
    op_on<NS, O>(f, type, 0, c0, onep);   // `type` is a constant here: the chain folds
  }
}

// (int)v for 0 <= v < 2^31 without the conversion pipe: v + 2^52 rounded toward -inf leaves
// floor(v) in the low mantissa word.  Only used on in-domain samples (v >= 0).
__device__ __forceinline__ int trunc_nonneg(double v) {
  return __double2loint(__dadd_rd(v, 4503599627370496.0));
}

// flat bin of one transformed sample, -1 when outside the PDF domain (src/pdfz.cpp:386-398).
// NaN compares false on both tests => outside (the reference's (int)NaN is undefined).
template <int NOBS>
__device__ __forceinline__ int flat_bin(const double (&x)[NOBS], const Geo& g) {
  bool in = true;
  int b = 0;
#pragma unroll
  for (int o = 0; o < NOBS; ++o) {
    in = in && (x[o] >= g.lo[o]) && (x[o] < g.hi[o]);
    const int i = trunc_nonneg(__dmul_rn(__dadd_rn(x[o], -g.lo[o]), g.sc[o]));
    b += (NOBS == 1 ? i : i * g.st[o]);
  }
  return in ? b : -1;
}

// four consecutive samples of a lane's sorted run: usually all in the current bin
template <bool SMEM>
__device__ __forceinline__ void run_push4(RunAcc& r, const int (&b)[4], unsigned* hist, int total_nbins) {
  if (b[0] == r.cur && b[1] == r.cur && b[2] == r.cur && b[3] == r.cur) {
    r.cnt += 4;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) run_push<SMEM>(r, b[k], hist, total_nbins);
  }
}

// ---- mbarrier / bulk-copy (TMA unit, SASS UBLKCP) primitives ---------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned ok = 0;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ float4 lds128(unsigned addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr));
  return v;
}

// One persistent CTA per SM, FAST_WARPS warps.  Work unit = a "chunk": R consecutive rows
// (R x 512 B per column) of a warp-tile.  The chunk sequence of all jobs is cut evenly over
// the CTAs; inside a CTA each job segment is cut into one CONTIGUOUS chunk range per warp.
// Every warp runs its own S-stage pipeline: lane 0 issues one bulk async copy (TMA unit)
// per column per chunk into the warp's private shared-memory stages, completion is signalled
// on a per-stage mbarrier, and the warp consumes the stage with conflict-free LDS.128 --
// the loads in flight live in shared memory, not in registers, and no CTA-wide barrier is
// needed in steady state.
template <int NOBS, int NS, bool SMEM, int NOPS, unsigned CODES>
__global__ void __launch_bounds__(FAST_THREADS, 1)
bin_fast_kernel(const __grid_constant__ FastParams P) {
  const Geo& geo = P.geo;
  const int R = P.R, S = P.S;
  const unsigned stage_off = P.stage_off;
  const int njobs = P.njobs;
  const unsigned long long* chunk_prefix = P.prefix;   // constant bank: no global round trip
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ SystRt s_syst[SXMC_MAX_SYST];
  __shared__ __align__(8) unsigned long long s_bar[FAST_WARPS * FAST_MAX_STAGES];
  unsigned* s_hist = reinterpret_cast<unsigned*>(smem_raw);

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const unsigned col_bytes = (unsigned)R * 512u;          // one column of one chunk
  const unsigned stage_bytes = NS * col_bytes;
  const unsigned my_stage0 = smem_u32(smem_raw + stage_off) + (unsigned)(warp * S) * stage_bytes;
  const unsigned my_bar0 = smem_u32(&s_bar[warp * FAST_MAX_STAGES]);
  if (lane == 0) {
    for (int s = 0; s < S; ++s) mbar_init(my_bar0 + 8u * s, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  const unsigned long long total = chunk_prefix[njobs];
  unsigned long long t0 = total * blockIdx.x / gridDim.x;
  const unsigned long long t1 = total * (blockIdx.x + 1ull) / gridDim.x;
  if (t0 >= t1) return;
  int j = 0;
  while (chunk_prefix[j + 1] <= t0) ++j;
  unsigned it = 0;    // chunks this warp has consumed so far: stage = it % S, parity = (it / S) & 1

  while (t0 < t1) {
    const FastJob& job = P.job[j];
    const unsigned long long jbeg = chunk_prefix[j], jend = chunk_prefix[j + 1];
    const unsigned long long tend = t1 < jend ? t1 : jend;
    const int nsyst = job.nsyst;

    const unsigned char* col[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) col[c] = reinterpret_cast<const unsigned char*>(job.col[c]);

    // this warp's contiguous chunk range [ca, cb) of the job (chunk c starts at byte c * col_bytes)
    const unsigned long long n = tend - t0;
    const unsigned long long ca = (t0 - jbeg) + n * (unsigned)warp / FAST_WARPS;
    const unsigned long long cb = (t0 - jbeg) + n * (unsigned)(warp + 1) / FAST_WARPS;

    auto issue = [&](unsigned long long c, unsigned slot) {   // lane 0 only
      const unsigned bar = my_bar0 + 8u * slot;
      const unsigned dst = my_stage0 + slot * stage_bytes;
      mbar_expect_tx(bar, stage_bytes);
#pragma unroll
      for (int k = 0; k < NS; ++k) bulk_g2s(dst + k * col_bytes, col[k] + c * col_bytes, col_bytes, bar);
    };
    // first thing in the segment: get the copies going (everything they need is in the
    // constant bank); the parameter loads and the histogram clear overlap with them
    if (lane == 0)
      for (unsigned i = 0; i < (unsigned)S && ca + i < cb; ++i) issue(ca + i, (it + i) % (unsigned)S);

    if (threadIdx.x < nsyst) {   // p = 0 + par * pow(x, 0)  (src/pdfz.cpp:310-314, npars <= 1)
      SystRt& sr = s_syst[threadIdx.x];
      const double par = job.npars[threadIdx.x]
                             ? job.params[(long long)job.pidx[threadIdx.x] * job.param_stride] : 0.0;
      sr.c0 = __dadd_rn(0.0, __dmul_rn(par, 1.0));
      sr.onep = __dadd_rn(1.0, sr.c0);
    }
    if (SMEM)
      for (int b = threadIdx.x; b < geo.total_nbins; b += FAST_THREADS) s_hist[b] = 0;
    __syncthreads();

    // the segment's systematics, in registers
    int op_code[FAST_MAX_OPS];
    double op_c0[FAST_MAX_OPS], op_onep[FAST_MAX_OPS];
#pragma unroll
    for (int k = 0; k < FAST_MAX_OPS; ++k) {
      op_code[k] = 0; op_c0[k] = 0.0; op_onep[k] = 1.0;
      if (k < nsyst) {
        op_code[k] = job.code[k];
        op_c0[k] = s_syst[k].c0;
        op_onep[k] = s_syst[k].onep;
      }
    }

    unsigned* hist = SMEM ? s_hist : job.bins;

    RunAcc acc;
    acc.norm = 0;
    acc.cur = -1;
    acc.cnt = 0;
#pragma unroll 1
    for (unsigned long long c = ca; c < cb; ++c, ++it) {
      const unsigned slot = it % (unsigned)S;
      mbar_wait(my_bar0 + 8u * slot, (it / (unsigned)S) & 1u);
      const unsigned src = my_stage0 + slot * stage_bytes + (unsigned)lane * 16u;
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
        run_push4<SMEM>(acc, bq, hist, geo.total_nbins);
      }
      __syncwarp();                       // every lane has read the stage before it is refilled
      if (lane == 0 && c + (unsigned)S < cb) issue(c + (unsigned)S, slot);
    }
    run_flush<SMEM>(acc, hist, geo.total_nbins);

    const unsigned wnorm = __reduce_add_sync(0xffffffffu, acc.norm);
    if (lane == 0 && wnorm) atomicAdd(job.norm, wnorm);
    if (SMEM) {
      __syncthreads();
      for (int b = threadIdx.x; b < geo.total_nbins; b += FAST_THREADS) {
        const unsigned cnt = s_hist[b];
        if (cnt) atomicAdd(job.bins + b, cnt);
      }
    }
    __syncthreads();   // s_syst / s_hist are rewritten by the next segment
    t0 = tend;
    ++j;
  }
}

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

#define OPC(type, obs, ext) ((unsigned)((type) | ((obs) << 2) | ((ext) << 5)))
struct FastEntry {
  int nobs, ns, nops;
  unsigned codes;
  FastKernel k[2];   // [global-memory histogram, shared-memory histogram]
};
#define FE(NOBS, NS, NOPS, CODES)                                                   \
  { NOBS, NS, NOPS, (CODES),                                                        \
    { bin_fast_kernel<NOBS, NS, false, NOPS, (CODES)>, bin_fast_kernel<NOBS, NS, true, NOPS, (CODES)> } }
static const FastEntry g_fast_table[] = {
    // run-time dispatch of the systematics (any program of <= FAST_MAX_OPS single-coefficient ops)
    FE(1, 1, -1, 0u), FE(1, 2, -1, 0u), FE(1, 3, -1, 0u), FE(2, 2, -1, 0u), FE(2, 3, -1, 0u),
    FE(2, 4, -1, 0u), FE(3, 3, -1, 0u), FE(3, 4, -1, 0u), FE(3, 5, -1, 0u),
    // no systematics
    FE(1, 1, 0, 0u), FE(2, 2, 0, 0u), FE(3, 3, 0, 0u),
    // workload B / bench_pdfz: shift(energy)
    FE(1, 1, 1, OPC(SXMC_SYST_SHIFT, 0, 0)),
    FE(1, 1, 1, OPC(SXMC_SYST_SCALE, 0, 0)),
    // workload A: scale(energy), resolution_scale(energy; mc_energy)
    FE(1, 2, 2, OPC(SXMC_SYST_SCALE, 0, 0) | OPC(SXMC_SYST_RESOLUTION_SCALE, 0, 1) << 8),
    // workload C: the same on 2-D energy x radius
    FE(2, 3, 2, OPC(SXMC_SYST_SCALE, 0, 0) | OPC(SXMC_SYST_RESOLUTION_SCALE, 0, 2) << 8),
    // workload D: scale(energy), resolution_scale(energy; mc_energy), shift(radius) on 3-D
    FE(3, 4, 3, OPC(SXMC_SYST_SCALE, 0, 0) | OPC(SXMC_SYST_RESOLUTION_SCALE, 0, 3) << 8 |
                    OPC(SXMC_SYST_SHIFT, 1, 0) << 16),
};

// nops < 0: the dispatching instantiation
static FastKernel pick_fast(int nobs, int ns, int nops, unsigned codes, bool smem) {
  for (const FastEntry& e : g_fast_table)
    if (e.nobs == nobs && e.ns == ns && e.nops == nops && (nops <= 0 || e.codes == codes))
      return e.k[smem ? 1 : 0];
  return nullptr;
}

bool bin_fast_supported(int nobs, int nslots) {
  return pick_fast(nobs, nslots, -1, 0u, false) != nullptr;
}
bool bin_fast_has_program(int nobs, int nslots, int nops, unsigned codes) {
  return nops >= 0 && pick_fast(nobs, nslots, nops, codes, false) != nullptr;
}

// Shared-memory plan of the fast kernel: [histogram][FAST_WARPS x S stages of NS x R x 512 B].
// Returns false when not even 2 single-row stages fit beside the histogram.
bool bin_fast_plan(int nslots, int hist_bins, int* R_out, int* S_out, unsigned* stage_off_out,
                   size_t* smem_out) {
  const size_t limit = 227 * 1024 - 2048;      // opt-in maximum minus the static part
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

int launch_bin(const BinLaunch& L, cudaStream_t stream) {
  if (L.total_units == 0) return SXMC_OK;
  if (L.fast) {
    FastKernel fk = pick_fast(L.nobs, L.nslots, L.prog_nops, L.prog_codes, L.smem_hist);
    if (!fk) { set_error("launch_bin: no fast kernel for this shape"); return SXMC_ESTATE; }
    SXMC_CUDA(cudaFuncSetAttribute((const void*)fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem_bytes));
    // jobs travel in the kernel parameters (constant bank), FAST_MAX_JOBS per launch
    for (int g0 = 0; g0 < L.njobs; g0 += FAST_MAX_JOBS) {
      const int ng = L.njobs - g0 < FAST_MAX_JOBS ? L.njobs - g0 : FAST_MAX_JOBS;
      FastParams P;
      memset(&P, 0, sizeof P);
      for (int i = 0; i < 3; ++i) {
        P.geo.lo[i] = L.lower[i]; P.geo.hi[i] = L.upper[i]; P.geo.sc[i] = L.scale[i]; P.geo.st[i] = L.stride[i];
      }
      P.geo.total_nbins = L.max_total_nbins;
      P.R = L.rows_per_stage; P.S = L.nstages; P.stage_off = L.stage_off; P.njobs = ng;
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
      if (total == 0) continue;
      // one persistent CTA per SM; at least one chunk per warp
      unsigned long long grid = (unsigned long long)num_sms();
      const unsigned long long max_grid = (total + FAST_WARPS - 1) / FAST_WARPS;
      if (grid > max_grid) grid = max_grid;
      fk<<<(unsigned)grid, FAST_THREADS, L.smem_bytes, stream>>>(P);
      SXMC_CUDA(cudaGetLastError());
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

// sorted position s -> tiled position (see common.cuh)
__host__ __device__ __forceinline__ unsigned long long tiled_offset(unsigned long long s) {
  const unsigned long long wt = s / WT_SAMPLES;
  const unsigned r = (unsigned)(s % WT_SAMPLES);
  const unsigned lane = r / LANE_RUN, i = r % LANE_RUN;
  return ((wt * WT_ROWS + (i >> 2)) * 32ull + lane) * 4ull + (i & 3);
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
    else v = (float)((double)job.bins[b] / bin_norm);
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
