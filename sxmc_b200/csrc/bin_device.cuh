// bin_device.cuh -- device code of the fast transform + binning pass, shared by bin_kernels.cu
// (one launch per step) and chain_kernels.cu (persistent chain kernel).
#pragma once
#include "common.cuh"

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

// x^3 rounded once from a double-double product (error ~2^-104 before the final rounding):
// equals the correctly rounded cube except when the exact value lies within ~2^-50 ulp of a
// rounding boundary.  glibc's pow (<0.53 ulp, the reference's `pow(x, 3)`) returns the correctly
// rounded value except within ~2^-15 ulp of a boundary, so the two agree on all but ~1e-5 of
// arguments; CUDA's own pow (up to 1 ulp off) does not.  Explicit FMAs: -fmad=false only stops
// contraction.  Overflow, inf, NaN and signed zero come out as pow(x, 3) gives them.
__device__ __forceinline__ double cube_rn(double x) {
  const double p = __dmul_rn(x, x);
  const double pe = __fma_rn(x, x, -p);        // x*x = p + pe exactly
  const double q = __dmul_rn(p, x);
  const double qe = __fma_rn(p, x, -q);        // p*x = q + qe exactly
  const double r = __dadd_rn(q, __dadd_rn(qe, __dmul_rn(pe, x)));
  return (r == r && fabs(q) <= 1.7976931348623157e308) ? r : q;   // inf - inf above: keep q
}

// p = sum_i par_i * pow(x, i) accumulated left to right (src/pdfz.cpp:310-314).
// pow(x,0)=1 and pow(x,1)=x are exact; pow(x,2.0)==x*x for glibc; i=3 is the
// once-rounded cube above (measured against the reference: tests/test_gpu_parity.py, P4 case).
__device__ __forceinline__ double syst_poly(const SystRt& s, double x) {
  double p = s.c0;
  if (s.npars > 1) {
    p = __dadd_rn(p, __dmul_rn(s.c1, x));
    if (s.npars > 2) {
      p = __dadd_rn(p, __dmul_rn(s.c2, __dmul_rn(x, x)));
      if (s.npars > 3) p = __dadd_rn(p, __dmul_rn(s.c3, cube_rn(x)));
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
// the CTAs; inside a CTA each job segment is dealt to the warps round robin (one contiguous range
// per warp measured 7 % slower on workload C: the CTA's copies in flight then read 16 distant
// addresses per column instead of neighbouring ones).
// Every warp runs its own S-stage pipeline: lane 0 issues one bulk async copy (TMA unit)
// per column per chunk into the warp's private shared-memory stages, completion is signalled
// on a per-stage mbarrier, and the warp consumes the stage with conflict-free LDS.128 --
// the loads in flight live in shared memory, not in registers, and no CTA-wide barrier is
// needed in steady state.
// Per-warp pipeline state that outlives one pass (the persistent chain kernel runs many).
struct FastCtx {
  unsigned my_stage0, my_bar0;   // shared-memory addresses of this warp's stages / mbarriers
  unsigned it;                   // chunks consumed so far: stage = it % S, parity = (it / S) & 1
  bool resident;                 // all of this warp's chunks fit in its S stages: load them once
  bool loaded;
};

template <int NS>
__device__ __forceinline__ void fast_ctx_init(FastCtx& cx, const FastParams& P, unsigned char* smem_raw,
                                              unsigned long long* s_bar) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned stage_bytes = NS * (unsigned)P.R * 512u;
  cx.my_stage0 = smem_u32(smem_raw + P.stage_off) + (unsigned)(warp * P.S) * stage_bytes;
  cx.my_bar0 = smem_u32(&s_bar[warp * FAST_MAX_STAGES]);
  cx.it = 0;
  cx.resident = false;
  cx.loaded = false;
  if (lane == 0) {
    for (int s = 0; s < P.S; ++s) mbar_init(cx.my_bar0 + 8u * s, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
}

// This CTA's chunk range [t0, t1) in the concatenated chunk sequence and the job t0 falls in.
__device__ __forceinline__ void fast_partition(const FastParams& P, unsigned long long& t0,
                                               unsigned long long& t1, int& j) {
  const unsigned long long* chunk_prefix = P.prefix;
  j = 0;
  if (P.aligned) {
    const unsigned b = blockIdx.x;
    while (j < P.njobs && P.cta_first[j + 1] <= b) ++j;
    if (j >= P.njobs) { t0 = t1 = 0; return; }
    const unsigned long long nj = chunk_prefix[j + 1] - chunk_prefix[j];
    const unsigned cj = P.cta_first[j + 1] - P.cta_first[j], k = b - P.cta_first[j];
    t0 = chunk_prefix[j] + nj * k / cj;
    t1 = chunk_prefix[j] + nj * (k + 1ull) / cj;
    return;
  }
  const unsigned long long total = chunk_prefix[P.njobs];
  t0 = total * blockIdx.x / gridDim.x;
  t1 = total * (blockIdx.x + 1ull) / gridDim.x;
  if (t0 >= t1) return;
  while (chunk_prefix[j + 1] <= t0) ++j;
}

// One pass over this CTA's share of all jobs.  hist_base != nullptr redirects the outputs to a
// concatenated buffer [njobs][total_nbins] + norms[njobs]; params_override != nullptr replaces
// every job's parameter buffer (may point to shared memory).
template <int NOBS, int NS, bool SMEM, int NOPS, unsigned CODES>
__device__ __forceinline__ void fast_bin_pass(const FastParams& P, FastCtx& cx, unsigned* s_hist,
                                              SystRt* s_syst, unsigned* s_norm, unsigned* hist_base,
                                              const double* params_override,
                                              unsigned long long* dbg = nullptr) {
#define BIN_DBG(slot)                                                                  \
  do {                                                                                 \
    if (dbg && threadIdx.x == 0) {                                                     \
      unsigned long long t_;                                                           \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                           \
      dbg[slot] = t_;                                                                  \
    }                                                                                  \
  } while (0)
  BIN_DBG(0);
  unsigned long long* const xd = P.xdbg ? P.xdbg + 32ull * blockIdx.x : nullptr;
#define BIN_XDBG(slot, who)                                                            \
  do {                                                                                 \
    if (xd && (who)) {                                                                 \
      unsigned long long t_;                                                           \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                           \
      xd[slot] = t_;                                                                   \
    }                                                                                  \
  } while (0)
  int xseg = 0;
  BIN_XDBG(1, threadIdx.x == 0);
  const Geo& geo = P.geo;
  const int R = P.R, S = P.S;
  const int njobs = P.njobs;
  const unsigned long long* chunk_prefix = P.prefix;   // constant bank: no global round trip
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const unsigned col_bytes = (unsigned)R * 512u;          // one column of one chunk
  const unsigned stage_bytes = NS * col_bytes;
  const unsigned my_stage0 = cx.my_stage0, my_bar0 = cx.my_bar0;
  const bool reuse = cx.resident && cx.loaded;            // the stages already hold this warp's chunks
  unsigned it = cx.resident ? 0u : cx.it;

  unsigned long long t0, t1;
  int j = 0;
  fast_partition(P, t0, t1, j);
  if (t0 >= t1) { pdl_wait(); pdl_trigger(); return; }
  bool first_segment = true;

  while (t0 < t1) {
    const FastJob& job = P.job[j];
    const unsigned long long jbeg = chunk_prefix[j], jend = chunk_prefix[j + 1];
    const unsigned long long tend = t1 < jend ? t1 : jend;
    const int nsyst = job.nsyst;

    const unsigned char* col[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) col[c] = reinterpret_cast<const unsigned char*>(job.col[c]);

    // this warp's chunks ca, ca + cstep, ... < cb of the job (chunk c starts at byte c * col_bytes)
    const unsigned long long n = tend - t0;
    const bool rr = P.cta_split != 1;   // chunks dealt to the warps round robin (1: contiguous, measurement)
    const unsigned NW = (unsigned)P.nwarps;   // == blockDim.x / 32
    const unsigned cstep = rr ? NW : 1u;
    const unsigned long long ca = rr ? (t0 - jbeg) + (unsigned)warp : (t0 - jbeg) + n * (unsigned)warp / NW;
    const unsigned long long cb = rr ? (tend - jbeg) : (t0 - jbeg) + n * (unsigned)(warp + 1) / NW;

    // (stage, parity) of chunk `it` kept as counters: it % S and it / S would be two integer
    // divisions per chunk in a kernel that is short of issue slots
    unsigned slot = it % (unsigned)S, par = (it / (unsigned)S) & 1u;
    // lane 0's issue cursor: the next chunk to request and its byte offset inside every column
    unsigned long long ci = ca, ci_off = ca * col_bytes;
    const unsigned long long step_bytes = (unsigned long long)cstep * col_bytes;
    auto issue = [&](unsigned sl) {   // lane 0 only
      const unsigned bar = my_bar0 + 8u * sl;
      const unsigned dst = my_stage0 + sl * stage_bytes;
      mbar_expect_tx(bar, stage_bytes);
#pragma unroll
      for (int k = 0; k < NS; ++k) bulk_g2s(dst + k * col_bytes, col[k] + ci_off, col_bytes, bar);
      ci += cstep;
      ci_off += step_bytes;
    };
    // first thing in the segment: get the copies going (everything they need is in the
    // constant bank); the parameter loads and the histogram clear overlap with them
    if (lane == 0 && !reuse) {
      unsigned sl = slot;
      for (unsigned i = 0; i < (unsigned)S && ci < cb; ++i) {
        issue(sl);
        if (++sl == (unsigned)S) sl = 0;
      }
    }

    // Programmatic dependent launch: everything above (barriers, the first bulk copies of the
    // sample columns) ran while the previous kernel of the stream -- the step before's lookup --
    // was still finishing; the parameters it proposes are read from here on.
    if (first_segment) { pdl_wait(); pdl_trigger(); first_segment = false; }
    if (threadIdx.x < nsyst) {   // p = 0 + par * pow(x, 0)  (src/pdfz.cpp:310-314, npars <= 1)
      SystRt& sr = s_syst[threadIdx.x];
      const double* pp = params_override ? params_override : job.params;
      const long long ps = params_override ? 1 : job.param_stride;
      const double par = job.npars[threadIdx.x] ? pp[(long long)job.pidx[threadIdx.x] * ps] : 0.0;
      sr.c0 = __dadd_rn(0.0, __dmul_rn(par, 1.0));
      sr.onep = __dadd_rn(1.0, sr.c0);
    }
    if (SMEM)
      for (int b = threadIdx.x; b < geo.total_nbins; b += blockDim.x) s_hist[b] = 0;
    __syncthreads();
    BIN_DBG(1);
    if (xseg < 4) BIN_XDBG(4 + 3 * xseg, threadIdx.x == 0);

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

    unsigned* job_bins = hist_base ? hist_base + (size_t)j * geo.total_nbins : job.bins;
    unsigned* job_norm = hist_base ? hist_base + (size_t)njobs * geo.total_nbins + j : job.norm;
    unsigned* hist = SMEM ? s_hist : job_bins;

    RunAcc acc;
    acc.norm = 0;
    acc.cur = -1;
    acc.cnt = 0;
#pragma unroll 1
    for (unsigned long long c = ca; c < cb; c += cstep, ++it) {
      mbar_wait(my_bar0 + 8u * slot, par);
      if (xseg == 0 && c == ca) BIN_XDBG(2, threadIdx.x == 0);
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
      if (lane == 0 && !reuse && ci < cb) issue(slot);
      if (++slot == (unsigned)S) { slot = 0; par ^= 1u; }
    }
    run_flush<SMEM>(acc, hist, geo.total_nbins);
    BIN_DBG(2);
    if (xseg < 4) BIN_XDBG(5 + 3 * xseg, threadIdx.x == 0);
    BIN_XDBG(16 + warp, lane == 0);

    // in-domain count: one global atomic per CTA and segment (all CTAs of a signal hit the
    // same address; per-warp atomics serialise in L2 and dominate small fits)
    const unsigned wnorm = __reduce_add_sync(0xffffffffu, acc.norm);
    if (lane == 0 && wnorm) atomicAdd(s_norm, wnorm);
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned cn = *s_norm;
      if (cn) atomicAdd(job_norm, cn);
      *s_norm = 0u;
    }
    if (SMEM) {
      for (int b = threadIdx.x; b < geo.total_nbins; b += blockDim.x) {
        const unsigned cnt = s_hist[b];
        if (cnt) atomicAdd(job_bins + b, cnt);
      }
    }
    __syncthreads();   // s_syst / s_hist / s_norm are rewritten by the next segment
    BIN_DBG(3);
    if (xseg < 4) BIN_XDBG(6 + 3 * xseg, threadIdx.x == 0);
    ++xseg;
    t0 = tend;
    ++j;
  }
  if (xd && threadIdx.x == 0) xd[3] = (unsigned long long)xseg;
  cx.it = it;
  cx.loaded = true;
}

// The same pass for launches whose histograms live in GLOBAL memory (more bins than fit beside the
// pipeline stages: workload D's 25 000, the 200 000-bin stress case).  Without a shared-memory
// histogram nothing ties a CTA to one signal at a time, so no warp ever waits for another: the
// concatenated chunk sequence of all signals is cut evenly over the CTAs, a CTA's range is dealt to
// its 16 warps round robin, and every warp streams its chunks through its own pipeline -- across
// signal boundaries without draining (flush the run, one atomic for the in-domain count, resolve
// the next signal's parameters by shuffle).  No CTA-wide barrier, no segment switch, no
// whole-CTAs-per-signal imbalance: the 1/8 shard of workload D bins in 59 us instead of 66
// (profiles/r02_bin_split_ab.txt).
template <int NOBS, int NS, int NOPS, unsigned CODES>
__device__ __forceinline__ void fast_bin_pass_warp(const FastParams& P, const FastCtx& cx) {
  const Geo& geo = P.geo;
  const int R = P.R, S = P.S;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned col_bytes = (unsigned)R * 512u, stage_bytes = NS * col_bytes;
  const unsigned my_stage0 = cx.my_stage0, my_bar0 = cx.my_bar0;
  unsigned long long* const xd = P.xdbg ? P.xdbg + 32ull * blockIdx.x : nullptr;
  const unsigned long long total = P.prefix[P.njobs];
  // the CTA takes a contiguous range and deals its chunks to the warps round robin: its 16 x S copies
  // in flight read neighbouring addresses of the same columns (measurement knob cta_split = 2: one
  // contiguous range per warp of the grid -- 3-5 % slower on workload D)
  unsigned long long t0, t1;
  unsigned tstep;
  if (P.cta_split == 0) {
    t0 = total * blockIdx.x / gridDim.x + (unsigned)warp;
    t1 = total * (blockIdx.x + 1ull) / gridDim.x;
    tstep = (unsigned)P.nwarps;
  } else {
    const unsigned long long nw = (unsigned long long)gridDim.x * P.nwarps, w = (unsigned long long)blockIdx.x * P.nwarps + warp;
    t0 = total * w / nw;
    t1 = total * (w + 1ull) / nw;
    tstep = 1;
  }
  if (t0 >= t1) { pdl_wait(); pdl_trigger(); return; }

  // issue cursor (lane 0 only): chunk index, end of its signal, the chunk's address in every column
  unsigned long long ti = t0, ji_end = 0;
  int ji = 0;
  const unsigned char* icol[NS];
  const unsigned long long step_bytes = (unsigned long long)tstep * col_bytes;
  auto issue = [&](unsigned slot) {
    if (ti >= ji_end) {
      while (P.prefix[ji + 1] <= ti) ++ji;
      ji_end = P.prefix[ji + 1];
      const unsigned long long off = (ti - P.prefix[ji]) * col_bytes;
#pragma unroll
      for (int k = 0; k < NS; ++k) icol[k] = reinterpret_cast<const unsigned char*>(P.job[ji].col[k]) + off;
    }
    const unsigned bar = my_bar0 + 8u * slot;
    const unsigned dst = my_stage0 + slot * stage_bytes;
    mbar_expect_tx(bar, stage_bytes);
#pragma unroll
    for (int k = 0; k < NS; ++k) {
      bulk_g2s(dst + k * col_bytes, icol[k], col_bytes, bar);
      icol[k] += step_bytes;
    }
    ti += tstep;
  };
  if (lane == 0)
    for (int i = 0; i < S && ti < t1; ++i) issue((unsigned)i);
  // the sample columns never depend on the previous kernel of the stream; the parameters do
  pdl_wait();
  pdl_trigger();

  int jc = -1;
  unsigned long long jc_end = 0;
  int op_code[FAST_MAX_OPS];
  double op_c0[FAST_MAX_OPS], op_onep[FAST_MAX_OPS];
  int nsyst = 0;
  unsigned* gbins = nullptr;
  RunAcc acc;
  acc.norm = 0; acc.cur = -1; acc.cnt = 0;
  auto close_job = [&]() {
    run_flush<false>(acc, gbins, geo.total_nbins);
    const unsigned wnorm = __reduce_add_sync(0xffffffffu, acc.norm);
    if (lane == 0 && wnorm) atomicAdd(P.job[jc].norm, wnorm);
    acc.norm = 0; acc.cur = -1; acc.cnt = 0;
  };
  unsigned it = 0, slot = 0, par = 0;      // (stage, parity) as counters, not it % S and it / S
#pragma unroll 1
  for (unsigned long long tc = t0; tc < t1; tc += tstep, ++it) {
    if (tc >= jc_end) {                                // this warp enters a (new) signal
      if (jc >= 0) close_job();
      jc = jc < 0 ? 0 : jc + 1;
      while (P.prefix[jc + 1] <= tc) ++jc;
      jc_end = P.prefix[jc + 1];
      const FastJob& job = P.job[jc];
      nsyst = job.nsyst;
      // p = 0 + par * pow(x, 0) (src/pdfz.cpp:310-314): lane k resolves systematic k, then broadcast
      double c0 = 0.0;
      if (lane < nsyst && lane < FAST_MAX_OPS) {
        const double par = job.npars[lane] ? job.params[(long long)job.pidx[lane] * job.param_stride] : 0.0;
        c0 = __dadd_rn(0.0, __dmul_rn(par, 1.0));
      }
#pragma unroll
      for (int k = 0; k < FAST_MAX_OPS; ++k) {
        op_code[k] = job.code[k];
        op_c0[k] = __shfl_sync(0xffffffffu, c0, k);
        op_onep[k] = __dadd_rn(1.0, op_c0[k]);
      }
      gbins = job.bins;
    }
    mbar_wait(my_bar0 + 8u * slot, par);
    if (xd && it == 0 && threadIdx.x == 0) {
      unsigned long long t_;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
      xd[2] = t_;
    }
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
      run_push4<false>(acc, bq, gbins, geo.total_nbins);
    }
    __syncwarp();                       // every lane has read the stage before it is refilled
    if (lane == 0 && ti < t1) issue(slot);
    if (++slot == (unsigned)S) { slot = 0; par ^= 1u; }
  }
  close_job();
  if (xd && lane == 0) {
    unsigned long long t_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
    xd[16 + warp] = t_;
    if (warp == 0) xd[3] = 1ull;
  }
}

template <int NOBS, int NS, bool SMEM, int NOPS, unsigned CODES>
__global__ void __launch_bounds__(FAST_THREADS, 1)
bin_fast_kernel(const __grid_constant__ FastParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ SystRt s_syst[SXMC_MAX_SYST];
  __shared__ __align__(8) unsigned long long s_bar[FAST_WARPS * FAST_MAX_STAGES];
  __shared__ unsigned s_norm;
  if (threadIdx.x == 0) s_norm = 0u;
  if (P.xdbg && threadIdx.x == 0) {
    unsigned long long t_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
    P.xdbg[32ull * blockIdx.x] = t_;
    unsigned smid_;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid_));
    P.xdbg[32ull * blockIdx.x + 11] = (unsigned long long)smid_ + 1ull;
  }
  FastCtx cx;
  fast_ctx_init<NS>(cx, P, smem_raw, s_bar);
  if (SMEM || (P.cta_split & 1)) {   // (1, 3: measurement knobs)
    fast_bin_pass<NOBS, NS, SMEM, NOPS, CODES>(P, cx, reinterpret_cast<unsigned*>(smem_raw), s_syst,
                                               &s_norm, nullptr, nullptr);
  } else {
    fast_bin_pass_warp<NOBS, NS, NOPS, CODES>(P, cx);
    __syncthreads();                    // (profiling stamp below: the CTA's last warp)
  }
  if (P.xdbg && threadIdx.x == 0) {
    unsigned long long t_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
    P.xdbg[32ull * blockIdx.x + 31] = t_;
  }
}

// the programs kernels are instantiated for (bin_kernels.cu and chain_kernels.cu)
#define OPC(type, obs, ext) ((unsigned)((type) | ((obs) << 2) | ((ext) << 5)))
#define SXMC_FAST_TABLE(FE)                                                                        \
  /* run-time dispatch of the systematics (any program of <= FAST_MAX_OPS single-coefficient ops) */ \
  FE(1, 1, -1, 0u), FE(1, 2, -1, 0u), FE(1, 3, -1, 0u), FE(2, 2, -1, 0u), FE(2, 3, -1, 0u),        \
  FE(2, 4, -1, 0u), FE(3, 3, -1, 0u), FE(3, 4, -1, 0u), FE(3, 5, -1, 0u),                          \
  /* no systematics */                                                                             \
  FE(1, 1, 0, 0u), FE(2, 2, 0, 0u), FE(3, 3, 0, 0u),                                               \
  /* workload B / bench_pdfz: shift(energy); scale(energy) */                                      \
  FE(1, 1, 1, OPC(SXMC_SYST_SHIFT, 0, 0)), FE(1, 1, 1, OPC(SXMC_SYST_SCALE, 0, 0)),                \
  /* workload A: scale(energy), resolution_scale(energy; mc_energy) */                             \
  FE(1, 2, 2, OPC(SXMC_SYST_SCALE, 0, 0) | OPC(SXMC_SYST_RESOLUTION_SCALE, 0, 1) << 8),            \
  /* workload C: the same on 2-D energy x radius */                                                \
  FE(2, 3, 2, OPC(SXMC_SYST_SCALE, 0, 0) | OPC(SXMC_SYST_RESOLUTION_SCALE, 0, 2) << 8),            \
  /* workload D: scale(energy), resolution_scale(energy; mc_energy), shift(radius) on 3-D */       \
  FE(3, 4, 3, OPC(SXMC_SYST_SCALE, 0, 0) | OPC(SXMC_SYST_RESOLUTION_SCALE, 0, 3) << 8 |            \
                  OPC(SXMC_SYST_SHIFT, 1, 0) << 16)


}  // namespace sxmc
