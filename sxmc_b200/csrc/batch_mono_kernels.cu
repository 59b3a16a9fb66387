// batch_mono_kernels.cu -- the chain batch kernel for MONOTONE fits: one observable whose
// systematics are all SHIFT / SCALE / CTSCALE (workloads B and E of SURVEY 8: the ensemble
// configuration of BASELINE.json).
//
// For such a fit a sample x reaches its bin through monotone roundings only,
//     x -> x' = syst(x) -> (int)((x' - lo) * scale)            (src/pdfz.cpp:306-331, :386-398)
// so over samples SORTED by x the bin index is a monotone step function, and a histogram is
// fully described by the positions at which the index changes.  Every CTA keeps a contiguous
// range of the sorted samples in shared memory (plain float array, loaded once per launch).
// For one chain, the two ends of the range give the ranks it spans, and every bin edge in
// between is found by one thread: a float guess, then a few evaluations of the reference's
// exact arithmetic on actual samples -- per EDGE instead of per SAMPLE, so the counts are
// bit-identical to the sample-by-sample pass.  The (chain, edge) pairs of all chains form one
// flat list shared by the CTA's 1024 threads.
// The host cuts the ranges so that every CTA sees about the same number of edges (sparse tails
// get few samples, the dense core is limited by shared-memory capacity) -- plan_chain_batch_mono.
//
// Preconditions (checked on the host): samples sorted by the observable, no NaN / -inf among
// them.  Fits that do not qualify run chain_batch_kernel (batch_kernels.cu).
#include "common.cuh"
#include "kernels.h"
#include "bin_device.cuh"
#include "nll_device.cuh"
#include "grid_sync.cuh"
// Threads per CTA.  The bin phase is a chain of dependent shared-memory loads and FP64
// evaluations per warp, so the warps in flight set its speed: 32 warps (64 registers each, no spills)
// against the 16 of the streaming kernels (E: 184 -> 210 -> 236 experiments/s at 16 / 24 / 32 warps).
#ifndef SXMC_MONO_THREADS
#define SXMC_MONO_THREADS 1024
#endif
#define BATCH_THREADS SXMC_MONO_THREADS
#include "batch_device.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

namespace sxmc {

constexpr int MONO_THREADS = SXMC_MONO_THREADS;
constexpr int MONO_WARPS = MONO_THREADS / 32;

// code of a transformed sample: -1 below the domain (or NaN), -2 at/above the upper edge, else
// the bin index (which can be total_nbins for x' within an ulp of the edge: counted in the norm
// but not in a bin, as the sample-by-sample kernels do)
template <int NOPS, unsigned CODES>
__device__ __forceinline__ int mono_code(float xf, const int (&op_code)[FAST_MAX_OPS],
                                         const double (&op_c0)[FAST_MAX_OPS],
                                         const double (&op_onep)[FAST_MAX_OPS], int nsyst, const Geo& geo) {
  double f[1][4];
  f[0][0] = f[0][1] = f[0][2] = f[0][3] = (double)xf;
  if constexpr (NOPS < 0) {
#pragma unroll
    for (int k = 0; k < FAST_MAX_OPS; ++k)
      if (k < nsyst) apply_op<1>(f, op_code[k], op_c0[k], op_onep[k]);
  } else {
    if constexpr (NOPS > 0) apply_static<1, (CODES >> 0) & 255>(f, op_c0[0], op_onep[0]);
    if constexpr (NOPS > 1) apply_static<1, (CODES >> 8) & 255>(f, op_c0[1], op_onep[1]);
    if constexpr (NOPS > 2) apply_static<1, (CODES >> 16) & 255>(f, op_c0[2], op_onep[2]);
    if constexpr (NOPS > 3) apply_static<1, (CODES >> 24) & 255>(f, op_c0[3], op_onep[3]);
  }
  const double x = f[0][0];
  if (!(x >= geo.lo[0])) return -1;
  if (!(x < geo.hi[0])) return -2;
  return trunc_nonneg(__dmul_rn(__dadd_rn(x, -geo.lo[0]), geo.sc[0]));
}

// position of a code along the x' axis: below < bin 0 < ... < bin tnb < at/above the edge
__device__ __forceinline__ int mono_rank(int code, int tnb) {
  return code == -1 ? 0 : (code == -2 ? tnb + 2 : code + 1);
}

// rank of an untransformed value (guesses only; never decides a count)
// `fine` sub-divisions per bin: rank 0 below the domain, then fine ranks per bin, then above.
constexpr int MONO_TAB = 512;
__device__ __forceinline__ int nominal_rank(float xf, const Geo& geo, int tnb, int fine) {
  const double x = (double)xf;
  if (!(x >= geo.lo[0])) return 0;
  if (!(x < geo.hi[0])) return (tnb + 1) * fine + 1;
  return min((int)((x - geo.lo[0]) * (geo.sc[0] * fine)), (tnb + 1) * fine - 1) + 1;
}

// Per-launch scratch in front of the samples in dynamic shared memory.
struct MonoShared {
  double par[BATCH_MAX_CHAINS][FAST_MAX_OPS];   // every chain's systematic coefficients
  double inv_a[BATCH_MAX_CHAINS];               // the chain's program as an affine map x' = a x + b:
  double aff_b[BATCH_MAX_CHAINS];               //   1 / a and b (guesses only; never decide a count)
  int rf[BATCH_MAX_CHAINS];                     // rank of the range's first / last sample under the chain
  int rl[BATCH_MAX_CHAINS];
  int off[BATCH_MAX_CHAINS + 1];                // exclusive scan of the chains' edge counts
  int wsum[8];
  int tab[MONO_TAB];                            // nominal edge table
};
constexpr size_t MONO_SHARED_BYTES = (sizeof(MonoShared) + 127) & ~(size_t)127;

template <int NOPS, unsigned CODES>
__global__ void __launch_bounds__(MONO_THREADS, 1)
chain_batch_mono_kernel(const __grid_constant__ MonoParams M, const __grid_constant__ BatchParams Q) {
  extern __shared__ __align__(128) unsigned char s_dyn[];
  MonoShared& X = *reinterpret_cast<MonoShared*>(s_dyn);
  float* const s_x = reinterpret_cast<float*>(s_dyn + MONO_SHARED_BYTES);   // this CTA's sorted samples
  __shared__ BatchShared S;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const unsigned G = gridDim.x;
  const int ns = Q.ns, nsources = Q.nsources, nch = Q.nchains;
  const int tnb = M.geo.total_nbins;
  const Geo& geo = M.geo;
  unsigned* const bar = Q.sync + 0;
  unsigned* const abort_flag = Q.sync + 3;
  unsigned bar_count = 0;

  const int n = (int)M.n[blockIdx.x];
  const int job_j = M.jobof[blockIdx.x];
  const MonoJob& job = M.job[job_j];
  {
    const unsigned long long s0 = M.s0[blockIdx.x];
    for (int p = tid; p < n; p += MONO_THREADS) s_x[p] = __ldg(job.col + tiled_offset(s0 + p));
  }
  const int mine = blockIdx.x < (unsigned)nch ? (int)blockIdx.x : -1;
  batch_owner_init(S, Q, mine, tid);
  __syncthreads();
  // Nominal edge table: tab[k] = first position whose rank WITHOUT systematics reaches
  // r0_nom + k, on a grid of up to 16 sub-divisions per bin.  A lane's threshold falls into one
  // cell, so its float bisection starts from that cell's samples instead of the whole range:
  // fewer steps, and the lanes no longer probe positions a power-of-two stride apart (for n near
  // k * 4096 those shared one bank: the CTA with n = 28671 took twice as long over its guesses
  // as its neighbours).
  int fine = 16, r0_nom = 0, ntab = 0;
  for (; fine >= 1 && n > 0; fine >>= 1) {          // the finest table that fits
    r0_nom = nominal_rank(s_x[0], geo, tnb, fine);
    const int span = nominal_rank(s_x[n - 1], geo, tnb, fine) - r0_nom + 2;
    if (span <= MONO_TAB) { ntab = span; break; }
  }
  if (fine < 1) fine = 1;                             // ntab == 0: span too wide, no table
  for (int k = tid; k < ntab; k += MONO_THREADS) {
    int a0 = 0, a1 = n;
    while (a0 < a1) {
      const int mid = (a0 + a1) >> 1;
      if (nominal_rank(s_x[mid], geo, tnb, fine) < r0_nom + k) a0 = mid + 1; else a1 = mid;
    }
    X.tab[k] = a0;
  }
  __syncthreads();

  const int nsyst = job.nsyst;
  const double inv_sc = 1.0 / geo.sc[0];
  int op_code[FAST_MAX_OPS];
#pragma unroll
  for (int k = 0; k < FAST_MAX_OPS; ++k) op_code[k] = k < nsyst ? job.code[k] : 0;
  const unsigned total_super = Q.nsteps + (Q.do_init ? 1u : 0u);

  // chain e's program constants, as bin_samples forms them (src/pdfz.cpp:310-314)
  auto load_ops = [&](int e, double (&op_c0)[FAST_MAX_OPS], double (&op_onep)[FAST_MAX_OPS]) {
#pragma unroll
    for (int k = 0; k < FAST_MAX_OPS; ++k) {
      op_c0[k] = 0.0; op_onep[k] = 1.0;
      if (k < nsyst) {
        op_c0[k] = __dadd_rn(0.0, __dmul_rn(X.par[e][k], 1.0));
        op_onep[k] = __dadd_rn(1.0, op_c0[k]);
      }
    }
  };

  // the owned chain's event multiplicities do not change during a launch: bin `tid`'s in a register
  const bool have_m0 = mine >= 0 && Q.mult != nullptr && tid < tnb;
  const unsigned m0 = have_m0 ? __ldg(Q.mult + (size_t)mine * tnb + tid) : 0u;

#define BKM_STAMP(slot)                                                              \
  do {                                                                               \
    if (Q.stamps && tid == 0 && s == 20u) Q.stamps[blockIdx.x * 8 + (slot)] = globaltimer_ns(); \
  } while (0)
  for (unsigned s = 0; s < total_super; ++s) {
    const int mode = (Q.do_init && s == 0) ? NLL_MODE_INIT : NLL_MODE_STEP;
    const bool bin_now = Q.rebin || mode == NLL_MODE_INIT;
    BKM_STAMP(0);

    // ---- bin ----------------------------------------------------------------------------------
    // A chain's histogram over this range is the set of positions where its rank steps.  The
    // (chain, edge) pairs of ALL chains form one flat list that the 1024 threads share, one
    // pair per thread per pass -- no lane idles because a chain has fewer than 32 edges here.
    // Edge at position p between the run of rank r and the run of rank r': +p to r, -p to r'
    // (unsigned wrap-around: the sums are exact once every edge has been added).
    if (bin_now) {
      if (n > 0) {
        // (1) all chains' coefficients in one round of loads
        for (int q = tid; q < nch * FAST_MAX_OPS; q += MONO_THREADS) {
          const int e = q / FAST_MAX_OPS, k = q % FAST_MAX_OPS;
          double par = 0.0;
          if (k < nsyst && job.npars[k]) par = __ldcg(&Q.st[e].proposed[nsources + job.pidx[k]]);
          X.par[e][k] = par;
        }
        __syncthreads();
        // (2) every chain's rank at the two ends of the range, and its affine map
        for (int q = tid; q < 2 * nch; q += MONO_THREADS) {
          const int e = q >> 1;
          double op_c0[FAST_MAX_OPS], op_onep[FAST_MAX_OPS];
          load_ops(e, op_c0, op_onep);
          const int r = mono_rank(mono_code<NOPS, CODES>(s_x[(q & 1) ? n - 1 : 0], op_code, op_c0, op_onep, nsyst, geo), tnb);
          if (q & 1) {
            X.rl[e] = r;
          } else {
            X.rf[e] = r;
            double aff_a = 1.0, aff_b = 0.0;
#pragma unroll
            for (int k = 0; k < FAST_MAX_OPS; ++k) {
              if (k < nsyst) {
                const int type = op_code[k] & 15;
                if (type == SXMC_SYST_SHIFT) aff_b += op_c0[k];
                else if (type == SXMC_SYST_SCALE) { aff_a *= op_onep[k]; aff_b *= op_onep[k]; }
                else if (type == SXMC_SYST_CTSCALE) { aff_a *= op_onep[k]; aff_b = 1.0 + (aff_b - 1.0) * op_onep[k]; }
              }
            }
            X.inv_a[e] = 1.0 / aff_a;
            X.aff_b[e] = aff_b;
          }
        }
        __syncthreads();
        // (3) edge counts, their exclusive scan, and the run that is still open at the end of
        // the range (rank rl, ends at n)
        int ne = 0, incl = 0;
        if (tid < 160) {
          if (tid < nch) {
            const int rl = X.rl[tid];
            ne = abs(rl - X.rf[tid]);
            if (rl >= 1 && rl <= tnb + 1) {
              unsigned* const hist = Q.hist + (size_t)tid * Q.hist_words;
              atomicAdd(hist + (size_t)ns * tnb + job_j, (unsigned)n);
              if (rl <= tnb) atomicAdd(hist + (size_t)job_j * tnb + (rl - 1), (unsigned)n);
            }
          }
          incl = ne;
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
          }
          if (lane == 31) X.wsum[warp] = incl;
        }
        __syncthreads();
        if (tid < 160) {
          int base = 0;
          for (int w = 0; w < warp; ++w) base += X.wsum[w];
          if (tid <= nch) X.off[tid] = base + incl - ne;
        }
        __syncthreads();
        const int total_edges = X.off[nch];
        // (4) one (chain, edge) pair per thread per pass
        for (int g = tid; g < total_edges; g += MONO_THREADS) {
          int e = 0;
          {
            int a = 0, b = nch - 1;                  // largest e with off[e] <= g
            while (a < b) {
              const int mid = (a + b + 1) >> 1;
              if (X.off[mid] <= g) a = mid; else b = mid - 1;
            }
            e = a;
          }
          const int k = g - X.off[e];
          double op_c0[FAST_MAX_OPS], op_onep[FAST_MAX_OPS];
          load_ops(e, op_c0, op_onep);
          auto rank_at = [&](int p) {
            return mono_rank(mono_code<NOPS, CODES>(s_x[p], op_code, op_c0, op_onep, nsyst, geo), tnb);
          };
          const int r_first = X.rf[e];
          const int dir = X.rl[e] >= r_first ? 1 : -1;
          const int t = r_first + dir * (k + 1);
          // first position whose rank has reached t (exists: the last sample has rank rl).
          // A guess comes from inverting the affine map in plain floating point and a float-only
          // bisection of the samples; the answer is then settled with the exact arithmetic:
          // gallop away from the guess, bisect what is left (usually two evaluations in all).
          int lo = 0, hi = n - 1;
          auto pred = [&](int p) { return (rank_at(p) - t) * dir >= 0; };
          const int rb = dir > 0 ? t : t + 1;          // rank whose lower boundary is crossed
          const double bv = rb <= tnb + 1 ? geo.lo[0] + (double)(rb - 1) * inv_sc : geo.hi[0];
          const float thr = (float)((bv - X.aff_b[e]) * X.inv_a[e]);
          int p0 = n >> 1;
          if (thr == thr) {
            int a0 = 0, a1 = n;                        // first sample >= thr (floats only)
            if (ntab) {
              const int kk = min(max(nominal_rank(thr, geo, tnb, fine) - r0_nom, 0), ntab - 1);
              a0 = X.tab[kk];
              a1 = kk + 1 < ntab ? X.tab[kk + 1] : n;
            }
            while (a0 < a1) {
              const int mid = (a0 + a1) >> 1;
              if (s_x[mid] < thr) a0 = mid + 1; else a1 = mid;
            }
            p0 = min(a0, n - 1);
          }
          // the guess and its two neighbours, evaluated together (the answer is almost always
          // one of them)
          const int pm = max(p0 - 1, 0), pp = min(p0 + 1, n - 1);
          const bool f_m = pred(pm), f_0 = pred(p0), f_p = pred(pp);
          if (f_0) {
            hi = p0;
            if (p0 == 0) {
              lo = 0;
            } else if (!f_m) {
              lo = p0;
            } else {
              hi = pm;
              int step = 1;
              while (hi > 0) {
                const int q = max(hi - step, 0);
                if (pred(q)) { hi = q; step <<= 1; } else { lo = q + 1; break; }
              }
              if (hi == 0) lo = 0;
            }
          } else if (f_p) {
            lo = hi = pp;
          } else {
            lo = pp + 1;
            int step = 1;
            while (lo < hi) {
              const int q = min(lo + step - 1, hi);
              if (pred(q)) { hi = q; break; }
              lo = q + 1;
              step <<= 1;
            }
          }
          while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (pred(mid)) hi = mid; else lo = mid + 1;
          }
          // the run of rank t - dir ends at lo, the run of rank t starts there
          unsigned* const hist = Q.hist + (size_t)e * Q.hist_words;
          const unsigned pos = (unsigned)lo;
          const int ra = t - dir;
          const bool in_a = ra >= 1 && ra <= tnb + 1, in_t = t >= 1 && t <= tnb + 1;
          if (in_a && ra <= tnb) atomicAdd(hist + (size_t)job_j * tnb + (ra - 1), pos);
          if (in_t && t <= tnb) atomicAdd(hist + (size_t)job_j * tnb + (t - 1), 0u - pos);
          if (in_a != in_t) atomicAdd(hist + (size_t)ns * tnb + job_j, in_a ? pos : 0u - pos);
        }
      }
      BKM_STAMP(1);
      cta_arrive(bar);
      ++bar_count;
      if (!cta_wait_ge(bar, bar_count * G, abort_flag, &S.ok)) break;
      BKM_STAMP(2);
    }

    // ---- lookup + event likelihood + tail of the chain this CTA owns ---------------------------
    if (mine >= 0)
      batch_lookup_tail(S, Q, mine, mode, tnb, tid, Q.rebin && s + 1 < total_super,
                        (Q.stamps && s == 20u) ? Q.stamps + blockIdx.x * 8 : nullptr, have_m0, m0);
    BKM_STAMP(4);
    if (Q.rebin && s + 1 < total_super) {      // proposals published, histograms cleared
      cta_arrive(bar);
      ++bar_count;
      if (!cta_wait_ge(bar, bar_count * G, abort_flag, &S.ok)) break;
    }
    BKM_STAMP(5);
  }
}

// ---------------------------------------------------------------------------------------------
typedef void (*MonoKernel)(const MonoParams, const BatchParams);
struct MonoEntry {
  int nops;
  unsigned codes;
  MonoKernel k;
};
static const MonoEntry g_mono_table[] = {
    {-1, 0u, chain_batch_mono_kernel<-1, 0u>},
    {0, 0u, chain_batch_mono_kernel<0, 0u>},
    {1, OPC(SXMC_SYST_SHIFT, 0, 0), chain_batch_mono_kernel<1, OPC(SXMC_SYST_SHIFT, 0, 0)>},
    {1, OPC(SXMC_SYST_SCALE, 0, 0), chain_batch_mono_kernel<1, OPC(SXMC_SYST_SCALE, 0, 0)>},
};

constexpr size_t MONO_STATIC_RESERVE = 56 * 1024;   // BatchShared (static) + MonoShared (dynamic) + slack
constexpr size_t MONO_CAP_SAMPLES = (227 * 1024 - MONO_STATIC_RESERVE) / sizeof(float);

// nominal (no systematics) rank of a sample, on the host: only used to balance the ranges
static int host_rank(float xf, const BinLaunch& L) {
  const double x = (double)xf;
  if (!(x >= L.lower[0])) return 0;
  if (!(x < L.upper[0])) return L.max_total_nbins + 2;
  return (int)((x - L.lower[0]) * L.scale[0]) + 1;
}

// Cut every job's sorted samples into contiguous ranges, one per CTA, so that no range holds
// more than `cap` samples and the largest number of bin edges inside a range is as small as
// possible.  nsamples[j] = real (unpadded) samples of job j.
int plan_chain_batch_mono(const BinLaunch& L, const unsigned long long* nsamples, int nchains, MonoPlan* plan) {
  if (!L.fast || L.nobs != 1 || L.nslots != 1 || L.njobs > FAST_MAX_JOBS || nchains < 1 ||
      nchains > BATCH_MAX_CHAINS) {
    set_error("monotone chain batch: fit shape not supported");
    return SXMC_ENOTSMALL;
  }
  int dev = 0, sms = 0;
  SXMC_CUDA(cudaGetDevice(&dev));
  SXMC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  if (sms > MONO_MAX_CTAS) sms = MONO_MAX_CTAS;
  if (nchains > sms) { set_error("monotone chain batch: more chains than SMs"); return SXMC_EINVAL; }
  const int nj = L.njobs;
  const int nranks = L.max_total_nbins + 3;
  // sorted samples of every job on the host, and the first position of every nominal rank
  std::vector<std::vector<unsigned>> edge(nj);
  std::vector<int> span(nj, 1);
  std::vector<unsigned long long> need(nj);
  unsigned long long need_total = 0, span_total = 0;
  for (int j = 0; j < nj; ++j) {
    const BinJob& job = L.jobs_host[j];
    const unsigned long long n = nsamples[j];
    if (n >= (1ull << 31)) { set_error("monotone chain batch: too many samples"); return SXMC_ENOTSMALL; }
    std::vector<float> tiled((size_t)job.nwtiles * WT_SAMPLES);
    if (!tiled.empty())
      SXMC_CUDA(cudaMemcpy(tiled.data(), job.cols[job.slot_field[0]], sizeof(float) * tiled.size(),
                           cudaMemcpyDeviceToHost));
    std::vector<int> rk((size_t)n);
    for (unsigned long long s = 0; s < n; ++s) rk[(size_t)s] = host_rank(tiled[(size_t)tiled_offset(s)], L);
    if (!std::is_sorted(rk.begin(), rk.end())) {
      set_error("monotone chain batch: samples are not sorted by the observable");
      return SXMC_ESTATE;
    }
    edge[j].resize(nranks + 1);
    for (int t = 0; t <= nranks; ++t)
      edge[j][t] = (unsigned)(std::lower_bound(rk.begin(), rk.end(), t) - rk.begin());
    if (n) span[j] = std::max(1, rk[(size_t)n - 1] - rk[0] + 1);
    need[j] = (n + MONO_CAP_SAMPLES - 1) / MONO_CAP_SAMPLES;
    need_total += need[j];
    span_total += n ? span[j] : 0;
  }
  if (need_total > (unsigned long long)sms) {
    set_error("chain batch: the samples do not fit in shared memory (not a small fit)");
    return SXMC_ENOTSMALL;
  }
  // greedy cut of job j for an edge budget E: the ranges needed (optionally listed as pos, count)
  auto cut = [&](int j, int E, std::vector<unsigned>* out) {
    const unsigned n = (unsigned)nsamples[j];
    int count = 0;
    unsigned pos = 0;
    while (pos < n) {
      // rank at pos = largest t with edge[t] <= pos
      const int r = (int)(std::upper_bound(edge[j].begin(), edge[j].end(), pos) - edge[j].begin()) - 1;
      const int lim = std::min(nranks, r + E);
      unsigned next = edge[j][lim];
      if (next <= pos) next = n;
      next = std::min<unsigned>(next, pos + (unsigned)MONO_CAP_SAMPLES);
      next = std::min(next, n);
      if (out) { out->push_back(pos); out->push_back(next - pos); }
      pos = next;
      ++count;
    }
    return count;
  };
  // smallest edge budget with which job j fits into c CTAs (nranks + 1: it does not fit at all)
  auto budget = [&](int j, int c) {
    if (!nsamples[j]) return 0;
    if (c < (int)need[j]) return nranks + 1;
    int lo = 1, hi = nranks;
    while (lo < hi) {
      const int mid = (lo + hi) / 2;
      if (cut(j, mid, nullptr) <= c) hi = mid; else lo = mid + 1;
    }
    return lo;
  };
  // CTAs per job: its capacity minimum, the rest in proportion to the bin edges it spans; then
  // CTAs move from the job with the smallest edge budget to the one with the largest while that
  // lowers the largest (capacity-limited ranges in a dense core hold few edges, so a job's
  // budget is not simply span / CTAs: the super-step waits for the slowest CTA)
  std::vector<int> ctas(nj);
  int used = 0;
  for (int j = 0; j < nj; ++j) { ctas[j] = (int)need[j]; used += ctas[j]; }
  while (used < sms) {
    int best = -1;
    double score = 0;
    for (int j = 0; j < nj; ++j) {
      if (!nsamples[j]) continue;
      const double sc = (double)span[j] / ctas[j];
      if (best < 0 || sc > score) { best = j; score = sc; }
    }
    if (best < 0) break;
    ctas[best] += 1;
    used += 1;
  }
  std::vector<int> bud(nj);
  for (int j = 0; j < nj; ++j) bud[j] = budget(j, ctas[j]);
  for (int round = 0; round < 4 * sms; ++round) {
    int jmax = -1, jmin = -1;
    for (int j = 0; j < nj; ++j) {
      if (!nsamples[j]) continue;
      if (jmax < 0 || bud[j] > bud[jmax]) jmax = j;
    }
    int best_after = jmax < 0 ? 0 : bud[jmax];
    int give_bud = 0, take_bud = 0;
    if (jmax < 0) break;
    const int after_take = budget(jmax, ctas[jmax] + 1);
    for (int j = 0; j < nj; ++j) {
      if (j == jmax || !nsamples[j] || ctas[j] <= (int)need[j] || ctas[j] <= 1) continue;
      const int after_give = budget(j, ctas[j] - 1);
      int worst = std::max(after_take, after_give);
      for (int q = 0; q < nj; ++q) if (q != j && q != jmax && nsamples[q]) worst = std::max(worst, bud[q]);
      if (worst < best_after) { best_after = worst; jmin = j; give_bud = after_give; take_bud = after_take; }
    }
    if (jmin < 0) break;
    ctas[jmin] -= 1; bud[jmin] = give_bud;
    ctas[jmax] += 1; bud[jmax] = take_bud;
  }
  MonoParams& M = plan->params;
  memset(&M, 0, sizeof M);
  for (int i = 0; i < 3; ++i) { M.geo.lo[i] = L.lower[i]; M.geo.hi[i] = L.upper[i]; M.geo.sc[i] = L.scale[i]; M.geo.st[i] = L.stride[i]; }
  M.geo.total_nbins = L.max_total_nbins;
  M.njobs = nj;
  unsigned max_n = 0;
  int cta = 0, max_edges = 0;
  for (int j = 0; j < nj; ++j) {
    const BinJob& src = L.jobs_host[j];
    MonoJob& d = M.job[j];
    d.col = src.cols[src.slot_field[0]];
    d.nsyst = src.nsyst;
    for (int k = 0; k < src.nsyst && k < FAST_MAX_OPS; ++k) {
      d.code[k] = (unsigned short)(src.syst[k].type | (src.syst[k].obs_slot << 4) | (src.syst[k].extra_slot << 8));
      d.pidx[k] = src.syst[k].pidx[0];
      d.npars[k] = (unsigned char)(src.syst[k].npars > 0);
    }
    const unsigned n = (unsigned)nsamples[j];
    if (!n) continue;
    const int lo = bud[j];
    std::vector<unsigned> ranges;
    const int got = cut(j, lo, &ranges);
    if (got > ctas[j]) { set_error("monotone chain batch: partition failed"); return SXMC_ESTATE; }
    max_edges = std::max(max_edges, lo);
    for (int q = 0; q < ctas[j]; ++q, ++cta) {
      M.jobof[cta] = (unsigned char)j;
      if (q < got) { M.s0[cta] = ranges[2 * q]; M.n[cta] = ranges[2 * q + 1]; max_n = std::max(max_n, M.n[cta]); }
    }
  }
  plan->grid = sms;
  plan->smem_bytes = MONO_SHARED_BYTES + (((size_t)max_n * sizeof(float) + 127) & ~(size_t)127);
  plan->max_edges = max_edges;
  plan->prog_nops = L.jit ? -1 : L.prog_nops;      // NVRTC-compiled programs exist for bin_fast_kernel only
  plan->prog_codes = L.jit ? 0u : L.prog_codes;
  return SXMC_OK;
}

int launch_chain_batch_mono(const MonoPlan& plan, const BatchParams& Q0, cudaStream_t stream) {
  MonoKernel fk = nullptr;
  for (const MonoEntry& e : g_mono_table)
    if (e.nops == plan.prog_nops && (plan.prog_nops <= 0 || e.codes == plan.prog_codes)) fk = e.k;
  if (!fk) fk = g_mono_table[0].k;      // run-time dispatch of the systematics
  const size_t smem = std::max<size_t>(plan.smem_bytes, 128);
  SXMC_CUDA(cudaFuncSetAttribute((const void*)fk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  SXMC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)fk, MONO_THREADS, smem));
  if (occ < 1) { set_error("monotone chain batch kernel does not fit on an SM"); return SXMC_ENOTSMALL; }
  BatchParams Q = Q0;
  void* args[2] = {(void*)&plan.params, (void*)&Q};
  SXMC_CUDA(cudaLaunchCooperativeKernel((const void*)fk, dim3((unsigned)plan.grid), dim3(MONO_THREADS), args, smem,
                                        stream));
  return SXMC_OK;
}

}  // namespace sxmc
