// ensemble_kernels.cu -- the two device pieces either side of a chain in the ensemble loop
// (reference src/sxmc.cpp:62-134):
//
//   * fake-data generation: make_fake_dataset (src/generator.cpp:10-48) =
//     Signal::get_efficiency (src/signal.cpp:172-197) + EvalHist::RandomSample
//     (src/pdfz.cpp:817-922).  The reference samples ROOT histograms (TH1::GetRandom,
//     TH2::GetRandom2, TH3::GetRandom3: inversion of the bin CDF, uniform inside the bin) with
//     gRandom; ROOT is not vendored, so this is a restatement of that published algorithm on
//     Philox4x32-10 counters -- RNG-dependent results are statistically, not bitwise, comparable.
//   * the Contour error estimator over the resident chain rows
//     (src/error_estimators/contour.cpp:18-69 + LikelihoodSpace::get_contour,
//     src/likelihood.cpp:88-101): min NLL, rows within delta of it, per-parameter extents.
#include "common.cuh"
#include "kernels.h"
#include "nll_device.cuh"

namespace sxmc {

// ---- counter-based random numbers of the generator ---------------------------------------
// key = (seed, stream id = experiment index); counter = (index, attempt, purpose, signal)
__device__ __forceinline__ uint4 gen_random(unsigned long long seed, unsigned long long stream_id,
                                            unsigned a, unsigned b, unsigned c, unsigned d) {
  const uint2 key = make_uint2((unsigned)seed ^ (unsigned)(stream_id * 0x9E3779B9u) ^ 0xA511E9B3u,
                               (unsigned)(seed >> 32) ^ (unsigned)(stream_id >> 7) ^ 0x63D83595u);
  return philox4x32_10(make_uint4(a, b, c, d), key);
}

// Poisson(mean): multiplication method below 10, Hoermann's transformed rejection (PTRS,
// 1993) above.  Exact for every mean; TRandom::Poisson draws from the same distribution with
// a different algorithm (src/pdfz.cpp:838).
__device__ long long poisson_draw(double mean, unsigned long long seed, unsigned long long stream_id,
                                  unsigned sig) {
  if (!(mean > 0)) return 0;
  unsigned n = 0;
  auto next = [&]() {
    const uint4 r = gen_random(seed, stream_id, n++, 0u, 7u, sig);
    return make_double2(u01(r.x, r.y), u01(r.z, r.w));
  };
  if (mean < 10) {
    const double limit = exp(-mean);
    double prod = 1;
    long long k = -1;
    do {
      const double2 u = next();
      prod *= u.x;
      ++k;
      if (prod <= limit) break;
      prod *= u.y;
      ++k;
    } while (prod > limit);
    return k;
  }
  const double slam = sqrt(mean), loglam = log(mean);
  const double b = 0.931 + 2.53 * slam;
  const double a = -0.059 + 0.02483 * b;
  const double invalpha = 1.1239 + 1.1328 / (b - 3.4);
  const double vr = 0.9277 - 3.6224 / (b - 2);
  for (int guard = 0; guard < 100000; ++guard) {
    const double2 uv = next();
    const double U = uv.x - 0.5, V = uv.y;
    const double us = 0.5 - fabs(U);
    const double k = floor((2 * a / us + b) * U + mean + 0.43);
    if (us >= 0.07 && V <= vr) return (long long)k;
    if (k < 0 || (us < 0.013 && V > us)) continue;
    if (log(V) + log(invalpha) - log(a / (us * us) + b) <= -mean + k * loglam - lgamma(k + 1)) return (long long)k;
  }
  return (long long)rint(mean);
}

// One thread per signal: events to draw (src/pdfz.cpp:836-842) and their offsets.
// counts[0..n) = observed, counts[n] = total; offsets[0..n] exclusive prefix.
__global__ void gen_counts_kernel(const GenSignal* __restrict__ sigs, int n, int poisson,
                                  unsigned long long seed, unsigned long long stream_id,
                                  unsigned long long* counts, unsigned long long* offsets) {
  __shared__ unsigned long long s_obs[256];
  const int j = threadIdx.x;
  if (j < n) {
    const GenSignal& g = sigs[j];
    // eff = 1.0 * norm / n_mc (src/signal.cpp:196); nevents = nexpected * eff (generator.cpp:33)
    const double eff = 1.0 * (double)*g.norm / (double)g.n_mc;
    const double mean = g.nexpected * eff;
    long long obs;
    if (poisson) obs = poisson_draw(mean, seed, stream_id, (unsigned)j);
    else obs = (long long)(unsigned)rintf((float)mean);   // nint(float): TMath::Nint, half to even
    s_obs[j] = obs > 0 ? (unsigned long long)obs : 0ull;
    counts[j] = s_obs[j];
  }
  __syncthreads();
  if (j == 0) {
    unsigned long long t = 0;
    for (int k = 0; k < n; ++k) { offsets[k] = t; t += s_obs[k]; }
    offsets[n] = t;
    counts[n] = t;
  }
}

// first index k in [0, nb) with cum[k] > t  (cum = inclusive prefix sums of the bin counts)
__device__ __forceinline__ int cdf_search(const unsigned* __restrict__ cum, int nb, double t) {
  int lo = 0, hi = nb - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((double)cum[mid] > t) hi = mid; else lo = mid + 1;
  }
  return lo;
}

// One thread per event: which signal, then inversion of that signal's bin CDF and a uniform
// position inside the bin; redrawn while outside the observables' limits (:851-855 etc.).
__global__ void gen_events_kernel(const GenSignal* __restrict__ sigs, int n, GenGeo geo,
                                  unsigned long long seed, unsigned long long stream_id,
                                  const unsigned long long* __restrict__ offsets, float* __restrict__ events,
                                  unsigned long long cap_events) {
  const unsigned long long total = offsets[n];
  const int nobs = geo.nobs;
  for (unsigned long long e = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; e < total;
       e += (unsigned long long)gridDim.x * blockDim.x) {
    if (e >= cap_events) return;
    int j = 0;
    while (j + 1 < n && offsets[j + 1] <= e) ++j;
    const GenSignal& g = sigs[j];
    const unsigned idx = (unsigned)(e - offsets[j]);
    const unsigned totalw = geo.total_nbins > 0 ? g.cum[geo.total_nbins - 1] : 0u;
    double x[3] = {0, 0, 0};
    if (totalw == 0) {
      // empty histogram: TH1::GetRandom returns NaN, which passes the limit test
      for (int o = 0; o < nobs; ++o) x[o] = __longlong_as_double(0x7ff8000000000000ll);
    } else {
      for (unsigned attempt = 0; attempt < 4096u; ++attempt) {
        const uint4 r = gen_random(seed, stream_id, idx, attempt, 1u, (unsigned)j);
        const double r1 = u01(r.x, r.y);
        const double t = r1 * (double)totalw;
        const int b = cdf_search(g.cum, geo.total_nbins, t);
        const double c0 = b ? (double)g.cum[b - 1] : 0.0, c1 = (double)g.cum[b];
        if (nobs == 1) {
          // TH1::GetRandom: the same r1 places the event inside the bin
          x[0] = geo.lower[0] + geo.width[0] * ((double)b + (t - c0) / (c1 - c0));
        } else {
          const uint4 q = gen_random(seed, stream_id, idx, attempt, 2u, (unsigned)j);
          const double u[3] = {u01(r.z, r.w), u01(q.x, q.y), u01(q.z, q.w)};
          int rem = b;
          for (int o = 0; o < nobs; ++o) {          // row-major flat index, last observable fastest
            const int k = rem / geo.stride[o];
            rem -= k * geo.stride[o];
            x[o] = geo.lower[o] + geo.width[o] * ((double)k + u[o]);
          }
        }
        bool ok = true;
        for (int o = 0; o < geo.nlimits && o < nobs; ++o)
          if (x[o] > (double)geo.lim_hi[o] || x[o] < (double)geo.lim_lo[o]) ok = false;
        if (ok) break;
      }
    }
    float* row = events + e * (unsigned long long)(nobs + 1);
    for (int o = 0; o < nobs; ++o) row[o] = (float)x[o];     // events.push_back(double) -> float
    row[nobs] = (float)g.dataset;
  }
}

int launch_generate(const GenSignal* sigs_dev, int n, const GenGeo& geo, int poisson,
                    unsigned long long seed, unsigned long long stream_id, unsigned long long* counts_dev,
                    unsigned long long* offsets_dev, float* events_dev, unsigned long long cap_events,
                    cudaStream_t stream) {
  if (n <= 0 || n > 256) { set_error("generator: 1..256 signals supported"); return SXMC_EINVAL; }
  gen_counts_kernel<<<1, 256, 0, stream>>>(sigs_dev, n, poisson, seed, stream_id, counts_dev, offsets_dev);
  SXMC_CUDA(cudaGetLastError());
  unsigned long long blocks = (cap_events + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  gen_events_kernel<<<(unsigned)blocks, 256, 0, stream>>>(sigs_dev, n, geo, seed, stream_id, offsets_dev,
                                                          events_dev, cap_events);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

// events per bin of one dataset (grouped lookup of the chain batch kernels)
__global__ void event_mult_kernel(const int* __restrict__ read_bins, long long n, int tnb, unsigned* mult) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int b = read_bins[i];
    if (b >= 0 && b < tnb) atomicAdd(mult + b, 1u);
  }
}
int launch_event_mult(const int* read_bins_dev, long long nevents, int tnb, unsigned* mult_dev, cudaStream_t stream) {
  SXMC_CUDA(cudaMemsetAsync(mult_dev, 0, sizeof(unsigned) * (size_t)tnb, stream));
  if (nevents <= 0) return SXMC_OK;
  long long blocks = (nevents + 255) / 256;
  if (blocks > 148 * 4) blocks = 148 * 4;
  event_mult_kernel<<<(unsigned)blocks, 256, 0, stream>>>(read_bins_dev, nevents, tnb, mult_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

// inclusive prefix sums of one histogram, single block (histograms are small and this runs once
// per generator, not per experiment)
__global__ void cum_kernel(const unsigned* __restrict__ bins, int nb, unsigned* __restrict__ cum) {
  __shared__ unsigned s_part[1024];
  const int t = threadIdx.x, nt = blockDim.x;
  const int per = (nb + nt - 1) / nt;
  const int a = min(nb, t * per), b = min(nb, a + per);
  unsigned s = 0;
  for (int i = a; i < b; ++i) s += bins[i];
  s_part[t] = s;
  __syncthreads();
  if (t == 0) {
    unsigned run = 0;
    for (int k = 0; k < nt; ++k) { const unsigned v = s_part[k]; s_part[k] = run; run += v; }
  }
  __syncthreads();
  unsigned run = s_part[t];
  for (int i = a; i < b; ++i) { run += bins[i]; cum[i] = run; }
}

int launch_cum(const unsigned* bins_dev, int nb, unsigned* cum_dev, cudaStream_t stream) {
  if (nb <= 0) return SXMC_OK;
  cum_kernel<<<1, 1024, 0, stream>>>(bins_dev, nb, cum_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

// =========================================================================================
// Contour estimator over the resident rows [nrows][np+1] (last column = NLL)
// =========================================================================================
__device__ __forceinline__ float atomic_min_f(float* addr, float v) {   // v is never NaN here
  return v >= 0 ? __int_as_float(atomicMin((int*)addr, __float_as_int(v)))
                : __uint_as_float(atomicMax((unsigned*)addr, __float_as_uint(v)));
}
__device__ __forceinline__ float atomic_max_f(float* addr, float v) {
  return v >= 0 ? __int_as_float(atomicMax((int*)addr, __float_as_int(v)))
                : __uint_as_float(atomicMin((unsigned*)addr, __float_as_uint(v)));
}

// out[0] = min over rows of the NLL column (TTree::GetMinimum("likelihood"))
__global__ void rows_min_nll_kernel(const float* __restrict__ rows, long long nrows, int np, float* out) {
  float m = __int_as_float(0x7f800000);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nrows;
       i += (long long)gridDim.x * blockDim.x) {
    const float v = rows[i * (np + 1) + np];
    if (v < m) m = v;
  }
  for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_down_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m < __int_as_float(0x7f800000)) atomic_min_f(out, m);
}

// For every level l (thresholds thr[l], TTreeFormula "likelihood+<off><thr" evaluated in
// double): count of selected rows and per-parameter min/max over them.
// out layout: counts[nlevels] (as unsigned long long), then mins[nlevels][np], maxs[nlevels][np].
__global__ void rows_extents_kernel(const float* __restrict__ rows, long long nrows, int np, double off,
                                    ContourLevels lv, unsigned long long* counts, float* mins, float* maxs) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nrows;
       i += (long long)gridDim.x * blockDim.x) {
    const float* row = rows + i * (np + 1);
    const double d = (double)row[np] + off;
    for (int l = 0; l < lv.n; ++l) {
      if (d < lv.thr[l]) {
        atomicAdd(counts + l, 1ull);
        for (int k = 0; k < np; ++k) {
          const float v = row[k] + 0.0f;   // -0 -> +0 (the integer-ordered atomics below)
          if (v == v) { atomic_min_f(mins + (size_t)l * np + k, v); atomic_max_f(maxs + (size_t)l * np + k, v); }
        }
      }
    }
  }
}

int launch_rows_min_nll(const float* rows_dev, long long nrows, int np, float* out_dev, cudaStream_t stream) {
  if (nrows <= 0) return SXMC_OK;
  long long blocks = (nrows + 255) / 256;
  if (blocks > 148 * 4) blocks = 148 * 4;
  rows_min_nll_kernel<<<(unsigned)blocks, 256, 0, stream>>>(rows_dev, nrows, np, out_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

int launch_rows_extents(const float* rows_dev, long long nrows, int np, double off, const ContourLevels& lv,
                        unsigned long long* counts_dev, float* mins_dev, float* maxs_dev, cudaStream_t stream) {
  if (nrows <= 0) return SXMC_OK;
  long long blocks = (nrows + 255) / 256;
  if (blocks > 148 * 4) blocks = 148 * 4;
  rows_extents_kernel<<<(unsigned)blocks, 256, 0, stream>>>(rows_dev, nrows, np, off, lv, counts_dev, mins_dev,
                                                            maxs_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

}  // namespace sxmc
