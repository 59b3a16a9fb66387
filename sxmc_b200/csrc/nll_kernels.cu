// nll_kernels.cu -- PDF lookup + event likelihood + NLL total + Metropolis + proposal as ONE
// launch per step.  Replaces eval_pdf (reference src/pdfz.cpp:411-436), nll_event_chunks
// (src/nll_kernels.cpp:89-116), nll_event_reduce (:120-146), nll_total (:150-188),
// jump_decider (:57-86), pick_new_vector (:31-53) and the host-side burn-in re-tune of
// src/mcmc.cpp:274-311.  RNG: Philox4x32-10 keyed by (seed, chain stream), counter =
// (step, parameter, purpose) -- replaces init_device_rngs / XORWOW (:19-26).
#include "common.cuh"
#include "kernels.h"

namespace sxmc {

constexpr int NLL_THREADS = 256;
constexpr int NLL_MAX_SIGNALS = 128;

// ---- Philox4x32-10 ----------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const unsigned M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const unsigned hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

__device__ __forceinline__ double u01(unsigned a, unsigned b) {   // (0,1), 53 bits
  const unsigned long long x = ((unsigned long long)a << 32) | b;
  return ((double)(x >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}

__device__ __forceinline__ uint4 chain_random(const ChainState* st, unsigned step, unsigned index,
                                              unsigned purpose) {
  const uint2 key = make_uint2((unsigned)st->seed ^ (unsigned)(st->stream_id * 0x9E3779B9u),
                               (unsigned)(st->seed >> 32) ^ (unsigned)(st->stream_id >> 7) ^ 0x5851F42Du);
  return philox4x32_10(make_uint4(step, index, purpose, (unsigned)st->stream_id), key);
}

__device__ __forceinline__ double chain_normal(const ChainState* st, unsigned step, unsigned k,
                                               unsigned purpose) {
  const uint4 r = chain_random(st, step, k, purpose);
  const double u1 = u01(r.x, r.y), u2 = u01(r.z, r.w);
  return sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
}

// ---- block-wide deterministic double sum -------------------------------------------------
__device__ __forceinline__ double block_sum(double v, double* s_warp) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  double t = 0;
  if (threadIdx.x == 0)
    for (int w = 0; w < NLL_THREADS / 32; ++w) t += s_warp[w];
  return t;   // valid on thread 0
}

// pick_new_vector_device (src/nll_kernels.cpp:31-53), executed by one warp.
__device__ void propose(ChainState* st, unsigned step_for_rng, unsigned purpose, int lane) {
  const int np = st->nparameters;
  unsigned long long tape_base = st->i_normal;
  for (int k0 = 0; k0 < np; k0 += 32) {
    const int k = k0 + lane;
    const bool floating = k < np && st->jump_width[k] > 0;
    const unsigned m = __ballot_sync(0xffffffffu, floating);
    if (k < np) {
      double v = st->current[k];
      if (floating) {
        double z;
        if (st->tape_normals) {
          const unsigned long long ti = tape_base + __popc(m & ((1u << lane) - 1));
          z = ti < st->n_normals ? st->tape_normals[ti] : 0.0;
        } else {
          z = chain_normal(st, step_for_rng, (unsigned)k, purpose);
        }
        v = __dadd_rn(v, __dmul_rn((double)st->jump_width[k], z));
      }
      st->proposed[k] = v;
    }
    tape_base += __popc(m);
  }
  __syncwarp();
  if (lane == 0 && st->tape_normals) st->i_normal = tape_base;
}

// Everything after the event sum, run by warp 0 of the last block to finish.
// s_term: scratch of nsignals + nparameters doubles filled by the whole block beforehand.
__device__ void finish_step(ChainState* st, double ev_new, int mode, const double* s_term,
                            int lane) {
  const int np = st->nparameters, ns = st->nsignals;

  // --- burn-in re-tune of the jump widths (src/mcmc.cpp:274-311), before this step's decision
  const unsigned step = st->step;
  if (mode == NLL_MODE_STEP && (step == st->burnin_steps || step == 2 * st->burnin_steps)) {
    const double n = st->stat_n;
    for (int k = lane; k < np; k += 32) {
      if (st->fixed[k]) continue;
      double rms = 0;
      if (n > 0) {
        const double m = st->stat_sx[k] / n;
        rms = sqrt(fabs(__dadd_rn(st->stat_sxx[k] / n, -__dmul_rn(m, m))));
      }
      const double fit_width = rms > 0 ? rms : (double)st->jump_width[k];
      st->jump_width[k] = (float)__dmul_rn((double)st->scale_factor, fit_width);
    }
    __syncwarp();
    if (!st->debug_mode) {
      for (int k = lane; k < np; k += 32) { st->stat_sx[k] = 0; st->stat_sxx[k] = 0; }
      if (lane == 0) { st->stat_n = 0; st->nrows = 0; }
    }
    __syncwarp();
  }

  // --- nll_total_device (src/nll_kernels.cpp:150-188), terms summed in the reference's order
  double nll = 0, ev = 0;
  if (lane == 0) {
    ev = ev_new;
    if (mode != NLL_MODE_EVAL_ONLY) {
      if (!isnan(ev_new)) st->event_total = ev_new;   // :113-115 keeps the old sum on NaN
      ev = st->event_total;
    }
    double sum = -ev;
    if (isnan(sum)) {
      nll = 1e18;
    } else {
      for (int i = 0; i < ns; ++i) sum = __dadd_rn(sum, s_term[i]);
      bool bad = false;
      for (int i = 0; i < np; ++i) {
        if (i < st->nsources && st->proposed[i] < 0) { bad = true; break; }
        if (st->sigmas[i] > 0) sum = __dadd_rn(sum, s_term[ns + i]);
      }
      nll = bad ? 1e18 : sum;
    }
  }
  nll = __shfl_sync(0xffffffffu, nll, 0);

  if (mode == NLL_MODE_EVAL_ONLY) {
    if (lane == 0) { st->last_nll = nll; st->last_event_sum = ev; }
    return;
  }
  if (mode == NLL_MODE_INIT) {
    if (lane == 0) { st->current_nll = nll; st->last_nll = nll; st->last_event_sum = ev; }
    __syncwarp();
    propose(st, 0xffffffffu, 1u, lane);   // the pick_new_vector before the loop (src/mcmc.cpp:252)
    return;
  }

  // --- jump_decider_device (src/nll_kernels.cpp:57-86)
  int accept = 0;
  if (lane == 0) {
    st->proposed_nll = nll;
    if (st->nll_trace) st->nll_trace[step] = nll;
    double u;
    if (st->tape_uniforms) {
      u = st->i_uniform < st->n_uniforms ? st->tape_uniforms[st->i_uniform] : 0.5;
      st->i_uniform += 1;
    } else {
      const uint4 r = chain_random(st, step, 0u, 2u);
      u = u01(r.x, r.y);
    }
    const double nc = st->current_nll;
    accept = (st->debug_mode || (nll < nc || u <= exp(nc - nll))) ? 1 : 0;
    if (accept) { st->current_nll = nll; st->accepted += 1; }
  }
  accept = __shfl_sync(0xffffffffu, accept, 0);
  const long long row = st->nrows;
  const bool store = row < st->row_cap;
  __syncwarp();   // every lane has read nrows before lane 0 advances it
  float* out = st->rows + row * (np + 1);
  for (int k = lane; k < np; k += 32) {
    if (st->prop_trace) st->prop_trace[(unsigned long long)step * np + k] = st->proposed[k];
    if (accept) st->current[k] = st->proposed[k];
    const float xf = (float)st->current[k];
    if (store) out[k] = xf;
    // running statistics for the burn-in RMS, accumulated in row order (ROOT TH1::GetStats)
    const double x = (double)xf;
    st->stat_sx[k] = __dadd_rn(st->stat_sx[k], x);
    st->stat_sxx[k] = __dadd_rn(st->stat_sxx[k], __dmul_rn(x, x));
  }
  if (lane == 0) {
    if (store) out[np] = (float)st->current_nll;
    st->nrows = row + 1;
    st->stat_n += 1;
    st->step = step + 1;
  }
  __syncwarp();
  propose(st, step, 1u, lane);
}

// =========================================================================================
__global__ void __launch_bounds__(NLL_THREADS)
lookup_nll_kernel(const NllSignal* __restrict__ sigs, int ns, long long ne, ChainState* st,
                  double* partials, unsigned* ticket, int mode, float* lut,
                  unsigned* zero_ptr, unsigned long long zero_words) {
  __shared__ double s_c[NLL_MAX_SIGNALS];
  __shared__ double s_bn[NLL_MAX_SIGNALS];
  __shared__ const unsigned* s_bins[NLL_MAX_SIGNALS];
  __shared__ const int* s_rb[NLL_MAX_SIGNALS];
  __shared__ int s_nb[NLL_MAX_SIGNALS];
  __shared__ double s_warp[NLL_THREADS / 32];
  __shared__ double s_term[NLL_MAX_SIGNALS + SXMC_MAX_PARAMETERS];
  __shared__ int s_last;

  const int tid = threadIdx.x;
  const long long gtid = blockIdx.x * (long long)NLL_THREADS + tid;
  const long long gstride = (long long)gridDim.x * NLL_THREADS;

  // clear the histogram buffer the NEXT step bins into (double-buffered; nobody reads it now)
  for (unsigned long long w = gtid; w < zero_words; w += gstride) zero_ptr[w] = 0u;

  const double* pars = st->proposed;
  for (int j = tid; j < ns; j += NLL_THREADS) {
    const NllSignal s = sigs[j];
    const unsigned norm = *s.norm;
    const float eff = (float)(1.0 * norm / s.n_mc);                 // nll_kernels.cpp:105
    s_c[j] = __dmul_rn(__dmul_rn(pars[s.source_id], s.nexpected), (double)eff);
    s_bn[j] = __dmul_rn((double)norm, s.bin_volume);                // pdfz.cpp:417
    s_bins[j] = s.bins;
    s_rb[j] = s.read_bins;
    s_nb[j] = s.total_nbins;
  }
  __syncthreads();

  double sum = 0;
  for (long long i = gtid; i < ne; i += gstride) {
    double s = 0;
    for (int j = 0; j < ns; ++j) {
      const int rb = s_rb[j][i];
      float v = 0.0f;
      float vlut = 0.0f;
      if (rb >= s_nb[j]) {
        // one-past-the-end index of a point within an ulp of `upper`: undefined in the reference
      } else if (rb >= 0) {
        v = (float)((double)s_bins[j][rb] / s_bn[j]);               // pdfz.cpp:429-432
        vlut = v;
        if (isnan(v)) v = 0.0f;                                     // nll_kernels.cpp:107
      } else if (rb == -1) {
        vlut = __int_as_float(0x7fc00000);
      }
      if (lut) lut[(long long)j * ne + i] = vlut;
      s = __dadd_rn(s, __dmul_rn(s_c[j], (double)v));
    }
    if (s > 0) sum += log(s);
  }
  const double bsum = block_sum(sum, s_warp);
  if (tid == 0) {
    partials[blockIdx.x] = bsum;
    __threadfence();
    const unsigned t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;

  // ---- last block: total the partials in a fixed order, then finish the step -------------
  __threadfence();
  double acc = 0;
  for (int b = tid; b < (int)gridDim.x; b += NLL_THREADS) acc += __ldcg(partials + b);
  const double total = block_sum(acc, s_warp);
  if (tid == 0) { s_warp[0] = total; *ticket = 0u; }

  // normalisation terms and Gaussian constraints, in parallel (nll_kernels.cpp:167-185)
  const int np = st->nparameters;
  for (int i = tid; i < ns; i += NLL_THREADS) {
    const NllSignal s = sigs[i];
    s_term[i] = __dmul_rn(__dmul_rn(pars[s.source_id], s.nexpected), (double)*s.norm) / (double)s.n_mc;
  }
  for (int i = tid; i < np; i += NLL_THREADS) {
    const double sg = st->sigmas[i];
    double t = 0;
    if (sg > 0) {
      const double x = __dadd_rn(pars[i], -st->means[i]) / sg;
      t = __dmul_rn(__dmul_rn(0.5, x), x);
    }
    s_term[ns + i] = t;
  }
  __syncthreads();
  if (tid < 32) finish_step(st, s_warp[0], mode, s_term, tid);
}

int nll_grid_for(long long nevents) {
  long long g = (nevents + NLL_THREADS - 1) / NLL_THREADS;
  if (g < 1) g = 1;
  if (g > 148 * 8) g = 148 * 8;
  return (int)g;
}

int launch_nll(const NllLaunch& L, cudaStream_t stream) {
  if (L.nsignals > NLL_MAX_SIGNALS) {
    set_error("more than 128 signals in one fit are not supported by the NLL kernel");
    return SXMC_EINVAL;
  }
  lookup_nll_kernel<<<L.grid, NLL_THREADS, 0, stream>>>(L.sigs_dev, L.nsignals, L.nevents, L.state,
                                                        L.partials, L.ticket, L.mode, L.lut,
                                                        L.zero_ptr, L.zero_words);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

// =========================================================================================
// standalone MCMC::nll over a materialised lut (src/mcmc.cpp:390-415), one block.
// =========================================================================================
__global__ void __launch_bounds__(1024)
nll_lut_kernel(const float* __restrict__ lut, size_t ne, size_t ns, const double* pars, size_t npars,
               size_t nsources, const double* means, const double* sigmas, const double* nexpected,
               const unsigned* n_mc, const short* source_id, const unsigned* norms, double* out2) {
  __shared__ double s_red[32];
  double sum = 0;
  for (size_t i = threadIdx.x; i < ne; i += blockDim.x) {
    double s = 0;
    for (size_t j = 0; j < ns; ++j) {
      const float v = lut[j * ne + i];
      const float eff = (float)(1.0 * norms[j] / n_mc[j]);
      const double c = __dmul_rn(__dmul_rn(pars[source_id[j]], nexpected[j]), (double)eff);
      s = __dadd_rn(s, __dmul_rn(c, (double)(isnan(v) ? 0.0f : v)));
    }
    if (s > 0) sum += log(s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double ev = 0;
    for (unsigned w = 0; w < blockDim.x / 32; ++w) ev += s_red[w];
    double t = -ev;
    double nll = t;
    if (isnan(t)) {
      nll = 1e18;
    } else {
      for (size_t i = 0; i < ns; ++i)
        t = __dadd_rn(t, __dmul_rn(__dmul_rn(pars[source_id[i]], nexpected[i]), (double)norms[i]) /
                             (double)n_mc[i]);
      nll = t;
      for (size_t i = 0; i < npars; ++i) {
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
}

int launch_nll_lut(const float* lut, size_t ne, size_t ns, const double* pars, size_t npars,
                   size_t nsources, const double* means, const double* sigmas,
                   const double* nexpected, const unsigned* n_mc, const short* source_id,
                   const unsigned* norms, double* out2_dev, cudaStream_t stream) {
  nll_lut_kernel<<<1, 1024, 0, stream>>>(lut, ne, ns, pars, npars, nsources, means, sigmas, nexpected,
                                         n_mc, source_id, norms, out2_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

}  // namespace sxmc
