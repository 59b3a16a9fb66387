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

// Shared-memory working copy of the mutable chain state.  Every block prefetches it at kernel
// start (the state only changes in the tail of the previous launch), so the tail of the last
// block runs out of shared memory instead of a chain of dependent global loads.
struct FinState {
  double current[SXMC_MAX_PARAMETERS];
  double proposed[SXMC_MAX_PARAMETERS];
  double means[SXMC_MAX_PARAMETERS];
  double sigmas[SXMC_MAX_PARAMETERS];
  double stat_sx[SXMC_MAX_PARAMETERS];
  double stat_sxx[SXMC_MAX_PARAMETERS];
  float jump_width[SXMC_MAX_PARAMETERS];
  int fixed[SXMC_MAX_PARAMETERS];
  double stat_n, current_nll, event_total;
  float scale_factor;
  int debug_mode;
  unsigned step, burnin_steps;
  long long nrows, row_cap, accepted;
  unsigned long long seed, stream_id;
  const double* tape_normals; unsigned long long n_normals, i_normal;
  const double* tape_uniforms; unsigned long long n_uniforms, i_uniform;
  float* rows;
  double* nll_trace;
  double* prop_trace;
};

__device__ __forceinline__ void prefetch_state(FinState& F, const ChainState* st, int np, int tid,
                                               int nthreads) {
  for (int k = tid; k < np; k += nthreads) {
    F.current[k] = st->current[k];
    F.proposed[k] = st->proposed[k];
    F.means[k] = st->means[k];
    F.sigmas[k] = st->sigmas[k];
    F.stat_sx[k] = st->stat_sx[k];
    F.stat_sxx[k] = st->stat_sxx[k];
    F.jump_width[k] = st->jump_width[k];
    F.fixed[k] = st->fixed[k];
  }
  // scalars: one per thread of the last warp, all in flight together
  const int t = tid - (nthreads - 32);
  switch (t) {
    case 0: F.stat_n = st->stat_n; break;
    case 1: F.current_nll = st->current_nll; break;
    case 2: F.event_total = st->event_total; break;
    case 3: F.scale_factor = st->scale_factor; break;
    case 4: F.debug_mode = st->debug_mode; break;
    case 5: F.step = st->step; break;
    case 6: F.burnin_steps = st->burnin_steps; break;
    case 7: F.nrows = st->nrows; break;
    case 8: F.row_cap = st->row_cap; break;
    case 9: F.accepted = st->accepted; break;
    case 10: F.seed = st->seed; break;
    case 11: F.stream_id = st->stream_id; break;
    case 12: F.tape_normals = st->tape_normals; break;
    case 13: F.n_normals = st->n_normals; break;
    case 14: F.i_normal = st->i_normal; break;
    case 15: F.tape_uniforms = st->tape_uniforms; break;
    case 16: F.n_uniforms = st->n_uniforms; break;
    case 17: F.i_uniform = st->i_uniform; break;
    case 18: F.rows = st->rows; break;
    case 19: F.nll_trace = st->nll_trace; break;
    case 20: F.prop_trace = st->prop_trace; break;
    default: break;
  }
}

__device__ __forceinline__ uint4 fin_random(const FinState& F, unsigned step, unsigned index,
                                            unsigned purpose) {
  const uint2 key = make_uint2((unsigned)F.seed ^ (unsigned)(F.stream_id * 0x9E3779B9u),
                               (unsigned)(F.seed >> 32) ^ (unsigned)(F.stream_id >> 7) ^ 0x5851F42Du);
  return philox4x32_10(make_uint4(step, index, purpose, (unsigned)F.stream_id), key);
}

// pick_new_vector_device (src/nll_kernels.cpp:31-53), executed by one warp; the new point goes
// to shared memory and to the global state (the next launch's binning kernel reads it there).
__device__ void propose(FinState& F, ChainState* st, int np, unsigned step_for_rng, unsigned purpose,
                        int lane) {
  unsigned long long tape_base = F.i_normal;
  for (int k0 = 0; k0 < np; k0 += 32) {
    const int k = k0 + lane;
    const bool floating = k < np && F.jump_width[k] > 0;
    const unsigned m = __ballot_sync(0xffffffffu, floating);
    if (k < np) {
      double v = F.current[k];
      if (floating) {
        double z;
        if (F.tape_normals) {
          const unsigned long long ti = tape_base + __popc(m & ((1u << lane) - 1));
          z = ti < F.n_normals ? F.tape_normals[ti] : 0.0;
        } else {
          const uint4 r = fin_random(F, step_for_rng, (unsigned)k, purpose);
          const double u1 = u01(r.x, r.y), u2 = u01(r.z, r.w);
          z = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
        }
        v = __dadd_rn(v, __dmul_rn((double)F.jump_width[k], z));
      }
      st->proposed[k] = v;
    }
    tape_base += __popc(m);
  }
  if (lane == 0 && F.tape_normals) st->i_normal = tape_base;
}

// Everything after the event sum, run by warp 0 of the last block to finish, on the shared
// copy F; results are written back to the global state with independent stores.
// s_term: nsignals normalisation terms + nparameters constraint terms, filled beforehand.
__device__ void finish_step(FinState& F, ChainState* st, int np, int ns, int nsources, double ev_new,
                            int mode, const double* s_term, int lane) {
  // --- burn-in re-tune of the jump widths (src/mcmc.cpp:274-311), before this step's decision
  const unsigned step = F.step;
  bool reset_stats = false;
  if (mode == NLL_MODE_STEP && (step == F.burnin_steps || step == 2 * F.burnin_steps)) {
    const double n = F.stat_n;
    for (int k = lane; k < np; k += 32) {
      if (F.fixed[k]) continue;
      double rms = 0;
      if (n > 0) {
        const double m = F.stat_sx[k] / n;
        rms = sqrt(fabs(__dadd_rn(F.stat_sxx[k] / n, -__dmul_rn(m, m))));
      }
      const double fit_width = rms > 0 ? rms : (double)F.jump_width[k];
      F.jump_width[k] = (float)__dmul_rn((double)F.scale_factor, fit_width);
      st->jump_width[k] = F.jump_width[k];
    }
    reset_stats = !F.debug_mode;
    __syncwarp();
  }
  const double stat_n0 = reset_stats ? 0.0 : F.stat_n;
  const long long row = reset_stats ? 0 : F.nrows;

  // --- nll_total_device (src/nll_kernels.cpp:150-188), terms summed in the reference's order
  double nll = 0, ev = 0;
  if (lane == 0) {
    ev = ev_new;
    if (mode != NLL_MODE_EVAL_ONLY) {
      if (!isnan(ev_new)) { F.event_total = ev_new; st->event_total = ev_new; }   // :113-115
      ev = F.event_total;
    }
    double sum = -ev;
    if (isnan(sum)) {
      nll = 1e18;
    } else {
      for (int i = 0; i < ns; ++i) sum = __dadd_rn(sum, s_term[i]);
      bool bad = false;
      for (int i = 0; i < np; ++i) {
        if (i < nsources && F.proposed[i] < 0) { bad = true; break; }
        if (F.sigmas[i] > 0) sum = __dadd_rn(sum, s_term[ns + i]);
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
    propose(F, st, np, 0xffffffffu, 1u, lane);   // the pick_new_vector before the loop (src/mcmc.cpp:252)
    return;
  }

  // --- jump_decider_device (src/nll_kernels.cpp:57-86)
  int accept = 0;
  if (lane == 0) {
    st->proposed_nll = nll;
    if (F.nll_trace) F.nll_trace[step] = nll;
    double u;
    if (F.tape_uniforms) {
      u = F.i_uniform < F.n_uniforms ? F.tape_uniforms[F.i_uniform] : 0.5;
      st->i_uniform = F.i_uniform + 1;
    } else {
      const uint4 r = fin_random(F, step, 0u, 2u);
      u = u01(r.x, r.y);
    }
    const double nc = F.current_nll;
    accept = (F.debug_mode || (nll < nc || u <= exp(nc - nll))) ? 1 : 0;
    if (accept) { F.current_nll = nll; st->current_nll = nll; st->accepted = F.accepted + 1; }
  }
  accept = __shfl_sync(0xffffffffu, accept, 0);
  const double cur_nll = __shfl_sync(0xffffffffu, F.current_nll, 0);
  const bool store = row < F.row_cap;
  float* out = F.rows + row * (np + 1);
  for (int k = lane; k < np; k += 32) {
    if (F.prop_trace) F.prop_trace[(unsigned long long)step * np + k] = F.proposed[k];
    if (accept) { F.current[k] = F.proposed[k]; st->current[k] = F.proposed[k]; }
    const float xf = (float)F.current[k];
    if (store) out[k] = xf;
    // running statistics for the burn-in RMS, accumulated in row order (ROOT TH1::GetStats)
    const double x = (double)xf;
    const double sx0 = reset_stats ? 0.0 : F.stat_sx[k];
    const double sxx0 = reset_stats ? 0.0 : F.stat_sxx[k];
    st->stat_sx[k] = __dadd_rn(sx0, x);
    st->stat_sxx[k] = __dadd_rn(sxx0, __dmul_rn(x, x));
  }
  if (lane == 0) {
    if (store) out[np] = (float)cur_nll;
    st->nrows = row + 1;
    st->stat_n = stat_n0 + 1;
    st->step = step + 1;
  }
  __syncwarp();
  propose(F, st, np, step, 1u, lane);
}

// =========================================================================================
__global__ void __launch_bounds__(NLL_THREADS)
lookup_nll_kernel(const NllSignal* __restrict__ sigs, const unsigned* __restrict__ norms, int ns,
                  int np, int nsources, long long ne, ChainState* st, double* partials,
                  unsigned* ticket, int mode, float* lut, unsigned* zero_ptr,
                  unsigned long long zero_words) {
  __shared__ double s_c[NLL_MAX_SIGNALS];
  __shared__ double s_bn[NLL_MAX_SIGNALS];
  __shared__ const unsigned* s_bins[NLL_MAX_SIGNALS];
  __shared__ const int* s_rb[NLL_MAX_SIGNALS];
  __shared__ int s_nb[NLL_MAX_SIGNALS];
  __shared__ double s_warp[NLL_THREADS / 32];
  __shared__ double s_term[NLL_MAX_SIGNALS + SXMC_MAX_PARAMETERS];
  __shared__ int s_last;
  __shared__ FinState F;

  const int tid = threadIdx.x;
  const long long gtid = blockIdx.x * (long long)NLL_THREADS + tid;
  const long long gstride = (long long)gridDim.x * NLL_THREADS;

  prefetch_state(F, st, np, tid, NLL_THREADS);
  // per-signal tables; norms[j] is addressed directly so that it does not wait for sigs[j]
  NllSignal my_sig;
  unsigned my_norm = 0;
  if (tid < ns) { my_sig = sigs[tid]; my_norm = norms[tid]; }

  // clear the histogram buffer the NEXT step bins into (double-buffered; nobody reads it now)
  for (unsigned long long w = gtid; w < zero_words; w += gstride) zero_ptr[w] = 0u;
  __syncthreads();

  if (tid < ns) {
    const int j = tid;
    const float eff = (float)(1.0 * my_norm / my_sig.n_mc);                 // nll_kernels.cpp:105
    const double rate = __dmul_rn(F.proposed[my_sig.source_id], my_sig.nexpected);
    s_c[j] = __dmul_rn(rate, (double)eff);
    s_bn[j] = __dmul_rn((double)my_norm, my_sig.bin_volume);                // pdfz.cpp:417
    s_bins[j] = my_sig.bins;
    s_rb[j] = my_sig.read_bins;
    s_nb[j] = my_sig.total_nbins;
    // normalisation term of nll_total (nll_kernels.cpp:167-171)
    s_term[j] = __dmul_rn(rate, (double)my_norm) / (double)my_sig.n_mc;
  }
  // Gaussian constraints (nll_kernels.cpp:176-185)
  for (int i = tid; i < np; i += NLL_THREADS) {
    const double sg = F.sigmas[i];
    double t = 0;
    if (sg > 0) {
      const double x = __dadd_rn(F.proposed[i], -F.means[i]) / sg;
      t = __dmul_rn(__dmul_rn(0.5, x), x);
    }
    s_term[ns + i] = t;
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
  __syncthreads();
  if (tid < 32) finish_step(F, st, np, ns, nsources, s_warp[0], mode, s_term, tid);
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
  lookup_nll_kernel<<<L.grid, NLL_THREADS, 0, stream>>>(L.sigs_dev, L.norms_dev, L.nsignals,
                                                        L.nparameters, L.nsources, L.nevents, L.state,
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
