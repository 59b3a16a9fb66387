// nll_device.cuh -- device code shared by the per-step NLL kernel (nll_kernels.cu) and the
// persistent chain kernel (chain_kernels.cu): Philox4x32-10, the block reduction, the
// shared-memory working copy of the chain state and the step tail (NLL total, Metropolis
// decision, row append, burn-in re-tune, next proposal).
#pragma once
#include "common.cuh"

namespace sxmc {

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
template <int NT>
__device__ __forceinline__ double block_sum(double v, double* s_warp) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  double t = 0;
  if (threadIdx.x == 0)
    for (int w = 0; w < NT / 32; ++w) t += s_warp[w];
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
  int accept_flag;          // finish_step2: the decision, handed from warp 0 to warp 1
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
__device__ __forceinline__ double fin_normal(const FinState& F, unsigned step, unsigned k, unsigned purpose) {
  const uint4 r = fin_random(F, step, k, purpose);
  const double u1 = u01(r.x, r.y), u2 = u01(r.z, r.w);
  return sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
}

// The random numbers of the step F.step is about to take (one warp; lane-strided over the
// parameters), so that a warp with nothing else to do can draw them off the serial tail.
// Not applicable with tapes (consumed in order) or outside NLL_MODE_STEP.
__device__ __forceinline__ bool pre_rng_applies(const FinState& F, int mode) {
  return mode == NLL_MODE_STEP && !F.tape_normals && !F.tape_uniforms;
}
__device__ __forceinline__ void pre_rng_draw(const FinState& F, int np, int lane, double* z, double* u) {
  for (int k = lane; k < np; k += 32)
    z[k] = F.jump_width[k] > 0 ? fin_normal(F, F.step, (unsigned)k, 1u) : 0.0;
  if (lane == 31) {
    const uint4 r = fin_random(F, F.step, 0u, 2u);
    *u = u01(r.x, r.y);
  }
}

// pre_z (optional): the step's normals, computed ahead of time by an idle warp
static __device__ void propose(FinState& F, ChainState* st, int np, unsigned step_for_rng, unsigned purpose,
                        int lane, const double* pre_z = nullptr) {
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
          z = pre_z ? pre_z[k] : fin_normal(F, step_for_rng, (unsigned)k, purpose);
        }
        v = __dadd_rn(v, __dmul_rn((double)F.jump_width[k], z));
      }
      F.proposed[k] = v;
      st->proposed[k] = v;
    }
    tape_base += __popc(m);
  }
  __syncwarp();
  if (lane == 0 && F.tape_normals) { F.i_normal = tape_base; st->i_normal = tape_base; }
}

// Everything after the event sum, run by warp 0 of the last block to finish, on the shared
// copy F; results are written back to the global state with independent stores.
// s_term: nsignals normalisation terms + nparameters constraint terms, filled beforehand.
static __device__ void finish_step(FinState& F, ChainState* st, int np, int ns, int nsources, double ev_new,
                            int mode, const double* s_term, int lane,
                            const double* pre_z = nullptr, const double* pre_u = nullptr,
                            long long* clk = nullptr) {
#define FIN_CLK(i) do { if (clk && lane == 0) clk[i] = clock64(); } while (0)
  FIN_CLK(0);
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
  FIN_CLK(1);

  if (mode == NLL_MODE_EVAL_ONLY) {
    if (lane == 0) { st->last_nll = nll; st->last_event_sum = ev; }
    return;
  }
  if (mode == NLL_MODE_INIT) {
    if (lane == 0) { F.current_nll = nll; st->current_nll = nll; st->last_nll = nll; st->last_event_sum = ev; }
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
      F.i_uniform += 1;
      st->i_uniform = F.i_uniform;
    } else {
      if (pre_u) {
        u = *pre_u;
      } else {
        const uint4 r = fin_random(F, step, 0u, 2u);
        u = u01(r.x, r.y);
      }
    }
    const double nc = F.current_nll;
    accept = (F.debug_mode || (nll < nc || u <= exp(nc - nll))) ? 1 : 0;
    if (accept) { F.current_nll = nll; st->current_nll = nll; F.accepted += 1; st->accepted = F.accepted; }
  }
  accept = __shfl_sync(0xffffffffu, accept, 0);
  FIN_CLK(2);
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
    F.stat_sx[k] = __dadd_rn(sx0, x);
    F.stat_sxx[k] = __dadd_rn(sxx0, __dmul_rn(x, x));
    st->stat_sx[k] = F.stat_sx[k];
    st->stat_sxx[k] = F.stat_sxx[k];
  }
  __syncwarp();   // every lane has read F.nrows / F.stat_n / F.step before lane 0 advances them
  if (lane == 0) {
    if (store) out[np] = (float)cur_nll;
    F.nrows = row + 1;
    F.stat_n = stat_n0 + 1;
    F.step = step + 1;
    st->nrows = F.nrows;
    st->stat_n = F.stat_n;
    st->step = F.step;
  }
  __syncwarp();
  FIN_CLK(3);
  propose(F, st, np, step, 1u, lane, pre_z);
  FIN_CLK(4);
#undef FIN_CLK
}

// The same step finished by TWO warps (threads 0..63 of the CTA call it with wid = warp index):
// the tail is a single warp's chain of dependent instructions (~900 of them, ~7 cycles each), and
// after the decision its two halves are independent -- warp 0 draws the next proposal while
// warp 1 appends the row and updates the running sums.  Both read what they need of
// current/proposed into registers, meet at a 64-thread barrier, then write.  Falls back to
// finish_step (warp 0 alone) outside NLL_MODE_STEP and for more than 64 parameters (a lane owns
// parameters lane and lane + 32).
__device__ __forceinline__ void fin_bar64() { asm volatile("bar.sync 1, 64;" ::: "memory"); }

static __device__ void finish_step2(FinState& F, ChainState* st, int np, int ns, int nsources, double ev_new,
                                    int mode, const double* s_term, int lane, int wid,
                                    const double* pre_z = nullptr, const double* pre_u = nullptr,
                                    long long* clk = nullptr) {
  if (np > 64 || mode != NLL_MODE_STEP) {
    if (wid == 0) finish_step(F, st, np, ns, nsources, ev_new, mode, s_term, lane, pre_z, pre_u, clk);
    return;
  }
#define FIN_CLK(i) do { if (clk && wid == 0 && lane == 0) clk[i] = clock64(); } while (0)
  FIN_CLK(0);
  const unsigned step = F.step;
  const bool retune = step == F.burnin_steps || step == 2 * F.burnin_steps;
  const bool reset_stats = retune && !F.debug_mode;
  // a lane owns parameters lane and lane + 32
  const int k0 = lane, k1 = lane + 32;
  if (wid == 0) {
    // --- burn-in re-tune of the jump widths (src/mcmc.cpp:274-311), before this step's decision
    if (retune) {
      const double n = F.stat_n;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int k = h ? k1 : k0;
        if (k < np && !F.fixed[k]) {
          double rms = 0;
          if (n > 0) {
            const double m = F.stat_sx[k] / n;
            rms = sqrt(fabs(__dadd_rn(F.stat_sxx[k] / n, -__dmul_rn(m, m))));
          }
          const double fit_width = rms > 0 ? rms : (double)F.jump_width[k];
          F.jump_width[k] = (float)__dmul_rn((double)F.scale_factor, fit_width);
          st->jump_width[k] = F.jump_width[k];
        }
      }
      __syncwarp();
    }
    if (lane == 0) {
      // --- nll_total_device (src/nll_kernels.cpp:150-188), terms summed in the reference's order
      if (!isnan(ev_new)) { F.event_total = ev_new; st->event_total = ev_new; }   // :113-115
      const double ev = F.event_total;
      double nll;
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
      FIN_CLK(1);
      // --- jump_decider_device (src/nll_kernels.cpp:57-86)
      st->proposed_nll = nll;
      if (F.nll_trace) F.nll_trace[step] = nll;
      double u;
      if (F.tape_uniforms) {
        u = F.i_uniform < F.n_uniforms ? F.tape_uniforms[F.i_uniform] : 0.5;
        F.i_uniform += 1;
        st->i_uniform = F.i_uniform;
      } else if (pre_u) {
        u = *pre_u;
      } else {
        const uint4 r = fin_random(F, step, 0u, 2u);
        u = u01(r.x, r.y);
      }
      const double nc = F.current_nll;
      const int accept = (F.debug_mode || (nll < nc || u <= exp(nc - nll))) ? 1 : 0;
      if (accept) { F.current_nll = nll; st->current_nll = nll; F.accepted += 1; st->accepted = F.accepted; }
      F.accept_flag = accept;
      FIN_CLK(2);
    }
  }
  fin_bar64();                                   // the decision is visible to both warps
  const int accept = F.accept_flag;
  if (wid == 0) {
    // --- pick_new_vector_device (src/nll_kernels.cpp:31-53) from the point the chain is at now
    const bool fl0 = k0 < np && F.jump_width[k0] > 0, fl1 = k1 < np && F.jump_width[k1] > 0;
    const unsigned m0 = __ballot_sync(0xffffffffu, fl0), m1 = __ballot_sync(0xffffffffu, fl1);
    double v[2] = {0, 0};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = h ? k1 : k0;
      if (k >= np) continue;
      double x = accept ? F.proposed[k] : F.current[k];
      if (h ? fl1 : fl0) {
        double z;
        if (F.tape_normals) {                    // normals are consumed in parameter order
          const unsigned long long ti = F.i_normal + (h ? __popc(m0) : 0) + __popc((h ? m1 : m0) & ((1u << lane) - 1));
          z = ti < F.n_normals ? F.tape_normals[ti] : 0.0;
        } else {
          z = pre_z ? pre_z[k] : fin_normal(F, step, (unsigned)k, 1u);
        }
        x = __dadd_rn(x, __dmul_rn((double)F.jump_width[k], z));
      }
      v[h] = x;
    }
    fin_bar64();                                 // warp 1 has read the old proposal
    if (k0 < np) { F.proposed[k0] = v[0]; st->proposed[k0] = v[0]; }
    if (k1 < np) { F.proposed[k1] = v[1]; st->proposed[k1] = v[1]; }
    if (lane == 0 && F.tape_normals) { F.i_normal += __popc(m0) + __popc(m1); st->i_normal = F.i_normal; }
    FIN_CLK(3);
    FIN_CLK(4);
  } else {
    // --- the ntuple row and the running sums of the burn-in RMS (src/nll_kernels.cpp:78-84)
    const double stat_n0 = reset_stats ? 0.0 : F.stat_n;
    const long long row = reset_stats ? 0 : F.nrows;
    const double cur_nll = F.current_nll;
    const bool store = row < F.row_cap;
    float* out = F.rows + row * (np + 1);
    double oldp[2] = {0, 0}, cur[2] = {0, 0}, sx0[2] = {0, 0}, sxx0[2] = {0, 0};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = h ? k1 : k0;
      if (k < np) {
        oldp[h] = F.proposed[k];
        cur[h] = accept ? oldp[h] : F.current[k];
        sx0[h] = reset_stats ? 0.0 : F.stat_sx[k];
        sxx0[h] = reset_stats ? 0.0 : F.stat_sxx[k];
      }
    }
    fin_bar64();                                 // warp 0 has read current / proposed
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = h ? k1 : k0;
      if (k < np) {
        if (F.prop_trace) F.prop_trace[(unsigned long long)step * np + k] = oldp[h];
        if (accept) { F.current[k] = cur[h]; st->current[k] = cur[h]; }
        const float xf = (float)cur[h];
        if (store) out[k] = xf;
        const double x = (double)xf;
        F.stat_sx[k] = __dadd_rn(sx0[h], x);
        F.stat_sxx[k] = __dadd_rn(sxx0[h], __dmul_rn(x, x));
        st->stat_sx[k] = F.stat_sx[k];
        st->stat_sxx[k] = F.stat_sxx[k];
      }
    }
    if (lane == 0) {
      if (store) out[np] = (float)cur_nll;
      F.nrows = row + 1;
      F.stat_n = stat_n0 + 1;
      F.step = step + 1;
      st->nrows = F.nrows;
      st->stat_n = F.stat_n;
      st->step = F.step;
    }
  }
#undef FIN_CLK
}

}  // namespace sxmc
