// nll_kernels.cu -- PDF lookup + event likelihood + NLL total + Metropolis + proposal as ONE
// launch per step.  Replaces eval_pdf (reference src/pdfz.cpp:411-436), nll_event_chunks
// (src/nll_kernels.cpp:89-116), nll_event_reduce (:120-146), nll_total (:150-188),
// jump_decider (:57-86), pick_new_vector (:31-53) and the host-side burn-in re-tune of
// src/mcmc.cpp:274-311.  RNG: Philox4x32-10 keyed by (seed, chain stream), counter =
// (step, parameter, purpose) -- replaces init_device_rngs / XORWOW (:19-26).
#include "common.cuh"
#include "kernels.h"
#include "nll_device.cuh"

namespace sxmc {

constexpr int NLL_THREADS = 256;
constexpr int NLL_MAX_SIGNALS = 128;

// =========================================================================================
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
lookup_nll_kernel(const NllSignal* __restrict__ sigs, const unsigned* __restrict__ norms, int ns,
                  int np, int nsources, long long ne, ChainState* st, double* partials,
                  unsigned* ticket, int mode, float* lut, unsigned* zero_ptr,
                  unsigned long long zero_words, const unsigned* __restrict__ mult, int mult_nbins,
                  const double* __restrict__ ptable, unsigned long long* xdbg, int zero_early) {
  __shared__ double s_c[NLL_MAX_SIGNALS];
  __shared__ double s_bn[NLL_MAX_SIGNALS];
  __shared__ const unsigned* s_bins[NLL_MAX_SIGNALS];
  __shared__ const int* s_rb[NLL_MAX_SIGNALS];
  __shared__ int s_nb[NLL_MAX_SIGNALS];
  __shared__ double s_warp[THREADS / 32];
  __shared__ double s_term[NLL_MAX_SIGNALS + SXMC_MAX_PARAMETERS];
  __shared__ int s_last;
  __shared__ FinState F;
  __shared__ double s_z[SXMC_MAX_PARAMETERS];
  __shared__ double s_u;

  const int tid = threadIdx.x;
  const long long gtid = blockIdx.x * (long long)THREADS + tid;
  const long long gstride = (long long)gridDim.x * THREADS;
  // profiling hook (sxmc_nll_debug_stamps): [block < 1200][8] %globaltimer stamps, the block that
  // finishes the step also writes row 1200
#define NLL_DBG(k)                                                                         \
  do {                                                                                     \
    if (xdbg && tid == 0) {                                                                \
      unsigned long long t_;                                                               \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                               \
      xdbg[8ull * (blockIdx.x < 1200u ? blockIdx.x : 1199u) + (k)] = t_;                   \
    }                                                                                      \
  } while (0)
  NLL_DBG(0);

  // ---- prologue that does not depend on this step's histograms: under programmatic dependent
  // launch it runs while the binning kernel is still streaming.  The chain state was written by the
  // previous step's launch, which completed before the binning kernel let this one start.
  prefetch_state(F, st, np, tid, THREADS);
  NllSignal my_sig;
  unsigned my_norm = 0;
  if (tid < ns) my_sig = sigs[tid];
  // clear the histogram buffer the NEXT step bins into (double-buffered; nobody reads it now --
  // unless peers may still be summing it: then only after the wait)
  if (zero_early)
    for (unsigned long long w = gtid; w < zero_words; w += gstride) zero_ptr[w] = 0u;
  pdl_wait();
  pdl_trigger();
  if (!zero_early)
    for (unsigned long long w = gtid; w < zero_words; w += gstride) zero_ptr[w] = 0u;
  if (tid < ns) my_norm = norms[tid];
  __syncthreads();
  NLL_DBG(1);

  // block 0's last warp draws the step's random numbers while everybody looks up events; the
  // block that finishes last picks them up (ordered by block 0's ticket, as its partial sum is)
  const bool pre_rng = pre_rng_applies(F, mode);
  if (pre_rng && blockIdx.x == 0 && tid >= THREADS - 32) pre_rng_draw(F, np, tid & 31, st->pre_z, &st->pre_u);

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
  for (int i = tid; i < np; i += THREADS) {
    const double sg = F.sigmas[i];
    double t = 0;
    if (sg > 0) {
      const double x = __dadd_rn(F.proposed[i], -F.means[i]) / sg;
      t = __dmul_rn(__dmul_rn(0.5, x), x);
    }
    s_term[ns + i] = t;
  }
  __syncthreads();
  NLL_DBG(2);

  double sum = 0;
  if (mult && !lut) {
    // Grouped lookup: every signal reads the same dataset, so the events of one bin share a
    // likelihood and sum_i log(s_i) = sum_b m_b log(s_b), m_b = events in bin b (counted once when
    // the data were set).  One division per (signal, occupied bin) instead of one per
    // (signal, event); no gather.
    for (long long b = gtid; b < mult_nbins; b += gstride) {
      const unsigned m = mult[b];
      double s = 0;
      // up to 32 signals' counts in flight at once, issued together with the multiplicity: ONE L2
      // round trip per bin instead of one per group of signals (the loads do not depend on the
      // running sum; a rolled loop issues them one after the other).  Same order of additions.
      // (width 32 in the many-bin launch of 128-thread blocks; 8 elsewhere, where registers are short)
      constexpr int LW = THREADS == 128 ? 32 : 8;
      for (int j0 = 0; j0 < ns; j0 += LW) {
        unsigned h[LW];
#pragma unroll
        for (int q = 0; q < LW; ++q) h[q] = (j0 + q < ns) ? s_bins[j0 + q][b] : 0u;
        if (!m) break;
#pragma unroll
        for (int q = 0; q < LW; ++q) {
          if (j0 + q < ns) {
            float v = pdf_value(h[q], s_bn[j0 + q]);                         // pdfz.cpp:429-432
            if (isnan(v)) v = 0.0f;                                   // nll_kernels.cpp:107
            s = __dadd_rn(s, __dmul_rn(s_c[j0 + q], (double)v));
          }
        }
      }
      if (m && s > 0) sum += (double)m * log(s);
    }
  } else if (ptable && !lut) {
    // Per-event lookup against the step's PDF-value table (pdf_table_kernel: eval_pdf's division
    // done once per (signal, bin), NaN already mapped to 0): 4 B of read_bins per (event, signal)
    // from HBM, one gather from the L2-resident table (values widened to double once), two FP64 operations.  Four events per
    // thread (one 128-bit load per signal), four signals' loads in flight; the additions keep the
    // reference's order within an event.
    const int nb = mult_nbins;
    const long long ne4 = ne & ~3ll;
    for (long long i = gtid * 4; i < ne4; i += gstride * 4) {
      double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
      int j = 0;
      for (; j + 4 <= ns; j += 4) {
        int4 rb[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) rb[q] = __ldcs(reinterpret_cast<const int4*>(s_rb[j + q] + i));
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const double* t = ptable + (size_t)(j + q) * nb;
          const double c = s_c[j + q];
          const double v0 = (unsigned)rb[q].x < (unsigned)nb ? __ldg(t + rb[q].x) : 0.0;
          const double v1 = (unsigned)rb[q].y < (unsigned)nb ? __ldg(t + rb[q].y) : 0.0;
          const double v2 = (unsigned)rb[q].z < (unsigned)nb ? __ldg(t + rb[q].z) : 0.0;
          const double v3 = (unsigned)rb[q].w < (unsigned)nb ? __ldg(t + rb[q].w) : 0.0;
          s0 = __dadd_rn(s0, __dmul_rn(c, v0));
          s1 = __dadd_rn(s1, __dmul_rn(c, v1));
          s2 = __dadd_rn(s2, __dmul_rn(c, v2));
          s3 = __dadd_rn(s3, __dmul_rn(c, v3));
        }
      }
      for (; j < ns; ++j) {
        const int4 rb = __ldcs(reinterpret_cast<const int4*>(s_rb[j] + i));
        const double* t = ptable + (size_t)j * nb;
        const double c = s_c[j];
        const double v0 = (unsigned)rb.x < (unsigned)nb ? __ldg(t + rb.x) : 0.0;
        const double v1 = (unsigned)rb.y < (unsigned)nb ? __ldg(t + rb.y) : 0.0;
        const double v2 = (unsigned)rb.z < (unsigned)nb ? __ldg(t + rb.z) : 0.0;
        const double v3 = (unsigned)rb.w < (unsigned)nb ? __ldg(t + rb.w) : 0.0;
        s0 = __dadd_rn(s0, __dmul_rn(c, v0));
        s1 = __dadd_rn(s1, __dmul_rn(c, v1));
        s2 = __dadd_rn(s2, __dmul_rn(c, v2));
        s3 = __dadd_rn(s3, __dmul_rn(c, v3));
      }
      if (s0 > 0) sum += log(s0);
      if (s1 > 0) sum += log(s1);
      if (s2 > 0) sum += log(s2);
      if (s3 > 0) sum += log(s3);
    }
    for (long long i = ne4 + gtid; i < ne; i += gstride) {
      double s = 0;
      for (int j = 0; j < ns; ++j) {
        const int rb = s_rb[j][i];
        const double v = (unsigned)rb < (unsigned)nb ? __ldg(ptable + (size_t)j * nb + rb) : 0.0;
        s = __dadd_rn(s, __dmul_rn(s_c[j], v));
      }
      if (s > 0) sum += log(s);
    }
  } else
  for (long long i = gtid; i < ne; i += gstride) {
    double s = 0;
    for (int j = 0; j < ns; ++j) {
      const int rb = s_rb[j][i];
      float v = 0.0f;
      float vlut = 0.0f;
      if (rb >= s_nb[j]) {
        // one-past-the-end index of a point within an ulp of `upper`: undefined in the reference
      } else if (rb >= 0) {
        v = pdf_value(s_bins[j][rb], s_bn[j]);                      // pdfz.cpp:429-432
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
  NLL_DBG(3);
  const double bsum = block_sum<THREADS>(sum, s_warp);
  NLL_DBG(4);
  if (gridDim.x == 1) {
    // small fit, one block (launch_nll): no partials, no ticket round trip, no second reduction
    if (tid == 0) s_warp[0] = bsum;
    __syncthreads();                 // also orders the last warp's pre_z / pre_u stores before the loads below
    if (pre_rng) {
      for (int k = tid; k < np; k += THREADS) s_z[k] = __ldcg(st->pre_z + k);
      if (tid == 0) s_u = __ldcg(&st->pre_u);
    }
  } else {
    if (tid == 0) {
      partials[blockIdx.x] = bsum;
      __threadfence();
      const unsigned t = atomicAdd(ticket, 1u);
      s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    NLL_DBG(5);
    if (!s_last) return;

    // ---- last block: total the partials in a fixed order, then finish the step -------------
    __threadfence();
    double acc = 0;
    for (int b = tid; b < (int)gridDim.x; b += THREADS) acc += __ldcg(partials + b);
    const double total = block_sum<THREADS>(acc, s_warp);
    if (tid == 0) { s_warp[0] = total; *ticket = 0u; }
    if (pre_rng) {
      for (int k = tid; k < np; k += THREADS) s_z[k] = __ldcg(st->pre_z + k);
      if (tid == 0) s_u = __ldcg(&st->pre_u);
    }
  }
  __syncthreads();
  NLL_DBG(6);
  if (tid < 64)
    finish_step2(F, st, np, ns, nsources, s_warp[0], mode, s_term, tid & 31, tid >> 5, pre_rng ? s_z : nullptr,
                pre_rng ? &s_u : nullptr);
  if (xdbg && tid == 0) {
    unsigned long long t_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
    xdbg[8ull * 1200 + 7] = t_;
    for (int k = 0; k < 7; ++k) xdbg[8ull * 1200 + k] = xdbg[8ull * (blockIdx.x < 1200u ? blockIdx.x : 1199u) + k];
    xdbg[8ull * 1201] = blockIdx.x;
  }
#undef NLL_DBG
}

// eval_pdf's value (src/pdfz.cpp:429-432) for every (signal, bin) of this step, with the NaN of
// an empty PDF already replaced by 0 as nll_event_chunks does (src/nll_kernels.cpp:107): the
// per-event lookup then costs a gather instead of a double-precision division per (event, signal).
__global__ void pdf_table_kernel(const NllSignal* __restrict__ sigs, const unsigned* __restrict__ norms, int nb,
                                 double* __restrict__ table) {
  const int j = blockIdx.y;
  const double bn = __dmul_rn((double)norms[j], sigs[j].bin_volume);        // pdfz.cpp:417
  const unsigned* bins = sigs[j].bins;
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += gridDim.x * blockDim.x) {
    float v = pdf_value(bins[b], bn);
    if (isnan(v)) v = 0.0f;
    table[(size_t)j * nb + b] = (double)v;      // widened here once: no conversion per (event, signal)
  }
}

int launch_pdf_table(const NllSignal* sigs_dev, const unsigned* norms_dev, int nsignals, int nbins, double* table_dev,
                     cudaStream_t stream) {
  if (nsignals <= 0 || nbins <= 0) return SXMC_OK;
  int gx = (nbins + 255) / 256;
  if (gx > 64) gx = 64;
  pdf_table_kernel<<<dim3((unsigned)gx, (unsigned)nsignals), 256, 0, stream>>>(sigs_dev, norms_dev, nbins, table_dev);
  SXMC_CUDA(cudaGetLastError());
  return SXMC_OK;
}

static unsigned long long* g_nll_xdbg = nullptr;   // profiling hook, NULL in production
void set_nll_xdbg(unsigned long long* dev) { g_nll_xdbg = dev; }

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
  // Grouped lookup over a few thousand bins: one block does it all, which drops the partial-sum
  // hand-over (store, fence, ticket, last block's reload: ~3 us of dependent round trips) from a
  // kernel that is latency from end to end.  1024 threads: one bin per thread for workload B.
  if (L.mult && !L.lut && L.mult_nbins <= 4096 && L.zero_words <= 65536) {
    SXMC_CUDA(launch_kernel_pdl(lookup_nll_kernel<1024>, dim3(1), dim3(1024), 0, stream, L.pdl, L.sigs_dev, L.norms_dev, L.nsignals, L.nparameters, L.nsources,
                                                    L.nevents, L.state, L.partials, L.ticket, L.mode, L.lut,
                                                    L.zero_ptr, L.zero_words, L.mult, L.mult_nbins, L.ptable, g_nll_xdbg, L.zero_early ? 1 : 0));
  } else if (L.mult && !L.lut) {
    // grouped lookup over many bins: one bin per thread, 128-thread blocks so that every SM gets
    // one (25 000 bins: 196 blocks), no block without work
    int grid = (L.mult_nbins + 127) / 128;
    if (grid > 148 * 8) grid = 148 * 8;
    SXMC_CUDA(launch_kernel_pdl(lookup_nll_kernel<128>, dim3((unsigned)grid), dim3(128), 0, stream, L.pdl, L.sigs_dev, L.norms_dev, L.nsignals, L.nparameters, L.nsources,
                                                     L.nevents, L.state, L.partials, L.ticket, L.mode, L.lut,
                                                     L.zero_ptr, L.zero_words, L.mult, L.mult_nbins, L.ptable, g_nll_xdbg, L.zero_early ? 1 : 0));
  } else {
    SXMC_CUDA(launch_kernel_pdl(lookup_nll_kernel<NLL_THREADS>, dim3((unsigned)L.grid), dim3(NLL_THREADS), 0, stream, L.pdl,
        L.sigs_dev, L.norms_dev, L.nsignals, L.nparameters, L.nsources, L.nevents, L.state, L.partials, L.ticket,
        L.mode, L.lut, L.zero_ptr, L.zero_words, L.mult, L.mult_nbins, L.ptable, g_nll_xdbg, L.zero_early ? 1 : 0));
  }
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
