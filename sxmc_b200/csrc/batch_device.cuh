// batch_device.cuh -- the lookup + tail phase shared by the chain batch kernels
// (batch_kernels.cu, batch_mono_kernels.cu): CTA e owns chain e of the batch.
#pragma once
#include "common.cuh"
#include "nll_device.cuh"
#include "grid_sync.cuh"

// threads per CTA of the including kernel (every thread takes part in the lookup + tail)
#ifndef BATCH_THREADS
#define BATCH_THREADS FAST_THREADS
#endif

namespace sxmc {

constexpr int BK_MAX_SIGNALS = 128;

struct BatchShared {
  double c[BK_MAX_SIGNALS];        // rate * efficiency of every signal at the proposed point
  double bn[BK_MAX_SIGNALS];       // norm * bin_volume
  double nexp[BK_MAX_SIGNALS];
  double binvol[BK_MAX_SIGNALS];
  const int* rb[BK_MAX_SIGNALS];
  unsigned nmc[BK_MAX_SIGNALS];
  short sid[BK_MAX_SIGNALS];
  double warp[BATCH_THREADS / 32];
  double term[BK_MAX_SIGNALS + SXMC_MAX_PARAMETERS];
  FinState F;
  int ok;
  long long ne;
  double pre_z[SXMC_MAX_PARAMETERS];   // the step's normals / uniform, drawn ahead of the serial tail
  double pre_u;
  float pdf[BATCH_PDF_TABLE];      // PDF values of every (signal, bin) at the proposed point (grouped lookup)
};

// once per launch: the owned chain's state and signal tables into shared memory
__device__ __forceinline__ void batch_owner_init(BatchShared& S, const BatchParams& Q, int mine, int tid) {
  if (mine < 0) return;
  prefetch_state(S.F, Q.st + mine, Q.np, tid, BATCH_THREADS);
  if (tid < Q.ns) {
    const NllSignal sg = Q.sigs[(size_t)mine * Q.ns + tid];
    S.nexp[tid] = sg.nexpected;
    S.binvol[tid] = sg.bin_volume;
    S.rb[tid] = sg.read_bins;
    S.nmc[tid] = sg.n_mc;
    S.sid[tid] = sg.source_id;
  }
  if (tid == 0) S.ne = Q.nevents[mine];
}

// One step of the owned chain after its histograms are complete: eval_pdf + nll_event_chunks +
// reduce (src/pdfz.cpp:411-436, src/nll_kernels.cpp:89-146), then finish_step; the histogram is
// cleared for the next super-step's bin phase.  All threads of the CTA call it.
__device__ __forceinline__ void batch_lookup_tail(BatchShared& S, const BatchParams& Q, int mine, int mode,
                                                  int tnb, int tid, bool clear,
                                                  unsigned long long* stamp = nullptr,
                                                  bool have_m0 = false, unsigned m0 = 0u) {
  const int ns = Q.ns, np = Q.np;
  unsigned* const my_hist = Q.hist + (size_t)mine * Q.hist_words;
  FinState& F = S.F;
  // the last warp draws the step's random numbers while the others look up the events
  const bool pre_rng = pre_rng_applies(F, mode);
  if (pre_rng && tid >= BATCH_THREADS - 32) pre_rng_draw(F, np, tid & 31, S.pre_z, &S.pre_u);
  // Small grouped fits (one bin per thread, a handful of signals): the bin's counts go straight
  // to registers, and the loads are issued before the norm-dependent stage instead of after it
  // (one L2 round trip and the shared-memory table less on a path that is latency end to end).
  constexpr int DIRECT_NS = 4;
  const bool direct = Q.mult != nullptr && tnb <= BATCH_THREADS && ns <= DIRECT_NS;
  unsigned hreg[DIRECT_NS];
  if (direct && tid < tnb) {
#pragma unroll
    for (int j = 0; j < DIRECT_NS; ++j) hreg[j] = j < ns ? __ldcg(my_hist + (size_t)j * tnb + tid) : 0u;
  }
  if (tid < ns) {
    const unsigned norm = __ldcg(my_hist + (size_t)ns * tnb + tid);
    const float eff = (float)(1.0 * norm / S.nmc[tid]);                    // nll_kernels.cpp:105
    const double rate = __dmul_rn(F.proposed[S.sid[tid]], S.nexp[tid]);
    S.c[tid] = __dmul_rn(rate, (double)eff);
    S.bn[tid] = __dmul_rn((double)norm, S.binvol[tid]);                    // pdfz.cpp:417
    S.term[tid] = __dmul_rn(rate, (double)norm) / (double)S.nmc[tid];      // nll_kernels.cpp:167-171
  }
  for (int i = tid; i < np; i += BATCH_THREADS) {       // Gaussian constraints (nll_kernels.cpp:176-185)
    const double sg = F.sigmas[i];
    double t = 0;
    if (sg > 0) {
      const double x = __dadd_rn(F.proposed[i], -F.means[i]) / sg;
      t = __dmul_rn(__dmul_rn(0.5, x), x);
    }
    S.term[ns + i] = t;
  }
  __syncthreads();
  const long long my_ne = S.ne;
  double sum = 0;
  if (direct) {
    if (tid < tnb) {
      const unsigned m = have_m0 ? m0 : Q.mult[(size_t)mine * tnb + tid];
      if (m) {
        double acc = 0;
#pragma unroll
        for (int j = 0; j < DIRECT_NS; ++j) {
          if (j < ns) {
            float v = pdf_value(hreg[j], S.bn[j]);                              // pdfz.cpp:429-432
            if (isnan(v)) v = 0.0f;                                              // nll_kernels.cpp:107
            acc = __dadd_rn(acc, __dmul_rn(S.c[j], (double)v));
          }
        }
        if (acc > 0) sum += (double)m * log(acc);
      }
    }
    if (stamp && tid == 0) stamp[3] = globaltimer_ns();
  } else if (Q.mult) {
    // Grouped lookup.  Every signal reads the same dataset, so events in the same bin have the
    // same likelihood: sum_i log(s_i) = sum_b m_b log(s_b), m_b = events of this chain in bin b
    // (counted once when the data were set).  eval_pdf's division runs once per (signal, bin)
    // instead of once per (signal, event), and the values live in shared memory.
    for (int q = tid; q < ns * tnb; q += BATCH_THREADS) {
      const int j = q / tnb;
      float v = pdf_value(__ldcg(my_hist + q), S.bn[j]);                     // pdfz.cpp:429-432
      if (isnan(v)) v = 0.0f;                                                  // nll_kernels.cpp:107
      S.pdf[q] = v;
    }
    __syncthreads();
    if (stamp && tid == 0) stamp[3] = globaltimer_ns();
    const unsigned* mult = Q.mult + (size_t)mine * tnb;
    for (int b = tid; b < tnb; b += BATCH_THREADS) {
      const unsigned m = (have_m0 && b == tid) ? m0 : mult[b];   // m0: the caller's register copy of mult[tid]
      if (!m) continue;
      double acc = 0;
      for (int j = 0; j < ns; ++j) acc = __dadd_rn(acc, __dmul_rn(S.c[j], (double)S.pdf[j * tnb + b]));
      if (acc > 0) sum += (double)m * log(acc);
    }
  } else
  for (long long i = tid; i < my_ne; i += BATCH_THREADS) {
    double acc = 0;
    for (int j = 0; j < ns; ++j) {
      const int rb = S.rb[j][i];
      float v = 0.0f;
      if (rb >= 0 && rb < tnb) {
        v = pdf_value(__ldcg(my_hist + (size_t)j * tnb + rb), S.bn[j]);          // pdfz.cpp:429-432
        if (isnan(v)) v = 0.0f;                                                  // nll_kernels.cpp:107
      }
      acc = __dadd_rn(acc, __dmul_rn(S.c[j], (double)v));
    }
    if (acc > 0) sum += log(acc);
  }
  const double total = block_sum<BATCH_THREADS>(sum, S.warp);   // ends past a CTA barrier: lookups done
  if (tid == 0) S.warp[0] = total;
  if (stamp && tid == 0) stamp[6] = globaltimer_ns();
  if (clear)     // rebuilt from zero by the next super-step's bin phase (the last one is kept)
    for (unsigned long long w = tid; w < Q.hist_words; w += BATCH_THREADS) my_hist[w] = 0u;
  __syncthreads();
  if (stamp && tid == 0) stamp[7] = globaltimer_ns();
  if (tid < 64)
    finish_step2(F, Q.st + mine, np, ns, Q.nsources, S.warp[0], mode, S.term, tid & 31, tid >> 5,
                pre_rng ? S.pre_z : nullptr, pre_rng ? &S.pre_u : nullptr,
                (stamp && mine == 0) ? reinterpret_cast<long long*>(stamp + BATCH_MAX_CHAINS * 8) : nullptr);
  __syncthreads();
}

}  // namespace sxmc
