/* oracle/sxmc_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, single-threaded CPU restatement of the sxmc per-step likelihood hot path,
 * written from the behaviour of the reference's CPU (hemi host-loop) build.  It is the
 * checker for the CUDA path in tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg; nothing under sxmc_b200/ may call it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against
 *   (1) the reference's own gtest golden vectors (test/test_pdfz.cpp:87-95,
 *       test/test_pdfz_2d.cpp:54-67, test/test_pdfz_syst.cpp:44-271, adapted as
 *       SURVEY.md section 4 describes) and
 *   (2) the unmodified reference sources compiled into oracle/_ref/libsxmc_ref.so
 *       (bins, norms, read_bins, lut bit-exact; NLL bit-exact; full chains row-exact
 *       when both are driven by the same random tape).
 *
 * Build: gcc -O2 -ffp-contract=off (x86-64 g++ -O2 emits no FMA for the reference
 * either), libm pow/log/exp/sqrt as the reference.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference).
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_MAX_NFIELDS 10 /* src/pdfz.cpp:17 */

/* systematic types, numeric values of pdfz::Systematic::Type (src/pdfz.h:111-116) */
enum { SYST_SHIFT = 0, SYST_SCALE = 1, SYST_RESOLUTION_SCALE = 2, SYST_CTSCALE = 3 };

/* A PDF's systematics as flat arrays (the reference's SystematicDescriptor,
 * src/pdfz.cpp:48-54): for systematic k, type[k], obs[k], extra[k], npars[k] and
 * npars[k] coefficient indices starting at pars[pars_start(k)]. */
typedef struct {
  int nsyst;
  const short* type;
  const short* obs;
  const short* extra;
  const short* npars;
  const short* pars; /* concatenated */
} oracle_systs;

/* src/pdfz.cpp:306-331 apply_systematic */
static void apply_systematic(short type, short obs, short extra, short npars,
                             const short* pars, double* fields,
                             const double* parameters, int param_stride) {
  double p = 0;
  short i;
  for (i = 0; i < npars; i++) {
    p += parameters[pars[i] * param_stride] * pow(fields[obs], (double)i);
  }
  switch (type) {
    case SYST_SHIFT: fields[obs] += p; break;
    case SYST_SCALE: fields[obs] *= (1 + p); break;
    case SYST_CTSCALE: fields[obs] = 1 + (fields[obs] - 1) * (1 + p); break;
    case SYST_RESOLUTION_SCALE: fields[obs] += (p * (fields[obs] - fields[extra])); break;
  }
}

/* src/pdfz.cpp:201-204: bin_volume starts as 1.0f and is multiplied in double. */
double oracle_bin_volume(int nobs, const double* lower, const double* upper,
                         const int* nbins) {
  double v = 1.0f;
  int i;
  for (i = 0; i < nobs; i++) v *= (upper[i] - lower[i]) / nbins[i];
  return v;
}

/* src/pdfz.cpp:206-213: row-major strides, last observable fastest; returns total_nbins. */
int oracle_bin_strides(int nobs, const int* nbins, int* stride) {
  int i;
  stride[nobs - 1] = 1;
  for (i = nobs - 2; i >= 0; i--) stride[i] = nbins[i + 1] * stride[i + 1];
  return stride[0] * nbins[0];
}

/* src/pdfz.cpp:334-346 zero_hist + :349-408 bin_samples (one "thread": offset 0, stride 1).
 * Samples whose flat index would fall outside [0,total_nbins) -- possible in the
 * reference only as undefined behaviour ("Reproduce quirks") -- are
 * counted in norm, as the reference does, but not written; *noob reports how many. */
void oracle_bin_samples(const float* data, long nsamples, int nfields, int nobs,
                        const double* lower, const double* upper, const int* nbins,
                        int nsyst, const short* stype, const short* sobs,
                        const short* sextra, const short* snpars, const short* spars,
                        const double* parameters, int param_stride,
                        unsigned* bins, unsigned* norm, long* noob) {
  double field_buffer[ORACLE_MAX_NFIELDS];
  double bin_scale[ORACLE_MAX_NFIELDS];
  int bin_stride[ORACLE_MAX_NFIELDS];
  int total_nbins = oracle_bin_strides(nobs, nbins, bin_stride);
  unsigned thread_norm = 0;
  long isample, oob = 0;
  int iobs, ifield, isyst;

  memset(bins, 0, sizeof(unsigned) * (size_t)total_nbins);
  *norm = 0;
  for (iobs = 0; iobs < nobs; iobs++)
    bin_scale[iobs] = nbins[iobs] / (upper[iobs] - lower[iobs]);

  for (isample = 0; isample < nsamples; isample++) {
    int in_pdf_domain = 1;
    int bin_id = 0;
    const short* pp = spars;
    for (ifield = 0; ifield < nfields; ifield++)
      field_buffer[ifield] = data[isample * nfields + ifield];
    for (isyst = 0; isyst < nsyst; isyst++) {
      apply_systematic(stype[isyst], sobs[isyst], sextra[isyst], snpars[isyst], pp,
                       field_buffer, parameters, param_stride);
      pp += snpars[isyst];
    }
    for (iobs = 0; iobs < nobs; iobs++) {
      double element = field_buffer[iobs];
      /* NaN: the reference's two compares are false and (int)NaN indexes out of bounds
       * (undefined; -ftrapv aborts).  Defined here, and in the CUDA path, as outside. */
      if (element != element || element < lower[iobs] || element >= upper[iobs]) { in_pdf_domain = 0; break; }
      bin_id += (int)((element - lower[iobs]) * bin_scale[iobs]) * bin_stride[iobs];
    }
    if (in_pdf_domain) {
      if (bin_id >= 0 && bin_id < total_nbins) bins[bin_id] += 1; else oob++;
      thread_norm += 1;
    }
  }
  *norm += thread_norm;
  if (noob) *noob = oob;
}

/* src/pdfz.cpp:245-302 EvalHist::SetEvalPoints.  Points are rows of nobs+1 floats, the
 * last being the dataset id.  -1 = outside the domain (wins), -2 = other dataset. */
void oracle_set_eval_points(const float* points, long npoints, int nobs,
                            const double* lower, const double* upper, const int* nbins,
                            unsigned dataset, int* read_bins) {
  double bin_scale[ORACLE_MAX_NFIELDS];
  int bin_stride[ORACLE_MAX_NFIELDS];
  long ipoint;
  int iobs;
  oracle_bin_strides(nobs, nbins, bin_stride);
  for (iobs = 0; iobs < nobs; iobs++) {
    double span = upper[iobs] - lower[iobs];
    bin_scale[iobs] = nbins[iobs] / span;
  }
  for (ipoint = 0; ipoint < npoints; ipoint++) {
    int in_pdf_domain = 1;
    int bin_id = 0;
    for (iobs = 0; iobs < nobs; iobs++) {
      double element = points[(nobs + 1) * ipoint + iobs];
      if (element != element || element < lower[iobs] || element >= upper[iobs]) { in_pdf_domain = 0; break; }
      bin_id += (int)((element - lower[iobs]) * bin_scale[iobs]) * bin_stride[iobs];
    }
    if (points[(nobs + 1) * ipoint + nobs] != dataset) bin_id = -2;
    read_bins[ipoint] = in_pdf_domain ? bin_id : -1;
  }
}

/* src/pdfz.cpp:411-436 eval_pdf.  Writes out[offset + stride*i] only. */
void oracle_eval_pdf(long npoints, const int* read_bins, const unsigned* bins,
                     unsigned norm, double bin_volume, float* out, int stride) {
  const double bin_norm = norm * bin_volume;
  long i;
  for (i = 0; i < npoints; i++) {
    int bin_id = read_bins[i];
    double pdf_value;
    if (bin_id == -2) pdf_value = 0.0;
    else if (bin_id < 0) pdf_value = nanf("");
    else pdf_value = bins[bin_id] / bin_norm;
    out[(long)stride * i] = pdf_value;
  }
}

/* src/nll_kernels.cpp:89-116 nll_event_chunks on one "thread".  `partial_io` is the
 * persistent event_partial_sums[0] slot: it is only overwritten when the new sum is
 * not NaN (:113-115). */
void oracle_nll_event_chunks(const float* lut, const double* pars, size_t ne, size_t ns,
                             const double* nexpected, const unsigned* n_mc,
                             const short* source_id, const unsigned* norms,
                             double* partial_io) {
  double sum = 0;
  size_t j;
  int i;
  for (i = 0; i < (int)ne; i++) {
    double s = 0;
    for (j = 0; j < ns; j++) {
      float v = lut[j * ne + i];
      float eff = 1.0 * norms[j] / n_mc[j];
      short sid = source_id[j];
      s += pars[sid] * nexpected[j] * eff * (!isnan(v) ? v : 0);
    }
    if (s > 0) sum += log(s);
  }
  if (!isnan(sum)) *partial_io = sum;
}

/* src/nll_kernels.cpp:150-188 nll_total_device */
double oracle_nll_total(size_t nparameters, size_t nsignals, size_t nsources,
                        const double* pars, const double* means, const double* sigmas,
                        double events_total, const double* nexpected,
                        const unsigned* n_mc, const short* source_id,
                        const unsigned* norms) {
  double sum = -events_total;
  unsigned i;
  if (isnan(sum)) return 1e18;
  for (i = 0; i < nsignals; i++) {
    short sid = source_id[i];
    sum += pars[sid] * nexpected[i] * norms[i] / n_mc[i];
  }
  for (i = 0; i < nparameters; i++) {
    if (i < nsources && pars[i] < 0) return 1e18;
    if (sigmas[i] > 0) {
      double x = (pars[i] - means[i]) / sigmas[i];
      sum += 0.5 * x * x;
    }
  }
  return sum;
}

/* MCMC::nll, src/mcmc.cpp:390-415 (chunks -> reduce over 1 partial -> total). */
double oracle_nll(const float* lut, size_t ne, size_t ns, const double* pars,
                  size_t npars, size_t nsources, const double* means,
                  const double* sigmas, const double* nexpected, const unsigned* n_mc,
                  const short* source_id, const unsigned* norms, double* evsum_out) {
  double partial = 0;
  oracle_nll_event_chunks(lut, pars, ne, ns, nexpected, n_mc, source_id, norms, &partial);
  if (evsum_out) *evsum_out = partial;
  return oracle_nll_total(npars, ns, nsources, pars, means, sigmas, partial, nexpected,
                          n_mc, source_id, norms);
}

/* ---- the chain: MCMC::MCMC (src/mcmc.cpp:28-130) + MCMC::operator() (:143-387) ---- */

/* One signal's PDF: an EvalHist (src/pdfz.cpp:177-236) with its systematics. */
typedef struct {
  const float* samples;
  long nsamples;
  int nfields;
  unsigned dataset;
  int nsyst;
  const short *stype, *sobs, *sextra, *snpars, *spars;
} oracle_signal;

typedef struct {
  const double* normals;  size_t n_normals,  i_normal;   /* standard-normal tape */
  const double* uniforms; size_t n_uniforms, i_uniform;  /* U(0,1) tape */
} oracle_tape;

static double tape_normal(oracle_tape* t) {
  return t->i_normal < t->n_normals ? t->normals[t->i_normal++] : 0.0;
}
static double tape_uniform(oracle_tape* t) {
  return t->i_uniform < t->n_uniforms ? t->uniforms[t->i_uniform++] : 0.5;
}

/* src/nll_kernels.cpp:31-53 pick_new_vector_device, CPU arm: gRandom->Gaus(mean, sigma)
 * = mean + sigma * z with z the next standard normal (ROOT TRandom::Gaus). */
static void pick_new_vector(int np, oracle_tape* t, const float* jump_width,
                            const double* current, double* proposed) {
  int i;
  for (i = 0; i < np; i++) {
    if (jump_width[i] > 0) proposed[i] = current[i] + jump_width[i] * tape_normal(t);
    else proposed[i] = current[i];
  }
}

/* Population RMS of column `col` over the float rows -- what ROOT's
 * TTree::Draw("x>>h") + TH1::GetRMS() yields (src/mcmc.cpp:288-293; unbinned
 * statistics fTsumw/fTsumwx/fTsumwx2 accumulated in double). */
static double column_rms(const float* rows, long nrows, int width, int col) {
  double n = 0, sx = 0, sxx = 0, m;
  long i;
  for (i = 0; i < nrows; i++) {
    double x = rows[i * width + col];
    n += 1; sx += x; sxx += x * x;
  }
  if (n <= 0) return 0;
  m = sx / n;
  return sqrt(fabs(sxx / n - m * m));
}

static void eval_all_pdfs(int nsignals, const oracle_signal* sig, int nobs,
                          const double* lower, const double* upper, const int* nbins,
                          const double* params_syst, unsigned** bins, unsigned* norms,
                          int** read_bins, long nevents, double bin_volume, float* lut) {
  int j;
  for (j = 0; j < nsignals; j++) {
    oracle_bin_samples(sig[j].samples, sig[j].nsamples, sig[j].nfields, nobs, lower, upper,
                       nbins, sig[j].nsyst, sig[j].stype, sig[j].sobs, sig[j].sextra,
                       sig[j].snpars, sig[j].spars, params_syst, 1, bins[j], &norms[j], 0);
    oracle_eval_pdf(nevents, read_bins[j], bins[j], norms[j], bin_volume,
                    lut + (size_t)j * nevents, 1);
  }
}

/* Returns the number of rows left in the chain "ntuple" (rows of nparameters+1 floats),
 * copied to chain_out (up to chain_cap rows).  If nll_trace != NULL it receives, per
 * step, the proposed-point NLL (nsteps doubles); if prop_trace != NULL the proposed
 * vectors (nsteps*nparameters doubles) -- used to replay proposals through the CUDA NLL.
 * accepted_out receives the total number of accepted jumps. */
long oracle_mcmc_run(int nsources, int nparameters, const double* means, const double* sigmas,
                     const int* fixed, int systematics_fixed,
                     int nobs, const double* lower, const double* upper, const int* nbins,
                     int nsignals, const oracle_signal* sig, const double* nexpected,
                     const unsigned* n_mc, const short* source_id,
                     const float* data, long nevents,
                     unsigned nsteps, float burnin_fraction, int debug_mode,
                     const double* normals, size_t n_normals,
                     const double* uniforms, size_t n_uniforms,
                     float* chain_out, long chain_cap, double* nll_trace,
                     double* prop_trace, long* accepted_out) {
  oracle_tape tape = { normals, n_normals, 0, uniforms, n_uniforms, 0 };
  const int np = nparameters, width = np + 1;
  unsigned burnin_steps = nsteps * burnin_fraction;  /* src/mcmc.cpp:152 (float product) */
  int nsystematics = np - nsources;                  /* only its being >0 matters (:264) */
  int nfloat = 0, total_nbins, strides[ORACLE_MAX_NFIELDS], j;
  unsigned i, k;
  double bin_volume = oracle_bin_volume(nobs, lower, upper, nbins);
  double* current = malloc(sizeof(double) * np);
  double* proposed = malloc(sizeof(double) * np);
  float* jump_width = malloc(sizeof(float) * np);
  unsigned* norms = calloc(nsignals, sizeof(unsigned));
  unsigned** bins = malloc(sizeof(unsigned*) * nsignals);
  int** read_bins = malloc(sizeof(int*) * nsignals);
  float* lut = malloc(sizeof(float) * (size_t)nsignals * (nevents > 0 ? nevents : 1));
  float* nt = malloc(sizeof(float) * (size_t)width * (nsteps > 0 ? nsteps : 1));
  long nt_rows = 0, accepted = 0;
  double current_nll, proposed_nll, partial = 0;
  float scale_factor;

  total_nbins = oracle_bin_strides(nobs, nbins, strides);
  for (j = 0; j < np; j++) { current[j] = means[j]; nfloat += fixed[j] ? 0 : 1; }

  /* initial jump widths, src/mcmc.cpp:199-228 */
  scale_factor = 2.4 * 2.4 / nfloat;
  for (j = 0; j < np; j++) {
    float mean, sigma, width_j = 0.1;
    if (fixed[j]) { jump_width[j] = -1; continue; }
    mean = means[j]; sigma = sigmas[j];
    if (sigma > 0) width_j = sigma;
    else if (j < nsignals) { float m = mean > 10.f ? mean : 10.f; width_j = sqrtf(m) / m; }
    else width_j = sqrtf(mean > 1.f ? mean : 1.f);
    jump_width[j] = 0.1 * width_j * scale_factor;
  }

  /* initial evaluation at the means with current_vector, :231-250 */
  for (j = 0; j < nsignals; j++) {
    bins[j] = malloc(sizeof(unsigned) * total_nbins);
    read_bins[j] = malloc(sizeof(int) * (nevents > 0 ? nevents : 1));
    oracle_set_eval_points(data, nevents, nobs, lower, upper, nbins, sig[j].dataset,
                           read_bins[j]);
  }
  eval_all_pdfs(nsignals, sig, nobs, lower, upper, nbins, current + nsources, bins, norms,
                read_bins, nevents, bin_volume, lut);
  oracle_nll_event_chunks(lut, current, nevents, nsignals, nexpected, n_mc, source_id,
                          norms, &partial);
  current_nll = oracle_nll_total(np, nsignals, nsources, current, means, sigmas, partial,
                                 nexpected, n_mc, source_id, norms);
  pick_new_vector(np, &tape, jump_width, current, proposed);  /* :252 */

  for (i = 0; i < nsteps; i++) {
    if (nsystematics > 0 && !systematics_fixed)  /* :264-271 */
      eval_all_pdfs(nsignals, sig, nobs, lower, upper, nbins, proposed + nsources, bins,
                    norms, read_bins, nevents, bin_volume, lut);

    if (i == burnin_steps || i == 2 * burnin_steps) {  /* :274-311 */
      for (j = 0; j < np; j++) {
        double rms, fit_width;
        if (fixed[j]) continue;
        rms = column_rms(nt, nt_rows, width, j);
        fit_width = rms > 0 ? rms : jump_width[j];
        jump_width[j] = scale_factor * fit_width;
      }
      if (!debug_mode) nt_rows = 0;
    }

    /* :314-348: event term, total, Metropolis, record, propose */
    oracle_nll_event_chunks(lut, proposed, nevents, nsignals, nexpected, n_mc, source_id,
                            norms, &partial);
    proposed_nll = oracle_nll_total(np, nsignals, nsources, proposed, means, sigmas,
                                    partial, nexpected, n_mc, source_id, norms);
    if (nll_trace) nll_trace[i] = proposed_nll;
    if (prop_trace) memcpy(prop_trace + (size_t)i * np, proposed, sizeof(double) * np);
    {  /* src/nll_kernels.cpp:57-86 jump_decider_device */
      double u = tape_uniform(&tape);
      if (debug_mode || (proposed_nll < current_nll || u <= exp(current_nll - proposed_nll))) {
        current_nll = proposed_nll;
        for (k = 0; k < (unsigned)np; k++) current[k] = proposed[k];
        accepted += 1;
      }
      for (k = 0; k < (unsigned)np; k++) nt[nt_rows * width + k] = current[k];
      nt[nt_rows * width + np] = current_nll;
      nt_rows++;
    }
    pick_new_vector(np, &tape, jump_width, current, proposed);
  }

  if (chain_out) {
    long ncopy = nt_rows < chain_cap ? nt_rows : chain_cap;
    memcpy(chain_out, nt, sizeof(float) * (size_t)width * ncopy);
  }
  if (accepted_out) *accepted_out = accepted;
  for (j = 0; j < nsignals; j++) { free(bins[j]); free(read_bins[j]); }
  free(bins); free(read_bins); free(lut); free(nt); free(norms);
  free(current); free(proposed); free(jump_width);
  return nt_rows;
}
