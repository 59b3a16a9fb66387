// oracle/ref_glue.cpp -- TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// A ctypes-loadable extern "C" face over the UNMODIFIED reference sources
// (/root/reference/src/{pdfz,nll_kernels,mcmc,...}.cpp, compiled where they lie by
// oracle/Makefile into oracle/_ref/libsxmc_ref.so with the reference's CPU flags:
// Makefile:20,31 -- g++ -O2 -DHEMI_CUDA_DISABLE, hemi host loops).
// Nothing here restates an algorithm: it only builds the reference's own objects,
// calls the reference's own functions and copies their outputs out.
//
// ROOT is absent; the few ROOT names the three hot TUs mention come from
// oracle/root_stubs/*.h.  gRandom is a "tape" generator (root_stubs/TRandom.h) so the
// reference chain can be driven by exactly the random numbers the CUDA chain draws.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include <TRandom.h>
#include <TDirectory.h>
#include <TNtuple.h>

#include <sxmc/pdfz.h>
#include <sxmc/nll_kernels.h>
#include <sxmc/mcmc.h>
#include <sxmc/signal.h>
#include <sxmc/source.h>
#include <sxmc/systematic.h>
#include <sxmc/observable.h>
#include <sxmc/likelihood.h>
#include <sxmc/utils.h>
#include <sxmc/ttree_io.h>
#include <sxmc/config.h>

// ---- symbols the reference expects from ROOT / from TUs that need real ROOT ----
static TRandom g_random_instance;
TRandom* gRandom = &g_random_instance;
static TDirectory g_directory_instance;
TDirectory* gDirectory = &g_directory_instance;

// src/utils.cpp:14 (TMath::Nint clone) -- utils.cpp itself needs real ROOT.
unsigned nint(float nexpected) { return (unsigned) nearbyintf(nexpected); }  // TMath::Nint: half-to-even

// src/likelihood.cpp needs TTree::Draw etc.; the chain sink only has to keep the ntuple.
LikelihoodSpace::LikelihoodSpace(TNtuple* _samples, float _cl, ErrorType _error_type)
    : error_type(_error_type), nll(0), cl(_cl), samples(_samples) {}
LikelihoodSpace::~LikelihoodSpace() { delete samples; }

// src/io/ttree_io.cpp needs ROOT TFile/TTree.  Same signature and contract
// (src/io/ttree_io.cpp:21-159: row-major float table, rank = {rows, columns}, column names), read
// from the flat-table form this repo's command lines accept: `<file>.npy` holding a C-order
// [n, k] '<f4' array next to a `<file>.npy.fields` text file with one column name per line.
// Lets the reference's own FitConfig / Signal constructors run on the test configurations.
namespace sxmc { namespace io {
int read_float_vector_ttree(const std::string& filename, std::vector<float>& data,
                            std::vector<unsigned int>& rank, std::vector<std::string>& ttree_fields) {
  data.clear(); rank.clear(); ttree_fields.clear();
  std::ifstream ff((filename + ".fields").c_str());
  std::string line;
  while (std::getline(ff, line)) {
    while (!line.empty() && (line[line.size() - 1] == '\r' || line[line.size() - 1] == ' ')) line.erase(line.size() - 1);
    if (!line.empty()) ttree_fields.push_back(line);
  }
  std::ifstream f(filename.c_str(), std::ios::binary);
  char magic[10];
  if (!f.read(magic, 10) || std::memcmp(magic, "\x93NUMPY", 6) != 0) return -1;
  size_t hlen = (unsigned char)magic[8] | ((size_t)(unsigned char)magic[9] << 8);
  if (magic[6] != 1) {   // version 2/3: 4-byte header length
    char more[2];
    if (!f.read(more, 2)) return -1;
    hlen |= ((size_t)(unsigned char)more[0] << 16) | ((size_t)(unsigned char)more[1] << 24);
  }
  std::string header(hlen, ' ');
  if (!f.read(&header[0], hlen)) return -1;
  if (header.find("'<f4'") == std::string::npos || header.find("'fortran_order': False") == std::string::npos) return -2;
  size_t a = header.find("'shape': ("), b = header.find(")", a);
  if (a == std::string::npos || b == std::string::npos) return -2;
  unsigned long rows = 0, cols = 0;
  if (std::sscanf(header.substr(a + 10, b - a - 10).c_str(), "%lu, %lu", &rows, &cols) != 2) return -2;
  if (cols != ttree_fields.size()) return -3;
  data.resize((size_t)rows * cols);
  if (!data.empty() && !f.read((char*)&data[0], sizeof(float) * data.size())) return -4;
  rank.push_back((unsigned)rows);
  rank.push_back((unsigned)cols);
  return 0;
}
} }

namespace {

// Exposes the protected state of the reference's EvalHist (src/pdfz.h:558-573).
class EvalHistProbe : public pdfz::EvalHist {
public:
  EvalHistProbe(const std::vector<float>& s, int nfields, int nobs,
                const std::vector<double>& lo, const std::vector<double>& hi,
                const std::vector<int>& nb, unsigned dataset)
      : pdfz::EvalHist(s, nfields, nobs, lo, hi, nb, dataset, false) {}
  int TotalBins() const { return total_nbins; }
  const unsigned* Bins() { return bins->readOnlyHostPtr(); }
  const int* ReadBins() { return read_bins ? read_bins->readOnlyHostPtr() : 0; }
  size_t NReadBins() { return read_bins ? read_bins->size() : 0; }
  double BinVolume() const { return bin_volume; }
  // the sample rows of ANY EvalHist (protected member read through this subclass)
  static std::vector<float> Rows(pdfz::EvalHist* h) {
    EvalHistProbe* p = static_cast<EvalHistProbe*>(h);
    const float* s = p->samples.readOnlyHostPtr();
    return std::vector<float>(s, s + p->samples.size());
  }
};

void set_err(char* err, int errlen, const std::string& msg) {
  if (err && errlen > 0) {
    std::strncpy(err, msg.c_str(), errlen - 1);
    err[errlen - 1] = 0;
  }
}

struct CoutSilencer {
  std::streambuf* old;
  CoutSilencer(bool on) : old(0) { if (on) old = std::cout.rdbuf(0); }
  ~CoutSilencer() { if (old) std::cout.rdbuf(old); }
};

}  // namespace

extern "C" {

void* ref_hist_create(const float* samples, size_t nfloats, int nfields, int nobs,
                      const double* lower, int nlower, const double* upper, int nupper,
                      const int* nbins, int nnbins, unsigned dataset,
                      char* err, int errlen) {
  try {
    std::vector<float> s(samples, samples + nfloats);
    std::vector<double> lo(lower, lower + nlower), hi(upper, upper + nupper);
    std::vector<int> nb(nbins, nbins + nnbins);
    return new EvalHistProbe(s, nfields, nobs, lo, hi, nb, dataset);
  } catch (pdfz::Error& e) {
    set_err(err, errlen, e.msg);
    return 0;
  }
}

void ref_hist_destroy(void* h) { delete (EvalHistProbe*)h; }

// type uses the reference's enum values (src/pdfz.h:111-116):
// 0 SHIFT, 1 SCALE, 2 RESOLUTION_SCALE, 3 CTSCALE.  `pars` is leaked on purpose,
// exactly as src/signal.cpp:139 does (the reference captures the raw pointer).
int ref_hist_add_systematic(void* h, int type, int obs, int extra, int npars,
                            const short* pars) {
  hemi::Array<short>* p = new hemi::Array<short>(npars, true);
  for (int i = 0; i < npars; i++) p->writeOnlyHostPtr()[i] = pars[i];
  EvalHistProbe* e = (EvalHistProbe*)h;
  try {
    switch (type) {
      case pdfz::Systematic::SHIFT: e->AddSystematic(pdfz::ShiftSystematic(obs, p)); break;
      case pdfz::Systematic::SCALE: e->AddSystematic(pdfz::ScaleSystematic(obs, p)); break;
      case pdfz::Systematic::RESOLUTION_SCALE:
        e->AddSystematic(pdfz::ResolutionScaleSystematic(obs, extra, p)); break;
      case pdfz::Systematic::CTSCALE:
        e->AddSystematic(pdfz::CosThetaScaleSystematic(obs, p)); break;
      default: return 1;
    }
  } catch (pdfz::Error&) { return 2; }
  return 0;
}

int ref_hist_set_eval_points(void* h, const float* pts, size_t nfloats, char* err, int errlen) {
  try {
    std::vector<float> p(pts, pts + nfloats);
    ((EvalHistProbe*)h)->SetEvalPoints(p);
    return 0;
  } catch (pdfz::Error& e) { set_err(err, errlen, e.msg); return 1; }
}

// One EvalAsync+EvalFinished.  pdf_io / norm_io are caller arrays whose previous
// contents are preserved outside the written slots (the offset/stride tests need that).
int ref_hist_eval(void* h, const double* params, int nparams, int param_offset,
                  int param_stride, int do_eval_pdf,
                  float* pdf_io, size_t pdf_len, int pdf_offset, int pdf_stride,
                  unsigned* norm_io, int norm_len, int norm_offset) {
  EvalHistProbe* e = (EvalHistProbe*)h;
  hemi::Array<double> pb(nparams > 0 ? nparams : 1, true);
  for (int i = 0; i < nparams; i++) pb.writeOnlyHostPtr()[i] = params[i];
  hemi::Array<unsigned> nb(norm_len, true);
  for (int i = 0; i < norm_len; i++) nb.writeOnlyHostPtr()[i] = norm_io[i];
  hemi::Array<float> out(pdf_len > 0 ? pdf_len : 1, true);
  for (size_t i = 0; i < pdf_len; i++) out.writeOnlyHostPtr()[i] = pdf_io[i];
  e->SetParameterBuffer(&pb, param_offset, param_stride);
  e->SetNormalizationBuffer(&nb, norm_offset);
  e->SetPDFValueBuffer(&out, pdf_offset, pdf_stride);
  e->EvalAsync(do_eval_pdf != 0);
  e->EvalFinished();
  for (int i = 0; i < norm_len; i++) norm_io[i] = nb.readOnlyHostPtr()[i];
  for (size_t i = 0; i < pdf_len; i++) pdf_io[i] = out.readOnlyHostPtr()[i];
  return 0;
}

int ref_hist_total_nbins(void* h) { return ((EvalHistProbe*)h)->TotalBins(); }
double ref_hist_bin_volume(void* h) { return ((EvalHistProbe*)h)->BinVolume(); }
void ref_hist_bins(void* h, unsigned* out) {
  EvalHistProbe* e = (EvalHistProbe*)h;
  std::memcpy(out, e->Bins(), sizeof(unsigned) * e->TotalBins());
}
size_t ref_hist_read_bins(void* h, int* out) {
  EvalHistProbe* e = (EvalHistProbe*)h;
  if (out && e->NReadBins()) std::memcpy(out, e->ReadBins(), sizeof(int) * e->NReadBins());
  return e->NReadBins();
}

// nll_event_chunks -> nll_event_reduce -> nll_total, the reference's own functions
// (CPU build: one "thread", src/mcmc.cpp:37-44,390-415).
double ref_nll(const float* lut, size_t ne, size_t ns, const double* pars, size_t npars,
               size_t nsources, const double* means, const double* sigmas,
               const double* nexpected, const unsigned* n_mc, const short* source_id,
               const unsigned* norms, double* evsum_out) {
  double partial = 0, total = 0, nll = 0;
  nll_event_chunks(lut, pars, ne, ns, nexpected, n_mc, source_id, norms, &partial);
  nll_event_reduce(1, &partial, &total);
  nll_total(npars, pars, ns, nsources, means, sigmas, &total, nexpected, n_mc,
            source_id, norms, &nll);
  if (evsum_out) *evsum_out = total;
  return nll;
}

void ref_set_tapes(const double* normals, size_t n_normals,
                   const double* uniforms, size_t n_uniforms, unsigned seed) {
  gRandom->normals = normals; gRandom->n_normals = n_normals; gRandom->i_normal = 0;
  gRandom->uniforms = uniforms; gRandom->n_uniforms = n_uniforms; gRandom->i_uniform = 0;
  gRandom->SetSeed(seed);
}
size_t ref_tape_normals_used() { return gRandom->i_normal; }
size_t ref_tape_uniforms_used() { return gRandom->i_uniform; }

// Full MCMC::operator() of the reference (src/mcmc.cpp:143-387) over already-built
// EvalHist handles.  Systematics must already be attached to every handle (with
// GLOBAL coefficient indices, src/config.cpp:118-120).  Returns the number of rows
// left in the chain ntuple; rows (nparameters+1 floats each) are copied to chain_out
// up to chain_cap rows.
long ref_mcmc_run(int nsources, const float* src_mean, const float* src_sigma,
                  const int* src_fixed,
                  int nsyst, const int* syst_npars, const double* syst_means,
                  const double* syst_sigmas, const int* syst_fixed,
                  int nobs,
                  int nsignals, void** hists, const double* nexpected,
                  const unsigned* n_mc, const int* source_index,
                  const float* data, size_t data_nfloats,
                  unsigned nsteps, float burnin_fraction, int debug_mode,
                  unsigned sync_interval, int verbose,
                  float* chain_out, long chain_cap) {
  std::vector<Source> sources(nsources);
  for (int i = 0; i < nsources; i++) {
    char nm[32]; snprintf(nm, sizeof nm, "s%d", i);
    sources[i] = Source(nm, i, src_mean[i], src_sigma[i], src_fixed[i] != 0);
  }
  std::vector<Systematic> systs(nsyst);
  std::vector<std::vector<double> > keep_means(nsyst), keep_sigmas(nsyst);
  size_t k = 0;
  for (int i = 0; i < nsyst; i++) {
    char nm[32]; snprintf(nm, sizeof nm, "y%d", i);
    systs[i].name = nm;
    systs[i].npars = syst_npars[i];
    keep_means[i].assign(syst_means + k, syst_means + k + syst_npars[i]);
    keep_sigmas[i].assign(syst_sigmas + k, syst_sigmas + k + syst_npars[i]);
    systs[i].means = &keep_means[i][0];
    systs[i].sigmas = &keep_sigmas[i][0];
    systs[i].fixed = syst_fixed[i] != 0;
    k += syst_npars[i];
  }
  std::vector<Observable> observables(nobs);
  std::vector<Signal> signals(nsignals);
  for (int i = 0; i < nsignals; i++) {
    char nm[32]; snprintf(nm, sizeof nm, "sig%d", i);
    signals[i].name = nm;
    signals[i].histogram = (EvalHistProbe*)hists[i];
    signals[i].nexpected = nexpected[i];
    signals[i].n_mc = n_mc[i];
    signals[i].source = sources[source_index[i]];
  }
  std::vector<float> d(data, data + data_nfloats);
  CoutSilencer quiet(!verbose);
  MCMC mcmc(sources, signals, systs, observables);
  LikelihoodSpace* ls = mcmc(d, nsteps, burnin_fraction, debug_mode != 0, sync_interval);
  const TNtuple* nt = ls->get_samples();
  long rows = nt->GetEntries();
  size_t width = nt->names.size();
  long ncopy = rows < chain_cap ? rows : chain_cap;
  if (chain_out && ncopy > 0)
    std::memcpy(chain_out, &nt->rows[0], sizeof(float) * width * (size_t)ncopy);
  delete ls;
  return rows;
}


// The reference's own FitConfig (src/config.cpp:19-297) run on a JSON file whose signals point at
// .npy tables, written out as JSON text: every ordering and rounding decision a drop-in FitConfig
// has to reproduce.  Returns the text length (truncated to cap), or -1 when FitConfig throws.
long ref_config_dump(const char* json_path, int verbose, char* out, size_t cap) {
  std::ostringstream o;
  o.precision(17);
  try {
    CoutSilencer quiet(!verbose);
    FitConfig fc(json_path);
    o << "{\"nexperiments\": " << fc.nexperiments << ", \"nsteps\": " << fc.nsteps
      << ", \"confidence\": " << (double)fc.confidence << ", \"burnin_fraction\": " << (double)fc.burnin_fraction
      << ", \"debug_mode\": " << (fc.debug_mode ? "true" : "false") << ", \"plots\": " << (fc.plots ? "true" : "false")
      << ", \"seed\": " << fc.seed << ", \"signal_name\": \"" << fc.signal_name << "\", \"output_prefix\": \""
      << fc.output_prefix << "\", \"error_type\": \"" << (fc.error_type == ERROR_PROJECTION ? "projection" : "contour")
      << "\", \"samples\": \"" << fc.samples << "\",\n \"sample_fields\": [";
    for (size_t i = 0; i < fc.sample_fields.size(); i++) o << (i ? ", " : "") << "\"" << fc.sample_fields[i] << "\"";
    o << "],\n \"datasets\": [";
    size_t nd = 0;
    for (std::set<unsigned>::const_iterator it = fc.datasets.begin(); it != fc.datasets.end(); ++it) o << (nd++ ? ", " : "") << *it;
    o << "],\n \"observables\": [";
    for (size_t i = 0; i < fc.observables.size(); i++) {
      const Observable& b = fc.observables[i];
      o << (i ? ", " : "") << "{\"name\": \"" << b.name << "\", \"field\": \"" << b.field << "\", \"field_index\": "
        << b.field_index << ", \"bins\": " << b.bins << ", \"lower\": " << (double)b.lower << ", \"upper\": "
        << (double)b.upper << "}";
    }
    o << "],\n \"cuts\": [";
    for (size_t i = 0; i < fc.cuts.size(); i++) {
      const Observable& b = fc.cuts[i];
      o << (i ? ", " : "") << "{\"name\": \"" << b.name << "\", \"field\": \"" << b.field << "\", \"lower\": "
        << (double)b.lower << ", \"upper\": " << (double)b.upper << "}";
    }
    o << "],\n \"sources\": [";
    for (size_t i = 0; i < fc.sources.size(); i++) {
      const Source& b = fc.sources[i];
      o << (i ? ", " : "") << "{\"name\": \"" << b.name << "\", \"index\": " << b.index << ", \"mean\": "
        << (double)b.mean << ", \"sigma\": " << (double)b.sigma << ", \"fixed\": " << (b.fixed ? "true" : "false") << "}";
    }
    o << "],\n \"systematics\": [";
    for (size_t i = 0; i < fc.systematics.size(); i++) {
      const Systematic& y = fc.systematics[i];
      o << (i ? ", " : "") << "{\"name\": \"" << y.name << "\", \"type\": " << (int)y.type
        << ", \"observable_field\": \"" << y.observable_field << "\", \"truth_field\": \"" << y.truth_field
        << "\", \"observable_field_index\": " << y.observable_field_index << ", \"truth_field_index\": "
        << (y.type == pdfz::Systematic::RESOLUTION_SCALE ? (long)y.truth_field_index : -1L)
        << ", \"npars\": " << y.npars << ", \"fixed\": " << (y.fixed ? "true" : "false") << ", \"means\": [";
      for (size_t k = 0; k < y.npars; k++) o << (k ? ", " : "") << y.means[k];
      o << "], \"sigmas\": [";
      for (size_t k = 0; k < y.npars; k++) o << (k ? ", " : "") << y.sigmas[k];
      o << "], \"pidx\": [";
      for (size_t k = 0; k < y.pidx.size(); k++) o << (k ? ", " : "") << y.pidx[k];
      o << "]}";
    }
    o << "],\n \"signals\": [";
    for (size_t i = 0; i < fc.signals.size(); i++) {
      const Signal& g = fc.signals[i];
      std::vector<float> rows;
      pdfz::EvalHist* h = dynamic_cast<pdfz::EvalHist*>(g.histogram);
      if (h) rows = EvalHistProbe::Rows(h);      // all sample_fields columns, as the Signal built them
      // order-sensitive checksum of the sample rows the PDF was built from (columns in
      // sample_fields order, cuts applied): sum of (index+1) * bit pattern, mod 2^64
      unsigned long long sum = 0;
      for (size_t k = 0; k < rows.size(); k++) {
        unsigned bits; std::memcpy(&bits, &rows[k], 4);
        sum += (unsigned long long)(k + 1) * bits;
      }
      o << (i ? ", " : "") << "{\"name\": \"" << g.name << "\", \"filename\": \"" << g.filename << "\", \"dataset\": "
        << g.dataset << ", \"source\": \"" << g.source.name << "\", \"source_index\": " << g.source.index
        << ", \"nexpected\": " << g.nexpected << ", \"n_mc\": " << g.n_mc << ", \"nsample_floats\": " << rows.size()
        << ", \"sample_checksum\": \"" << sum << "\", \"systematic_names\": [";
      for (size_t k = 0; k < g.systematic_names.size(); k++) o << (k ? ", " : "") << "\"" << g.systematic_names[k] << "\"";
      o << "]}";
    }
    o << "]}\n";
  } catch (...) {
    return -1;
  }
  const std::string t = o.str();
  if (out && cap) {
    std::strncpy(out, t.c_str(), cap - 1);
    out[cap - 1] = 0;
  }
  return (long)t.size();
}

}  // extern "C"
