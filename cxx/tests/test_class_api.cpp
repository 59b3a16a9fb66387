// Runs the sxmc class API (cxx/include/sxmc/*.h) on a GPU the way the reference's gtest suite
// drives it (test/test_pdfz.cpp, test_pdfz_2d.cpp, test_pdfz_syst.cpp: build an EvalHist from a
// fixture, bind buffers, EvalAsync/EvalFinished, compare norm and PDF values), plus one MCMC fit.
//
//   test_class_api <cases.txt> <fit.txt> <outdir>
//
// cases.txt and fit.txt are written by tests/test_cxx_shim.py from the same tables the Python
// parity tests use (tests/golden_vectors.py = the reference's ASSERT constants); the chain rows
// this program writes to <outdir>/rows.bin are compared there with the ctypes mirror's chain.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include <sxmc/mcmc.h>
#include <sxmc/pdfz.h>

#include "sxmc_b200.h"

static int g_failures = 0, g_checks = 0;
#define EXPECT(cond, what)                                                                   \
  do {                                                                                       \
    ++g_checks;                                                                              \
    if (!(cond)) { ++g_failures; std::cerr << "FAIL " << what << " (" #cond ") at line " << __LINE__ << std::endl; } \
  } while (0)

// gtest's ASSERT_FLOAT_EQ: within 4 units in the last place
static bool float_eq_4ulp(float a, float b) {
  if (std::isnan(a) || std::isnan(b)) return false;
  int32_t ia, ib;
  std::memcpy(&ia, &a, 4);
  std::memcpy(&ib, &b, 4);
  int64_t ba = ia < 0 ? (int64_t)0x80000000ll - ia : (int64_t)ia + 0x80000000ll;
  int64_t bb = ib < 0 ? (int64_t)0x80000000ll - ib : (int64_t)ib + 0x80000000ll;
  return std::llabs(ba - bb) <= 4;
}

static std::vector<float> read_floats(std::istream& in, size_t n) {
  std::vector<float> v(n);
  for (size_t i = 0; i < n; ++i) {
    std::string t;
    in >> t;
    v[i] = t == "nan" ? NAN : std::stof(t);
  }
  return v;
}

template <typename T>
static std::vector<T> read_vec(std::istream& in, size_t n) {
  std::vector<T> v(n);
  for (size_t i = 0; i < n; ++i) in >> v[i];
  return v;
}

static pdfz::Systematic* make_systematic(int type, int obs, int extra, hemi::Array<short>* pars) {
  switch (type) {
    case pdfz::Systematic::SHIFT: return new pdfz::ShiftSystematic(obs, pars);
    case pdfz::Systematic::SCALE: return new pdfz::ScaleSystematic(obs, pars);
    case pdfz::Systematic::CTSCALE: return new pdfz::CosThetaScaleSystematic(obs, pars);
    default: return new pdfz::ResolutionScaleSystematic(obs, extra, pars);
  }
}

// one golden case: "case name nsamples nfields nobs | samples | lower upper nbins | npoints | points |
//                   syst_type obs extra (type -1: none) | value | norm | expected[npoints]"
static void run_case(std::istream& in, const std::string& name) {
  size_t nsamples; int nfields, nobs;
  in >> nsamples >> nfields >> nobs;
  std::vector<float> samples = read_floats(in, nsamples * nfields);
  std::vector<double> lower = read_vec<double>(in, nobs), upper = read_vec<double>(in, nobs);
  std::vector<int> nbins = read_vec<int>(in, nobs);
  size_t npoints;
  in >> npoints;
  std::vector<float> points = read_floats(in, npoints * (nobs + 1));
  int stype, sobs, sextra;
  double value;
  unsigned want_norm;
  in >> stype >> sobs >> sextra >> value >> want_norm;
  std::vector<float> want = read_floats(in, npoints);

  pdfz::EvalHist* evaluator = new pdfz::EvalHist(samples, nfields, nobs, lower, upper, nbins);
  hemi::Array<float>* pdf_values = new hemi::Array<float>(npoints, true);
  hemi::Array<unsigned int>* norm = new hemi::Array<unsigned int>(1, true);
  hemi::Array<double>* params = new hemi::Array<double>(1, true);
  hemi::Array<short>* pars = new hemi::Array<short>(1, true);
  pars->writeOnlyHostPtr()[0] = 0;
  evaluator->SetEvalPoints(points);
  evaluator->SetPDFValueBuffer(pdf_values);
  evaluator->SetNormalizationBuffer(norm);
  evaluator->SetParameterBuffer(params);
  pdfz::Systematic* syst = NULL;
  if (stype >= 0) {
    syst = make_systematic(stype, sobs, sextra, pars);
    evaluator->AddSystematic(*syst);
  }
  params->writeOnlyHostPtr()[0] = value;
  evaluator->EvalAsync();
  evaluator->EvalFinished();
  EXPECT(*norm->readOnlyHostPtr() == want_norm, name + ": norm");
  float* results = pdf_values->hostPtr();
  for (size_t i = 0; i < npoints; ++i) {
    if (std::isnan(want[i])) EXPECT(std::isnan(results[i]), name + ": NaN expected");
    else EXPECT(float_eq_4ulp(want[i], results[i]), name + ": pdf value");
  }
  // the density integrates to one over the domain (CreateHistogram, src/pdfz.cpp:497-592)
  pdfz::HistogramData hd = evaluator->CreateHistogram();
  double integral = 0, vol = 1;
  for (int d = 0; d < nobs; ++d) vol *= (upper[d] - lower[d]) / nbins[d];
  for (size_t i = 0; i < hd.contents.size(); ++i) integral += hd.contents[i] * vol;
  EXPECT(hd.norm == want_norm && std::fabs(integral - 1.0) < 1e-12, name + ": CreateHistogram integral");
  // GetSamples hands back observables + dataset id, original order (src/pdfz.h:542-556)
  std::vector<float> sv;
  evaluator->GetSamples(sv);
  bool same = sv.size() == nsamples * (nobs + 1);
  for (size_t i = 0; same && i < nsamples; ++i)
    for (int j = 0; j < nobs; ++j) same = same && sv[i * (nobs + 1) + j] == samples[i * nfields + j];
  EXPECT(same, name + ": GetSamples");
  delete syst;
  delete evaluator;
  delete pdf_values; delete norm; delete params; delete pars;
}

// "error name nfloats nfields nobs nlower nupper nnbins nbins... | message": the constructor must throw
// pdfz::Error by value with the reference's text (test/test_pdfz.cpp:42-73)
static void run_ctor_error(std::istream& in, const std::string& name) {
  size_t nfloats; int nfields, nobs, nlower, nupper, nnbins;
  in >> nfloats >> nfields >> nobs >> nlower >> nupper >> nnbins;
  std::vector<int> nbins = read_vec<int>(in, nnbins);
  std::string msg;
  std::getline(in, msg);
  while (!msg.empty() && msg[0] == ' ') msg.erase(0, 1);
  std::vector<float> samples(nfloats, 0.25f);
  std::vector<double> lower(nlower, 0.0), upper(nupper, 1.0);
  bool thrown = false;
  try {
    pdfz::EvalHist e(samples, nfields, nobs, lower, upper, nbins);
  } catch (pdfz::Error err) {
    thrown = true;
    EXPECT(err.msg == msg, name + ": message '" + err.msg + "'");
  }
  EXPECT(thrown, name + ": pdfz::Error expected");
}

// pdf buffer offset/stride and norm slot (test/test_pdfz.cpp:98-126)
static void run_offset_stride() {
  const float s[7] = {0.1f, 0.2f, 0.3f, 0.4f, 0.5f, 1.1f, -0.1f};
  const float p[12] = {-0.1f, 0, 0.0f, 0, 0.25f, 0, 0.5f, 0, 0.75f, 0, 1.0f, 0};
  std::vector<float> samples(s, s + 7), points(p, p + 12);
  std::vector<double> lower(1, 0.0), upper(1, 1.0);
  std::vector<int> nbins(1, 2);
  pdfz::EvalHist e(samples, 1, 1, lower, upper, nbins);
  hemi::Array<float> pdf(20, true);
  hemi::Array<unsigned> norm(3, true);
  hemi::Array<double> params(1, true);
  float* ph = pdf.writeOnlyHostPtr();
  for (int i = 0; i < 20; ++i) ph[i] = -7.0f;
  unsigned* nh = norm.writeOnlyHostPtr();
  nh[0] = 77; nh[1] = 88; nh[2] = 99;
  params.writeOnlyHostPtr()[0] = 0;
  pdf.readOnlyDevicePtr();       // both sides valid, so the untouched slots survive the round trip
  norm.readOnlyDevicePtr();
  e.SetEvalPoints(points);
  e.SetPDFValueBuffer(&pdf, 3, 2);
  e.SetNormalizationBuffer(&norm, 1);
  e.SetParameterBuffer(&params);
  e.EvalAsync();
  e.EvalFinished();
  const unsigned* n = norm.readOnlyHostPtr();
  EXPECT(n[0] == 77 && n[1] == 5 && n[2] == 99, "offset_stride: norm slots");
  const float* r = pdf.readOnlyHostPtr();
  EXPECT(std::isnan(r[3]) && float_eq_4ulp(r[5], 1.6f) && float_eq_4ulp(r[7], 1.6f) && float_eq_4ulp(r[9], 0.4f) &&
             float_eq_4ulp(r[11], 0.4f) && std::isnan(r[13]), "offset_stride: values");
  bool untouched = true;
  for (int i = 0; i < 20; ++i)
    if (i < 3 || i > 13 || (i - 3) % 2) untouched = untouched && r[i] == -7.0f;
  EXPECT(untouched, "offset_stride: other slots untouched");
}

static void run_random_sample() {
  std::vector<float> samples;
  uint32_t x = 12345u;
  for (int i = 0; i < 20000; ++i) {      // triangular-ish distribution on [0, 1)
    x = x * 1664525u + 1013904223u; float a = (x >> 8) * (1.0f / 16777216.0f);
    x = x * 1664525u + 1013904223u; float b = (x >> 8) * (1.0f / 16777216.0f);
    samples.push_back(0.5f * (a + b));
  }
  std::vector<double> lower(1, 0.0), upper(1, 1.0);
  std::vector<int> nbins(1, 20);
  pdfz::EvalHist e(samples, 1, 1, lower, upper, nbins, 3);
  pdfz::EvalHist::SetRandomSeed(99);
  std::vector<float> events;
  int n = e.RandomSample(events, 5000.4, false, 3);
  EXPECT(n == 5000 && events.size() == 10000u, "RandomSample: nint(nexpected) events");   // src/pdfz.cpp:838-843
  bool ok = true;
  double mean = 0;
  for (size_t i = 0; i + 1 < events.size(); i += 2) {
    ok = ok && events[i] >= 0.0f && events[i] < 1.0f && events[i + 1] == 3.0f;
    mean += events[i];
  }
  mean /= n > 0 ? n : 1;
  EXPECT(ok, "RandomSample: events inside the domain, dataset tag kept");
  EXPECT(std::fabs(mean - 0.5) < 0.02, "RandomSample: mean of the sampled PDF");
  std::vector<float> ev2;
  int m = e.RandomSample(ev2, 5000.0, true, 0);
  EXPECT(std::abs(m - 5000) < 400 && ev2.size() == 2u * m, "RandomSample: Poisson count");
}

// fit.txt (see tests/test_cxx_shim.py) -> Signal/Source/Systematic/Observable -> MCMC -> rows.bin
static void run_fit(const char* path, const std::string& outdir) {
  std::ifstream in(path);
  std::string dir(path);
  dir = dir.substr(0, dir.find_last_of('/') + 1);
  int nobs, nfields;
  in >> nobs >> nfields;
  std::vector<Observable> observables(nobs);
  for (int i = 0; i < nobs; ++i) {
    in >> observables[i].lower >> observables[i].upper >> observables[i].bins;
    observables[i].field_index = i;
    observables[i].name = "o" + std::to_string(i);
  }
  int nsources;
  in >> nsources;
  std::vector<Source> sources(nsources);
  for (int i = 0; i < nsources; ++i) {
    int fixed;
    in >> sources[i].mean >> sources[i].sigma >> fixed;
    sources[i].fixed = fixed != 0;
    sources[i].index = i;
    sources[i].name = "s" + std::to_string(i);
  }
  int nsyst;
  in >> nsyst;
  std::vector<Systematic> systematics(nsyst);
  std::vector<std::vector<double> > means(nsyst), sigmas(nsyst);
  for (int i = 0; i < nsyst; ++i) {
    int type, fixed;
    in >> type >> systematics[i].observable_field_index >> systematics[i].truth_field_index >>
        systematics[i].npars >> fixed;
    systematics[i].type = (pdfz::Systematic::Type)type;
    systematics[i].fixed = fixed != 0;
    systematics[i].name = "y" + std::to_string(i);
    means[i] = read_vec<double>(in, systematics[i].npars);
    sigmas[i] = read_vec<double>(in, systematics[i].npars);
    systematics[i].pidx = read_vec<short>(in, systematics[i].npars);
    systematics[i].means = means[i].data();
    systematics[i].sigmas = sigmas[i].data();
  }
  int nsignals;
  in >> nsignals;
  std::vector<Signal> signals(nsignals);
  for (int j = 0; j < nsignals; ++j) {
    int source; size_t nrows; std::string file;
    in >> signals[j].nexpected >> source >> signals[j].n_mc >> nrows >> file;
    std::vector<float> samples(nrows * nfields);
    std::ifstream f((dir + file).c_str(), std::ios::binary);
    f.read((char*)samples.data(), samples.size() * sizeof(float));
    EXPECT((size_t)f.gcount() == samples.size() * sizeof(float), "fit: sample file " + file);
    signals[j].source = sources[source];
    signals[j].name = "sig" + std::to_string(j);
    const size_t n_mc = signals[j].n_mc;
    signals[j].build(samples, nfields, observables, systematics);
    signals[j].n_mc = n_mc;
  }
  size_t ndata; std::string datafile;
  in >> ndata >> datafile;
  std::vector<float> data(ndata);
  {
    std::ifstream f((dir + datafile).c_str(), std::ios::binary);
    f.read((char*)data.data(), ndata * sizeof(float));
  }
  unsigned nsteps; float burnin; uint64_t seed;
  in >> nsteps >> burnin >> seed;

  // Signal::get_efficiency at the systematics' means (src/signal.cpp:172-197)
  std::vector<double> eff(nsignals);
  for (int j = 0; j < nsignals; ++j) eff[j] = signals[j].get_efficiency(systematics);

  MCMC mcmc(sources, signals, systematics, observables);
  mcmc.SetSeed(seed, 0);
  LikelihoodSpace* ls = mcmc(data, nsteps, burnin);
  const NtupleRows* nt = ls->get_samples();
  EXPECT(nt->nrows() > 0, "fit: chain rows");
  EXPECT(nt->names.size() == (size_t)nsources + nsyst + 1 && nt->names.back() == "likelihood", "fit: ntuple columns");
  float nll = 0;
  std::map<std::string, Interval> best = ls->get_best_fit(nll);
  EXPECT(best.size() == nt->names.size() - 1, "fit: one interval per parameter");
  {
    std::ofstream f((outdir + "/rows.bin").c_str(), std::ios::binary);
    f.write((const char*)nt->rows.data(), nt->rows.size() * sizeof(float));
    std::ofstream g((outdir + "/fit_out.txt").c_str());
    g.precision(9);
    g << nt->nrows() << " " << mcmc.accepted() << " " << nll << "\n";
    for (int j = 0; j < nsignals; ++j) { g.precision(17); g << eff[j] << "\n"; }
    g.precision(9);
    for (size_t i = 0; i + 1 < nt->names.size(); ++i) {
      const Interval& iv = best[nt->names[i]];
      g << nt->names[i] << " " << iv.point_estimate << " " << iv.lower << " " << iv.upper << "\n";
    }
  }
  ls->print_best_fit();
  delete ls;
  for (int j = 0; j < nsignals; ++j) delete signals[j].histogram;
}

int main(int argc, char** argv) {
  if (argc < 4) { std::cerr << "usage: test_class_api cases.txt fit.txt outdir" << std::endl; return 2; }
  if (sxmc_init(0) != SXMC_OK) { std::cerr << "sxmc_init: " << sxmc_last_error() << std::endl; return 3; }
  try {
    std::ifstream in(argv[1]);
    std::string kind, name;
    while (in >> kind >> name) {
      if (kind == "case") run_case(in, name);
      else if (kind == "error") run_ctor_error(in, name);
      else { std::cerr << "unknown record " << kind << std::endl; return 2; }
    }
    run_offset_stride();
    run_random_sample();
    run_fit(argv[2], argv[3]);
  } catch (pdfz::Error e) {
    std::cerr << "pdfz::Error: " << e.msg << std::endl;
    return 4;
  } catch (const std::exception& e) {
    std::cerr << "exception: " << e.what() << std::endl;
    return 4;
  }
  std::cout << g_checks << " checks, " << g_failures << " failures" << std::endl;
  return g_failures ? 1 : 0;
}
