// sxmc -- the command line of the reference (src/sxmc.cpp:44-235): load the JSON configuration,
// run the ensemble of fits, print the upper limits and their median.
//
//   cxx/bin/sxmc fit_configuration.json output_path
//
// Without ROOT: sample files are float tables (cxx/include/sxmc/config.h), each experiment's
// likelihood space is written as <output_path>/<output_prefix>_<i>.npy (float32 rows: one
// column per parameter, then "likelihood") instead of a TNtuple in a .root file, there are no
// plots, and only the contour error estimator exists.  One process drives one GPU
// (SXMC_DEVICE selects it); experiment i uses Philox stream i, so SXMC_RANK / SXMC_WORLD split
// an ensemble over processes without changing any experiment's result.
#include <sys/stat.h>

#include <algorithm>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

#include <sxmc/config.h>
#include <sxmc/mcmc.h>
#include <sxmc/pdfz.h>

#include "sxmc_b200.h"

// src/sxmc.cpp:145-160
static int check_create_output(const std::string& path) {
  struct stat st;
  if (stat(path.c_str(), &st) != 0) {
    if (mkdir(path.c_str(), 0777) != 0 && errno != EEXIST) {
      std::cerr << "Error creating output directory, errno = " << errno << std::endl;
      return -1;
    }
  } else if (!S_ISDIR(st.st_mode)) {
    std::cerr << "Output path exists and is not a directory" << std::endl;
    return -1;
  }
  return 0;
}

// src/utils.h:77-89
static float median(std::vector<float> v) {
  std::sort(v.begin(), v.end());
  const size_t m = v.size() / 2;
  return v.size() % 2 ? v[m] : (v[m - 1] + v[m]) / 2;
}

static void write_npy(const std::string& path, const NtupleRows& nt) {
  const size_t ncol = nt.names.size();
  std::ostringstream h;
  h << "{'descr': '<f4', 'fortran_order': False, 'shape': (" << nt.nrows() << ", " << ncol << "), }";
  std::string header = h.str();
  while ((10 + header.size() + 1) % 64) header.push_back(' ');
  header.push_back('\n');
  std::ofstream f(path.c_str(), std::ios::binary);
  f.write("\x93NUMPY\x01\x00", 8);
  const unsigned short hl = (unsigned short)header.size();
  f.write((const char*)&hl, 2);
  f.write(header.data(), header.size());
  f.write((const char*)nt.rows.data(), nt.rows.size() * sizeof(float));
}

// what the loop of src/sxmc.cpp:110-134 does with a finished likelihood space
static void report_experiment(FitConfig& fc, unsigned i, bool first, LikelihoodSpace* ls,
                              const std::string& output_path, std::vector<float>& limits) {
  float lm = 0;
  std::map<std::string, Interval> best_fit = ls->extract_best_fit(lm, fc.confidence, fc.error_type);
  std::cout << "-- Best fit --" << std::endl;
  for (std::map<std::string, Interval>::const_iterator it = best_fit.begin(); it != best_fit.end(); ++it)
    std::cout << " " << it->first << ": " << it->second.str() << std::endl;

  if (fc.signal_name != "" && best_fit.find(fc.signal_name) != best_fit.end()) {
    Interval bfi = best_fit[fc.signal_name];
    std::cout << "ensemble: Signal " << fc.signal_name << ": " << bfi.str() << std::endl;
    if (!bfi.one_sided) std::cerr << "ensemble: Warning: Two-sided limit!" << std::endl;
    std::cout << "ensemble: lower = " << bfi.lower << ", upper = " << bfi.upper << ", coverage = " << bfi.coverage
              << std::endl;
    limits.push_back(bfi.upper);
  }
  std::ostringstream lsfile;
  lsfile << output_path << fc.output_prefix << "_" << i << ".npy";
  write_npy(lsfile.str(), *ls->get_samples());
  if (first) {
    std::ofstream cols((output_path + fc.output_prefix + "_columns.txt").c_str());
    for (size_t k = 0; k < ls->get_samples()->names.size(); ++k) cols << ls->get_samples()->names[k] << "\n";
  }
}

// src/sxmc.cpp:44-141
static std::vector<float> ensemble(FitConfig& fc, const std::string& output_path, unsigned rank, unsigned world) {
  std::vector<float> limits;
  if (fc.samples != "") throw std::runtime_error("fit.samples (re-reading a saved likelihood space) needs ROOT");
  unsigned done = 0;                 // experiments of this rank already fitted by the batch path
  if (fc.data.empty() && !std::getenv("SXMC_NO_BATCH")) {
    // Fake-data ensembles of small fits: the fits of up to 148 experiments advance together in one
    // cooperative launch, their data sets drawn on the device.  Experiment i is the chain the
    // one-at-a-time loop below produces (same Philox stream), only ~70x sooner.
    try {
      MCMCBatch batch(fc.sources, fc.signals, fc.systematics, fc.observables);
      std::vector<unsigned> mine;
      for (unsigned i = rank; i < fc.nexperiments; i += world) mine.push_back(i);
      for (size_t g0 = 0; g0 < mine.size(); g0 += (size_t)batch.max_chains()) {
        const int n = (int)std::min(mine.size() - g0, (size_t)batch.max_chains());
        for (int slot = 0; slot < n; ++slot) {
          std::vector<unsigned> observed;
          batch.GenerateData(slot, true, (uint64_t)fc.seed, mine[g0 + slot], &observed);
          batch.SetSeed(slot, (uint64_t)fc.seed, mine[g0 + slot]);
        }
        batch.Run(n, fc.nsteps, fc.burnin_fraction, fc.debug_mode);
        for (int slot = 0; slot < n; ++slot) {
          const unsigned i = mine[g0 + slot];
          std::cout << "Experiment " << i + 1 << " / " << fc.nexperiments << std::endl;
          LikelihoodSpace* ls = batch.GetLikelihoodSpace(slot);
          report_experiment(fc, i, i == rank, ls, output_path, limits);
          delete ls;
          ++done;
        }
      }
      return limits;
    } catch (const MCMCBatch::NotSmallFit& e) {
      if (done) throw;
      std::cout << "ensemble: one fit at a time (" << e.what() << ")" << std::endl;
    }
  }
  for (unsigned i = rank; i < fc.nexperiments; i += world) {
    std::cout << "Experiment " << i + 1 << " / " << fc.nexperiments << std::endl;
    std::vector<float> samples;
    if (fc.data.empty()) {
      std::cout << "ensemble: Sampling fake dataset " << i << std::endl;
      std::vector<unsigned> observed;
      samples = make_fake_dataset(fc.signals, fc.systematics, fc.observables, true, (uint64_t)fc.seed, i, &observed);
      for (size_t j = 0; j < observed.size(); ++j)
        std::cout << "make_fake_dataset: " << fc.signals[j].name << ": " << observed[j] << " events" << std::endl;
    } else {
      for (std::map<unsigned, std::vector<Signal> >::iterator it = fc.data.begin(); it != fc.data.end(); ++it) {
        if (i >= it->second.size()) throw std::runtime_error("ensemble: no data file for this experiment");
        std::cout << "ensemble: Loading dataset " << it->first << "." << i << " (" << it->second[i].filename << ")"
                  << std::endl;
        dynamic_cast<pdfz::EvalHist*>(it->second[i].histogram)->GetSamples(samples);
      }
    }
    MCMC mcmc(fc.sources, fc.signals, fc.systematics, fc.observables);
    mcmc.SetSeed((uint64_t)fc.seed, i);
    LikelihoodSpace* ls = mcmc(samples, fc.nsteps, fc.burnin_fraction, fc.debug_mode);
    report_experiment(fc, i, i == rank, ls, output_path, limits);
    delete ls;
  }
  return limits;
}

// `sxmc --dump-config fit.json`: what FitConfig made of the file, as JSON text in the schema of
// oracle/ref_glue.cpp::ref_config_dump (the same dump of the reference's own FitConfig) -- the
// parity tests compare the two key by key.
static void dump_config(const FitConfig& fc, std::ostream& o) {
  o.precision(17);
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
    if (g.histogram) {
      rows.resize(sxmc_hist_nsamples(g.histogram->handle()) * fc.sample_fields.size());
      if (!rows.empty() && sxmc_hist_get_samples(g.histogram->handle(), rows.data()) != SXMC_OK)
        throw std::runtime_error(sxmc_last_error());
    }
    unsigned long long sum = 0;      // order-sensitive checksum of the sample rows, as in ref_config_dump
    for (size_t k = 0; k < rows.size(); k++) {
      unsigned bits;
      std::memcpy(&bits, &rows[k], 4);
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
}

int main(int argc, char* argv[]) {
  const bool dump = argc == 3 && std::string(argv[1]) == "--dump-config";
  if (argc != 3) {
    std::cerr << "Usage: " << argv[0] << " fit_configuration.json output_path" << std::endl;
    return 1;
  }
  const char* dev = std::getenv("SXMC_DEVICE");
  const char* rk = std::getenv("SXMC_RANK");
  const char* wd = std::getenv("SXMC_WORLD");
  const unsigned rank = rk ? (unsigned)std::atoi(rk) : 0, world = wd ? (unsigned)std::max(1, std::atoi(wd)) : 1;
  try {
    if (sxmc_init(dev ? std::atoi(dev) : 0) != SXMC_OK) {      // no CUDA device: there is no CPU path
      std::cerr << "sxmc: " << sxmc_last_error() << std::endl;
      return 3;
    }
    if (dump) {
      std::streambuf* quiet = std::cout.rdbuf(NULL);      // FitConfig's progress lines
      FitConfig fc(argv[2]);
      std::cout.rdbuf(quiet);
      dump_config(fc, std::cout);
      return 0;
    }
    std::string output_path(argv[2]);
    if (output_path.empty() || output_path[output_path.size() - 1] != '/') output_path += "/";
    if (check_create_output(output_path) != 0) {
      std::cerr << "Error with output path." << std::endl;
      return 2;
    }
    std::cout << "sxmc: Loading configuration..." << std::endl;
    std::cout << "sxmc: Configuration: " << argv[1] << std::endl;
    FitConfig fc(argv[1]);
    fc.print();

    std::cout << "sxmc: Running ensemble..." << std::endl;
    std::vector<float> limits = ensemble(fc, output_path, rank, world);

    std::cout << "sxmc: Upper limits: ";
    for (size_t i = 0; i < limits.size(); i++) std::cout << limits[i] << ", ";
    std::cout << std::endl;
    std::cout << "sxmc: Median upper limit: " << (!limits.empty() ? median(limits) : -1) << std::endl;
  } catch (pdfz::Error e) {
    std::cerr << "sxmc: pdfz::Error: " << e.msg << std::endl;
    return 4;
  } catch (const std::exception& e) {
    std::cerr << "sxmc: " << e.what() << std::endl;
    return 4;
  }
  return 0;
}
