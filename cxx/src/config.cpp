// FitConfig, the float-table reader and make_fake_dataset (reference bodies: src/config.cpp:19-297,
// src/signal.cpp:11-108, src/io/ttree_io.cpp:21-159, src/generator.cpp:10-48).
#include <sxmc/config.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

#include <sxmc/json_min.h>
#include <sxmc/pdfz.h>

#include "sxmc_b200.h"

namespace {

void fail(const std::string& what) { throw std::runtime_error(what); }

std::string slurp(const std::string& filename, bool binary = false) {
  std::ifstream f(filename.c_str(), binary ? std::ios::binary : std::ios::in);
  if (!f) fail(filename + ": cannot open");
  std::ostringstream ss;
  ss << f.rdbuf();
  return ss.str();
}

std::vector<std::string> sidecar_fields(const std::string& filename) {
  const std::string path = filename + ".fields";
  std::ifstream f(path.c_str());
  if (!f) fail(filename + ": needs a '" + path + "' file naming its columns");
  std::vector<std::string> out;
  std::string line;
  while (std::getline(f, line)) {
    while (!line.empty() && std::isspace((unsigned char)line[line.size() - 1])) line.erase(line.size() - 1);
    size_t a = 0;
    while (a < line.size() && std::isspace((unsigned char)line[a])) ++a;
    if (a < line.size()) out.push_back(line.substr(a));
  }
  return out;
}

bool ends_with(const std::string& s, const std::string& suffix) {
  return s.size() >= suffix.size() && s.compare(s.size() - suffix.size(), suffix.size(), suffix) == 0;
}

// (plain pointers for the names: gcc's -Wdangling-reference flags any reference-returning call
// that is handed a temporary std::string)
const jsonmin::Value& need(const jsonmin::Value& obj, const char* key, const char* where) {
  if (!obj.isMember(key)) fail(std::string(where) + ": missing key '" + key + "'");
  return obj[key];
}

}  // namespace

namespace sxmc {
namespace io {

void read_float_table(const std::string& filename, std::vector<float>& data, std::vector<unsigned>& rank,
                      std::vector<std::string>& fields) {
  data.clear();
  fields.clear();
  if (ends_with(filename, ".root"))
    fail(filename + ": ROOT files cannot be read on this box (no ROOT); convert the tree to a .npy table");
  if (ends_with(filename, ".npy")) {
    const std::string raw = slurp(filename, true);
    if (raw.size() < 10 || raw.compare(0, 6, "\x93NUMPY") != 0) fail(filename + ": not a .npy file");
    const int major = (unsigned char)raw[6];
    size_t hlen, hoff;
    if (major == 1) { hlen = (unsigned char)raw[8] | ((unsigned char)raw[9] << 8); hoff = 10; }
    else { uint32_t h; std::memcpy(&h, raw.data() + 8, 4); hlen = h; hoff = 12; }
    const std::string header = raw.substr(hoff, hlen);
    const bool f8 = header.find("'<f8'") != std::string::npos;
    if (!f8 && header.find("'<f4'") == std::string::npos) fail(filename + ": dtype must be little-endian float32 or float64");
    if (header.find("'fortran_order': False") == std::string::npos) fail(filename + ": must be C-ordered");
    size_t p = header.find("'shape': (");
    if (p == std::string::npos) fail(filename + ": no shape in the header");
    p += 10;
    std::vector<size_t> shape;
    while (p < header.size() && header[p] != ')') {
      if (std::isdigit((unsigned char)header[p])) {
        size_t v = 0;
        while (p < header.size() && std::isdigit((unsigned char)header[p])) v = v * 10 + (header[p++] - '0');
        shape.push_back(v);
      } else {
        ++p;
      }
    }
    if (shape.size() != 2) fail(filename + ": expected a 2-D array [rows, columns]");
    fields = sidecar_fields(filename);
    if (fields.size() != shape[1]) fail(filename + ": column count does not match its .fields file");
    const size_t n = shape[0] * shape[1];
    const char* body = raw.data() + hoff + hlen;
    if (raw.size() < hoff + hlen + n * (f8 ? 8 : 4)) fail(filename + ": truncated");
    data.resize(n);
    if (f8) for (size_t i = 0; i < n; ++i) { double d; std::memcpy(&d, body + 8 * i, 8); data[i] = (float)d; }
    else std::memcpy(data.data(), body, 4 * n);
  } else if (ends_with(filename, ".f32")) {
    fields = sidecar_fields(filename);
    const std::string raw = slurp(filename, true);
    if (fields.empty() || raw.size() % (4 * fields.size())) fail(filename + ": size is not a multiple of its columns");
    data.resize(raw.size() / 4);
    std::memcpy(data.data(), raw.data(), raw.size());
  } else if (ends_with(filename, ".csv") || ends_with(filename, ".txt")) {
    std::ifstream f(filename.c_str());
    if (!f) fail(filename + ": cannot open");
    std::string line;
    std::getline(f, line);
    for (size_t i = 0; i < line.size(); ++i) if (line[i] == ',' || line[i] == '#') line[i] = ' ';
    std::istringstream hs(line);
    std::string tok;
    while (hs >> tok) fields.push_back(tok);
    while (std::getline(f, line)) {
      for (size_t i = 0; i < line.size(); ++i) if (line[i] == ',') line[i] = ' ';
      std::istringstream ls(line);
      double v;
      while (ls >> v) data.push_back((float)v);
    }
    if (fields.empty() || data.size() % fields.size()) fail(filename + ": ragged rows");
  } else {
    fail(filename + ": unknown sample file type (.npy, .f32, .csv)");
  }
  rank.assign(2, 0);
  rank[0] = (unsigned)(data.size() / fields.size());
  rank[1] = (unsigned)fields.size();
}

}  // namespace io
}  // namespace sxmc

void read_dataset_to_samples(std::vector<float>& samples, std::vector<float>& dataset, unsigned dataset_id,
                             std::vector<std::string>& sample_fields, std::vector<std::string>& dataset_fields,
                             std::vector<Observable>& cuts) {
  const size_t ncol = dataset_fields.size(), nout = sample_fields.size();
  std::vector<int> cut_of(ncol, -1);
  for (size_t i = 0; i < ncol; ++i)
    for (size_t j = 0; j < cuts.size(); ++j)
      if (cuts[j].field == dataset_fields[i]) cut_of[i] = (int)j;
  std::vector<size_t> src(nout - 1);
  for (size_t i = 0; i + 1 < nout; ++i) {
    const size_t idx = std::find(dataset_fields.begin(), dataset_fields.end(), sample_fields[i]) - dataset_fields.begin();
    if (idx >= ncol) fail("sample field '" + sample_fields[i] + "' is not a column of the file");
    src[i] = idx;
  }
  const size_t nrows = ncol ? dataset.size() / ncol : 0;
  samples.clear();
  samples.reserve(nrows * nout);
  for (size_t r = 0; r < nrows; ++r) {
    bool keep = true;
    for (size_t j = 0; j < ncol && keep; ++j) {
      if (cut_of[j] < 0) continue;
      const float v = dataset[r * ncol + j];
      const Observable& c = cuts[cut_of[j]];
      if (v < (double)c.lower || v > (double)c.upper) keep = false;       // both ends included (src/signal.cpp:86)
    }
    if (!keep) continue;
    for (size_t i = 0; i + 1 < nout; ++i) samples.push_back(dataset[r * ncol + src[i]]);
    samples.push_back((float)dataset_id);
  }
}

// ---- FitConfig -------------------------------------------------------------------------

static Observable make_observable(const std::string& name, const jsonmin::Value& c) {     // src/observable.cpp:8-34
  const std::string where = "observable '" + name + "'";
  Observable o;
  o.name = name;
  o.title = need(c, "title", where.c_str()).asString();
  o.field = need(c, "field", where.c_str()).asString();
  o.bins = (size_t)need(c, "bins", where.c_str()).asInt();
  o.lower = need(c, "min", where.c_str()).asFloat();
  o.upper = need(c, "max", where.c_str()).asFloat();
  o.units = c["units"].asString();
  o.logscale = c["logscale"].asBool(false);
  o.yrange.assign(2, -1.0f);
  if (c.isMember("yrange")) { o.yrange[0] = c["yrange"][(size_t)0].asFloat(); o.yrange[1] = c["yrange"][(size_t)1].asFloat(); }
  return o;
}

FitConfig::~FitConfig() {
  for (size_t i = 0; i < signals.size(); ++i) {
    delete signals[i].histogram;
    for (size_t k = 0; k < signals[i].owned_pars.size(); ++k) delete signals[i].owned_pars[k];
  }
  for (std::map<unsigned, std::vector<Signal> >::iterator it = data.begin(); it != data.end(); ++it)
    for (size_t i = 0; i < it->second.size(); ++i) delete it->second[i].histogram;
  for (size_t i = 0; i < owned.size(); ++i) delete owned[i];
}

FitConfig::FitConfig(std::string filename) {
  jsonmin::Value root;
  try {
    root = jsonmin::parse(slurp(filename));
  } catch (const std::runtime_error& e) {
    fail(std::string("FitConfig::FitConfig: ") + e.what());                                // src/config.cpp:28-33
  }
  std::string base = filename;
  const size_t slash = base.find_last_of('/');
  base = slash == std::string::npos ? std::string("") : base.substr(0, slash + 1);

  const jsonmin::Value& fit = need(root, "fit", "config");
  const jsonmin::Value& obs_params = need(need(root, "pdfs", "config"), "observables", "pdfs");
  const jsonmin::Value& sys_params = root["pdfs"]["systematics"];
  const jsonmin::Value& sig_params = need(root, "signals", "config");
  const jsonmin::Value& src_params = root["sources"];

  nexperiments = (unsigned)need(fit, "nexperiments", "fit").asInt();                       // :43-70
  nsteps = (unsigned)need(fit, "nsteps", "fit").asInt();
  samples = fit["samples"].asString("");
  const std::string err_type = fit["error_type"].asString("contour");
  if (err_type == "contour") error_type = ERROR_CONTOUR;
  else if (err_type == "projection") error_type = ERROR_PROJECTION;
  else fail("FitConfig::FitConfig: Unknown error type " + err_type);
  burnin_fraction = fit.isMember("burnin_fraction") ? fit["burnin_fraction"].asFloat() : 0.1f;
  debug_mode = fit["debug_mode"].asBool(false);
  output_prefix = fit["output_prefix"].asString("lspace");
  plots = fit.isMember("plots") ? fit["plots"].asBool() : true;
  seed = fit["seed"].asInt(0);
  confidence = fit.isMember("confidence") ? fit["confidence"].asFloat() : 0.683f;
  signal_name = fit["signal_name"].asString("");

  const jsonmin::Value& fobs = need(fit, "observables", "fit");                           // :74-91
  for (size_t i = 0; i < fobs.size(); ++i) {
    const std::string name = fobs[i].asString();
    if (!obs_params.isMember(name)) fail("fit.observables: '" + name + "' is not defined under pdfs.observables");
    observables.push_back(make_observable(name, obs_params[name]));
  }
  const jsonmin::Value& fcuts = fit["cuts"];
  for (size_t i = 0; i < fcuts.size(); ++i) {
    const std::string name = fcuts[i].asString();
    if (!obs_params.isMember(name)) fail("fit.cuts: '" + name + "' is not defined under pdfs.observables");
    for (size_t j = 0; j < observables.size(); ++j)
      if (observables[j].name == name) fail("fit.cuts: '" + name + "' is already an observable");
    cuts.push_back(make_observable(name, obs_params[name]));
  }

  // systematics and sources: union over all configured signals, in sorted key order (:93-151)
  short pidx = 0;
  for (std::map<std::string, jsonmin::Value>::const_iterator it = sig_params.obj.begin(); it != sig_params.obj.end(); ++it) {
    const std::string& sname = it->first;
    const jsonmin::Value& sc = it->second;
    const jsonmin::Value& slist = sc["systematics"];
    for (size_t q = 0; q < slist.size(); ++q) {
      const std::string yname = slist[q].asString();
      if (!sys_params.isMember(yname)) fail("signal '" + sname + "': systematic '" + yname + "' is not defined");
      bool exists = false;
      for (size_t k = 0; k < systematics.size(); ++k) exists = exists || systematics[k].name == yname;
      if (exists) continue;
      const jsonmin::Value& yc = sys_params[yname];                                       // src/systematic.cpp:10-66
      const std::string where = "systematic '" + yname + "'";
      Systematic y;
      y.name = yname;
      y.title = need(yc, "title", where.c_str()).asString();
      y.observable_field = need(yc, "observable_field", where.c_str()).asString();
      const std::string type = need(yc, "type", where.c_str()).asString();
      if (type == "shift") y.type = pdfz::Systematic::SHIFT;
      else if (type == "scale") y.type = pdfz::Systematic::SCALE;
      else if (type == "ctscale") y.type = pdfz::Systematic::CTSCALE;
      else if (type == "resolution_scale") {
        y.type = pdfz::Systematic::RESOLUTION_SCALE;
        y.truth_field = need(yc, "truth_field", where.c_str()).asString();
      } else fail("FitConfig::load_pdf_systematics: Unknown systematic type " + type);
      const jsonmin::Value& mean = need(yc, "mean", where.c_str());
      y.npars = mean.size();
      std::vector<double>* m = new std::vector<double>(y.npars, 0.0);
      std::vector<double>* sg = new std::vector<double>(y.npars, 0.0);
      owned.push_back(m);
      owned.push_back(sg);
      for (size_t j = 0; j < y.npars; ++j) (*m)[j] = mean[j].asDouble();
      if (yc.isMember("sigma")) {
        if (yc["sigma"].size() != y.npars) fail(where + ": 'sigma' must have as many entries as 'mean'");
        for (size_t j = 0; j < y.npars; ++j) (*sg)[j] = yc["sigma"][j].asDouble();
      }
      y.means = m->data();
      y.sigmas = sg->data();
      y.fixed = yc["fixed"].asBool(false);
      for (size_t j = 0; j < y.npars; ++j) y.pidx.push_back(pidx++);
      systematics.push_back(y);
    }
    if (sc.isMember("source")) {
      const std::string src = sc["source"].asString();
      if (!src_params.isMember(src)) fail("signal '" + sname + "': source '" + src + "' is not defined");
      bool exists = false;
      for (size_t k = 0; k < sources.size(); ++k) exists = exists || sources[k].name == src;
      if (!exists) {
        const jsonmin::Value& p = src_params[src];
        sources.push_back(Source(src, sources.size(), p.isMember("mean") ? p["mean"].asFloat() : 1.0f,
                                 p["sigma"].asFloat(0.0f), p["fixed"].asBool(false)));
      }
    } else {                                                   // the signal is a source for itself
      sources.push_back(Source(sname, sources.size(), sc.isMember("mean") ? sc["mean"].asFloat() : 1.0f,
                               sc["sigma"].asFloat(0.0f), sc["fixed"].asBool(false)));
    }
  }

  // sample columns: observables' fields, truth fields, DATASET (:153-194)
  for (size_t i = 0; i < observables.size(); ++i) {
    std::vector<std::string>::iterator it = std::find(sample_fields.begin(), sample_fields.end(), observables[i].field);
    if (it == sample_fields.end()) { sample_fields.push_back(observables[i].field); it = sample_fields.end() - 1; }
    observables[i].field_index = it - sample_fields.begin();
  }
  for (size_t i = 0; i < systematics.size(); ++i) {
    Systematic& y = systematics[i];
    const size_t index = std::find(sample_fields.begin(), sample_fields.end(), y.observable_field) - sample_fields.begin();
    if (index >= sample_fields.size())
      fail("systematic '" + y.name + "': field '" + y.observable_field + "' is not an observable of the fit");
    y.observable_field_index = index;
    if (y.type == pdfz::Systematic::RESOLUTION_SCALE) {
      std::vector<std::string>::iterator it = std::find(sample_fields.begin(), sample_fields.end(), y.truth_field);
      if (it == sample_fields.end()) { sample_fields.push_back(y.truth_field); it = sample_fields.end() - 1; }
      y.truth_field_index = it - sample_fields.begin();
    }
  }
  sample_fields.push_back("DATASET");

  // signals of the fit (:196-257 and Signal::Signal, src/signal.cpp:11-47)
  const jsonmin::Value& fsig = need(fit, "signals", "fit");
  signals.reserve(fsig.size());
  for (size_t i = 0; i < fsig.size(); ++i) {
    const std::string name = fsig[i].asString();
    if (!sig_params.isMember(name)) fail("fit.signals: '" + name + "' is not defined under signals");
    const jsonmin::Value& c = sig_params[name];
    const std::string where = "signal '" + name + "'";
    std::cout << "FitConfig: Loading signal: " << name << std::endl;
    Signal s;
    s.name = name;
    s.title = need(c, "title", where.c_str()).asString();
    s.filename = need(c, "filename", where.c_str()).asString();
    s.dataset = (unsigned)need(c, "dataset", where.c_str()).asInt();
    datasets.insert(s.dataset);
    if (c.isMember("rate") == c.isMember("scale")) fail(where + ": exactly one of 'rate' and 'scale' is required");
    // `float nexpected = -1.0 / config["scale"].asFloat()` (src/config.cpp:213-221): rounded to float
    s.nexpected = c.isMember("rate") ? c["rate"].asFloat() : (float)(-1.0 / c["scale"].asFloat());
    std::vector<Systematic> systs;
    const jsonmin::Value& slist = c["systematics"];
    for (size_t q = 0; q < slist.size(); ++q)
      for (size_t k = 0; k < systematics.size(); ++k)
        if (systematics[k].name == slist[q].asString()) systs.push_back(systematics[k]);
    const std::string source_name = c.isMember("source") ? c["source"].asString() : name;
    bool found = false;
    for (size_t k = 0; k < sources.size(); ++k)
      if (sources[k].name == source_name) { s.source = sources[k]; found = true; break; }
    if (!found) fail(where + ": no source");
    std::vector<float> table, samples_rows;
    std::vector<unsigned> rank;
    std::vector<std::string> fields;
    const std::string path = (!s.filename.empty() && s.filename[0] == '/') ? s.filename : base + s.filename;
    sxmc::io::read_float_table(path, table, rank, fields);
    const size_t n_mc = rank[0];                                            // rows before cuts (src/signal.cpp:27)
    if (s.nexpected < 0) s.nexpected *= -1.0 * n_mc;                        // 'scale' given (:32-36)
    read_dataset_to_samples(samples_rows, table, s.dataset, sample_fields, fields, cuts);
    s.build(samples_rows, (int)sample_fields.size(), observables, systs);
    s.n_mc = n_mc;
    signals.push_back(s);
  }

  // data: per dataset id a list of files, one per experiment (:259-296)
  const jsonmin::Value& dpar = root["data"];
  for (std::map<std::string, jsonmin::Value>::const_iterator it = dpar.obj.begin(); it != dpar.obj.end(); ++it) {
    const unsigned dataset = (unsigned)std::strtoul(it->first.c_str(), NULL, 10);
    std::vector<Observable> cc = observables;                   // observables act as cuts: clip to the PDF domain
    cc.insert(cc.end(), cuts.begin(), cuts.end());
    for (size_t q = 0; q < it->second.size(); ++q) {
      const jsonmin::Value& drow = it->second[q];
      Signal d;
      d.title = d.name = drow["title"].asString();
      d.filename = drow["filename"].asString();
      d.dataset = dataset;
      d.nexpected = -1.0;
      std::cout << "FitConfig: Loading data: " << d.filename << std::endl;
      std::vector<float> table, rows;
      std::vector<unsigned> rank;
      std::vector<std::string> fields;
      const std::string path = (!d.filename.empty() && d.filename[0] == '/') ? d.filename : base + d.filename;
      sxmc::io::read_float_table(path, table, rank, fields);
      read_dataset_to_samples(rows, table, dataset, sample_fields, fields, cc);
      std::vector<Systematic> none;
      d.build(rows, (int)sample_fields.size(), observables, none);
      d.n_mc = rank[0];
      data[dataset].push_back(d);
    }
  }
}

void FitConfig::print() const {                                                           // src/config.cpp:300-340
  std::cout << "Fit:" << std::endl
            << "  Fake experiments: " << nexperiments << std::endl
            << "  MCMC steps: " << nsteps << std::endl
            << "  Burn-in fraction: " << burnin_fraction << std::endl
            << "  Signal name: " << signal_name << std::endl
            << "  Output prefix: " << output_prefix << std::endl
            << "  Confidence level: " << confidence << std::endl;
  for (size_t i = 0; i < signals.size(); ++i)
    std::cout << "  " << signals[i].name << ": nexpected " << signals[i].nexpected << ", n_mc " << signals[i].n_mc
              << ", source " << signals[i].source.name << ", dataset " << signals[i].dataset << std::endl;
}

// ---- make_fake_dataset -----------------------------------------------------------------

static void check_abi(int rc) {
  if (rc == SXMC_OK) return;
  if (rc == SXMC_EINVAL) throw pdfz::Error(sxmc_last_error());
  throw std::runtime_error(sxmc_last_error());
}

std::vector<float> make_fake_dataset(std::vector<Signal>& signals, std::vector<Systematic>& systematics,
                                     std::vector<Observable>& observables, bool poisson, uint64_t seed,
                                     uint64_t experiment, std::vector<unsigned>* observed_out) {
  std::vector<double> syst_vals;
  for (size_t i = 0; i < systematics.size(); ++i)
    for (size_t j = 0; j < systematics[i].npars; ++j) syst_vals.push_back(systematics[i].means[j]);
  std::vector<float> upper, lower;
  for (size_t i = 0; i < observables.size(); ++i) { upper.push_back(observables[i].upper); lower.push_back(observables[i].lower); }
  std::vector<sxmc_hist*> hs(signals.size());
  std::vector<double> nexpected(signals.size());
  std::vector<unsigned> n_mc(signals.size());
  for (size_t i = 0; i < signals.size(); ++i) {
    hs[i] = signals[i].histogram->handle();
    nexpected[i] = signals[i].nexpected;
    n_mc[i] = (unsigned)signals[i].n_mc;
  }
  sxmc_generator* g = NULL;
  check_abi(sxmc_generator_create(hs.data(), (int)hs.size(), nexpected.data(), n_mc.data(),
                                  syst_vals.empty() ? NULL : syst_vals.data(), (int)syst_vals.size(), upper.data(),
                                  lower.data(), (int)observables.size(), &g));
  std::vector<unsigned> observed(signals.size(), 0);
  std::vector<float> events;
  size_t nfloats = 0;
  int rc = sxmc_generator_sample(g, poisson ? 1 : 0, seed, experiment, NULL, 0, observed.data(), &nfloats);
  if (rc == SXMC_OK && nfloats) {
    events.resize(nfloats);
    rc = sxmc_generator_sample(g, poisson ? 1 : 0, seed, experiment, events.data(), nfloats, observed.data(), &nfloats);
  }
  sxmc_generator_destroy(g);
  check_abi(rc);
  if (observed_out) *observed_out = observed;
  return events;
}
