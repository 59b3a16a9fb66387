// FitConfig -- the JSON fit configuration (reference: src/config.h:24-77, src/config.cpp:19-297),
// same public fields.  Sample files are float tables instead of ROOT trees (no ROOT on this
// image): `.npy` ([n, k] float32/float64, C order) or `.f32` (flat little-endian float32), each
// with a `<file>.fields` text file naming the columns, or `.csv` with a header line.
#ifndef SXMC_B200_CONFIG_H
#define SXMC_B200_CONFIG_H
#include <map>
#include <set>
#include <string>
#include <vector>

#include <sxmc/likelihood.h>
#include <sxmc/observable.h>
#include <sxmc/signal.h>
#include <sxmc/systematic.h>

namespace sxmc {
namespace io {
// stands in for read_float_vector_ttree (src/io/ttree_io.cpp:21-159): row-major table + column names;
// throws std::runtime_error naming the file
void read_float_table(const std::string& filename, std::vector<float>& data, std::vector<unsigned>& rank,
                      std::vector<std::string>& fields);
}  // namespace io
}  // namespace sxmc

class FitConfig {
 public:
  FitConfig(std::string filename);
  virtual ~FitConfig();
  void print() const;

  unsigned nexperiments;
  unsigned nsteps;
  float confidence;
  float burnin_fraction;
  bool debug_mode;
  bool plots;
  long seed;
  std::string signal_name;
  std::string output_prefix;
  std::vector<Signal> signals;
  std::vector<Systematic> systematics;
  std::vector<Observable> observables;
  std::vector<Observable> cuts;
  std::vector<Source> sources;
  std::map<unsigned, std::vector<Signal> > data;
  std::vector<std::string> sample_fields;
  std::set<unsigned> datasets;
  std::string samples;
  ErrorType error_type;

 private:
  std::vector<std::vector<double>*> owned;     // the systematics' means / sigmas arrays
};

// Signal::read_dataset_to_samples (src/signal.cpp:50-108)
void read_dataset_to_samples(std::vector<float>& samples, std::vector<float>& dataset, unsigned dataset_id,
                             std::vector<std::string>& sample_fields, std::vector<std::string>& dataset_fields,
                             std::vector<Observable>& cuts);

// make_fake_dataset (src/generator.cpp:10-48) for experiment `experiment` of an ensemble: every
// signal's Poisson count and events in one device launch, Philox stream = experiment index.
std::vector<float> make_fake_dataset(std::vector<Signal>& signals, std::vector<Systematic>& systematics,
                                     std::vector<Observable>& observables, bool poisson, uint64_t seed,
                                     uint64_t experiment, std::vector<unsigned>* observed = NULL);
#endif
