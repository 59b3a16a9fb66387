// pdfz -- the histogram-PDF evaluator class API of sxmc, kept source-compatible for its callers
// (reference: src/pdfz.h:93-574), implemented over the C ABI of libsxmc_b200 (include/sxmc_b200.h).
// No CUDA, ROOT or torch types appear here: a caller compiles this with plain g++ and links
// -lsxmc_cxx -lsxmc_b200.
//
// Differences from the reference header, all forced by the absence of ROOT on this image:
//   * CreateHistogram / CreateHistogramProjection / DefaultHistogram return a pdfz::HistogramData
//     (bin contents + axes) instead of a TH1*; the ROOT fill a maintainer keeps is two lines
//     on top of it.
//   * RandomSample draws with the device generator (Philox; sxmc_generator_*) instead of
//     TH1::GetRandom, so the events differ from a ROOT build draw by draw (same distribution).
#ifndef SXMC_B200_PDFZ_H
#define SXMC_B200_PDFZ_H

#include <cstdint>
#include <string>
#include <vector>

#include <hemi/array.h>

struct sxmc_hist;

namespace pdfz {

// src/pdfz.h:93-102 -- thrown BY VALUE
struct Error {
  Error(const std::string& _msg) : msg(_msg) {}
  std::string msg;
};

// src/pdfz.h:109-134
struct Systematic {
  enum Type { SHIFT, SCALE, RESOLUTION_SCALE, CTSCALE };
  Type type;
  Systematic(Type _type) : type(_type) {}
  virtual ~Systematic() {}
};

// x' = x + p(x)                                 src/pdfz.h:145-158
struct ShiftSystematic : public Systematic {
  ShiftSystematic(int _obs, hemi::Array<short>* _pars) : Systematic(SHIFT), obs(_obs), pars(_pars) {}
  int obs;
  hemi::Array<short>* pars;
};

// x' = x (1 + p(x))                             src/pdfz.h:168-181
struct ScaleSystematic : public Systematic {
  ScaleSystematic(int _obs, hemi::Array<short>* _pars) : Systematic(SCALE), obs(_obs), pars(_pars) {}
  int obs;
  hemi::Array<short>* pars;
};

// x' = 1 + (x - 1)(1 + p(x))                    src/pdfz.h:194-207
struct CosThetaScaleSystematic : public Systematic {
  CosThetaScaleSystematic(int _obs, hemi::Array<short>* _pars) : Systematic(CTSCALE), obs(_obs), pars(_pars) {}
  int obs;
  hemi::Array<short>* pars;
};

// x' = x + p(x) (x - x_true)                    src/pdfz.h:218-233
struct ResolutionScaleSystematic : public Systematic {
  ResolutionScaleSystematic(int _obs, int _true_obs, hemi::Array<short>* _pars)
      : Systematic(RESOLUTION_SCALE), obs(_obs), true_obs(_true_obs), pars(_pars) {}
  int obs;
  int true_obs;
  hemi::Array<short>* pars;
};

// What CreateHistogram* hand back in place of a TH1 (row-major, last observable fastest,
// src/pdfz.cpp:206-213).
struct HistogramData {
  std::vector<int> nbins;
  std::vector<double> lower, upper;
  std::vector<double> contents;      // bin counts (CreateHistogram) or densities (DefaultHistogram)
  unsigned norm;
};

// src/pdfz.h:246-395
class Eval {
 public:
  Eval(const std::vector<float>& samples, int nfields, int nobservables,
       const std::vector<double>& lower, const std::vector<double>& upper, unsigned dataset = 0);
  virtual ~Eval();

  virtual void SetEvalPoints(const std::vector<float>& points) = 0;
  virtual void SetPDFValueBuffer(hemi::Array<float>* output, int offset = 0, int stride = 1);
  virtual void SetNormalizationBuffer(hemi::Array<unsigned int>* norm, int offset = 0);
  virtual void SetParameterBuffer(hemi::Array<double>* params, int offset = 0, int stride = 1);
  virtual void AddSystematic(const Systematic& syst);
  virtual void EvalAsync(bool do_eval_pdf = true) = 0;
  virtual void EvalFinished() = 0;

  sxmc_hist* handle() const { return h_; }      // what MCMC hands to sxmc_chain_create

 protected:
  int nfields;
  int nobservables;
  std::vector<double> lower, upper;
  unsigned dataset;
  hemi::Array<float>* pdf_buffer;
  int pdf_offset, pdf_stride;
  hemi::Array<unsigned int>* norm_buffer;
  int norm_offset;
  hemi::Array<double>* param_buffer;
  int param_offset, param_stride;
  int nsystematics;
  sxmc_hist* h_;
};

// src/pdfz.h:402-574
class EvalHist : public Eval {
 public:
  EvalHist(const std::vector<float>& samples, int nfields, int nobservables,
           const std::vector<double>& lower, const std::vector<double>& upper,
           const std::vector<int>& nbins, unsigned dataset = 0, bool optimize = true);
  virtual ~EvalHist();

  virtual void SetEvalPoints(const std::vector<float>& points);
  virtual HistogramData CreateHistogram();
  virtual HistogramData CreateHistogramProjection(int observable_index);
  HistogramData DefaultHistogram();
  virtual void Optimize();
  virtual void OptimizeBin();
  virtual void OptimizeEval();
  virtual void EvalAsync(bool do_eval_pdf = true);
  virtual void EvalFinished();

  int RandomSample(std::vector<float>& events, double nexpected, std::vector<double>& syst_vals,
                   std::vector<float>& uppers, std::vector<float>& lowers, bool poisson = false,
                   unsigned dataset = 0);
  int RandomSample(std::vector<float>& events, double nexpected, bool poisson = false, unsigned dataset = 0);
  void GetSamples(std::vector<float>& sv);

  // seed of RandomSample's Philox stream (the reference draws from the global gRandom)
  static void SetRandomSeed(uint64_t seed);

 protected:
  std::vector<int> nbins;
  int total_nbins;
};

}  // namespace pdfz

#endif
