// Header-only stand-in for ROOT histograms (only what src/pdfz.cpp names:
// CreateHistogram at src/pdfz.cpp:498-594 and RandomSample at :817-922).
// Test infrastructure for oracle/_ref only.
#ifndef SXMC_STUB_TH1_H
#define SXMC_STUB_TH1_H
#include <vector>
#include <cmath>
#include <TRandom.h>

struct TClassStub { int id; };

class TH1 {
public:
  int nx, ny, nz; double xlo, xhi, ylo, yhi, zlo, zhi;
  std::vector<double> content, error;
  TH1(int _nx, double _xlo, double _xhi, int _ny=1, double _ylo=0, double _yhi=1,
      int _nz=1, double _zlo=0, double _zhi=1)
    : nx(_nx), ny(_ny), nz(_nz), xlo(_xlo), xhi(_xhi), ylo(_ylo), yhi(_yhi),
      zlo(_zlo), zhi(_zhi), content((size_t)(_nx+2)*(_ny+2)*(_nz+2), 0.0),
      error((size_t)(_nx+2)*(_ny+2)*(_nz+2), 0.0) {}
  virtual ~TH1() {}
  void Sumw2() {}
  virtual int GetBin(int x, int y=0, int z=0) const { return x + (nx+2) * (y + (ny+2) * z); }
  void SetBinContent(int b, double v) { content[b] = v; }
  void SetBinError(int b, double v) { error[b] = v; }
  double GetBinContent(int b) const { return content[b]; }
  int GetNbinsX() const { return nx; }
  int GetNbinsY() const { return ny; }
  int GetNbinsZ() const { return nz; }
  virtual const TClassStub* IsA() const = 0;
  virtual const char* ClassName() const = 0;
};

class TH1D : public TH1 {
public:
  TH1D(const char*, const char*, int n, double lo, double hi) : TH1(n, lo, hi) {}
  int GetBin(int x, int y=0, int z=0) const { return x; }
  static const TClassStub* Class() { static TClassStub c = {1}; return &c; }
  const TClassStub* IsA() const { return Class(); }
  const char* ClassName() const { return "TH1D"; }
  double GetRandom() const {  // inverse-CDF over bins + uniform inside the bin
    double tot = 0; for (int i=1;i<=nx;i++) tot += content[i];
    double r = gRandom->Uniform() * tot, acc = 0;
    for (int i=1;i<=nx;i++) { acc += content[i];
      if (r < acc || i == nx) return xlo + (xhi-xlo) * ((i-1) + gRandom->Uniform()) / nx; }
    return xlo;
  }
};
#endif
