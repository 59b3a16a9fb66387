#ifndef SXMC_STUB_TH2D_H
#define SXMC_STUB_TH2D_H
#include <TH1.h>
class TH2D : public TH1 {
public:
  TH2D(const char*, const char*, int n, double lo, double hi, int m, double ylo_, double yhi_)
    : TH1(n, lo, hi, m, ylo_, yhi_) {}
  int GetBin(int x, int y=0, int z=0) const { return x + (nx+2) * y; }
  static const TClassStub* Class() { static TClassStub c = {2}; return &c; }
  const TClassStub* IsA() const { return Class(); }
  const char* ClassName() const { return "TH2D"; }
  void GetRandom2(double& x, double& y) const {
    double tot = 0;
    for (int j=1;j<=ny;j++) for (int i=1;i<=nx;i++) tot += content[GetBin(i,j)];
    double r = gRandom->Uniform() * tot, acc = 0;
    for (int j=1;j<=ny;j++) for (int i=1;i<=nx;i++) { acc += content[GetBin(i,j)];
      if (r < acc || (i==nx && j==ny)) {
        x = xlo + (xhi-xlo) * ((i-1) + gRandom->Uniform()) / nx;
        y = ylo + (yhi-ylo) * ((j-1) + gRandom->Uniform()) / ny; return; } }
  }
};
#endif
