#ifndef SXMC_STUB_TH3D_H
#define SXMC_STUB_TH3D_H
#include <TH1.h>
class TH3D : public TH1 {
public:
  TH3D(const char*, const char*, int n, double lo, double hi, int m, double ylo_, double yhi_,
       int l, double zlo_, double zhi_) : TH1(n, lo, hi, m, ylo_, yhi_, l, zlo_, zhi_) {}
  static const TClassStub* Class() { static TClassStub c = {3}; return &c; }
  const TClassStub* IsA() const { return Class(); }
  const char* ClassName() const { return "TH3D"; }
  void GetRandom3(double& x, double& y, double& z) const {
    double tot = 0;
    for (int k=1;k<=nz;k++) for (int j=1;j<=ny;j++) for (int i=1;i<=nx;i++) tot += content[GetBin(i,j,k)];
    double r = gRandom->Uniform() * tot, acc = 0;
    for (int k=1;k<=nz;k++) for (int j=1;j<=ny;j++) for (int i=1;i<=nx;i++) { acc += content[GetBin(i,j,k)];
      if (r < acc || (i==nx && j==ny && k==nz)) {
        x = xlo + (xhi-xlo) * ((i-1) + gRandom->Uniform()) / nx;
        y = ylo + (yhi-ylo) * ((j-1) + gRandom->Uniform()) / ny;
        z = zlo + (zhi-zlo) * ((k-1) + gRandom->Uniform()) / nz; return; } }
  }
};
#endif
