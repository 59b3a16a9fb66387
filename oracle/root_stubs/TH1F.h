#ifndef SXMC_STUB_TH1F_H
#define SXMC_STUB_TH1F_H
struct TH1F {
  double rms;
  TH1F(double r=0) : rms(r) {}
  double GetRMS() const { return rms; }
  void Delete() { delete this; }
};
#endif
