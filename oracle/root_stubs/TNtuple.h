// In-memory stand-in for ROOT's TNtuple + the "Draw(name>>hsproj)" idiom that
// src/mcmc.cpp:288-302 uses to get the per-parameter RMS of the burn-in chain.
// ROOT semantics restated: TH1::GetRMS() of an auto-ranged Draw histogram is the
// population standard deviation sqrt(|sum(x^2)/n - mean^2|) of the float entries,
// accumulated in double (TH1::GetStats -> fTsumw, fTsumwx, fTsumwx2).
#ifndef SXMC_STUB_TNTUPLE_H
#define SXMC_STUB_TNTUPLE_H
#include <string>
#include <vector>
#include <map>
#include <cmath>
#include <TH1F.h>
#include <TDirectory.h>

class TNtuple {
public:
  std::vector<std::string> names;
  std::vector<float> rows;  // row-major, names.size() floats per entry
  TNtuple(const char*, const char*, const char* varlist) {
    std::string s(varlist), cur;
    for (size_t i=0;i<=s.size();i++) {
      if (i==s.size() || s[i]==':') { names.push_back(cur); cur.clear(); }
      else cur += s[i];
    }
  }
  void Fill(const float* v) { rows.insert(rows.end(), v, v + names.size()); }
  void Reset() { rows.clear(); }
  long GetEntries() const { return (long)(rows.size() / names.size()); }
  void Draw(const char* expr) {
    std::string e(expr);
    std::string name = e.substr(0, e.find(">>"));
    size_t col = 0;
    for (; col < names.size(); col++) if (names[col] == name) break;
    double n = 0, sx = 0, sxx = 0;
    for (long i=0; i<GetEntries(); i++) {
      double x = rows[i * names.size() + col];
      n += 1; sx += x; sxx += x * x;
    }
    double rms = 0;
    if (n > 0) { double m = sx / n; rms = std::sqrt(std::fabs(sxx / n - m * m)); }
    gDirectory->objects["hsproj"] = new TH1F(rms);
  }
};
#endif
