#ifndef SXMC_STUB_TSTOPWATCH_H
#define SXMC_STUB_TSTOPWATCH_H
#include <chrono>
struct TStopwatch {
  std::chrono::steady_clock::time_point t0, t1;
  void Start() { t0 = std::chrono::steady_clock::now(); }
  void Stop() { t1 = std::chrono::steady_clock::now(); }
  double RealTime() { Stop(); return std::chrono::duration<double>(t1 - t0).count(); }
};
#endif
