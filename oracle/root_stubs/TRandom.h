// Header-only stand-in for ROOT's TRandom, used ONLY to compile the unmodified
// reference hot-path sources into oracle/_ref (test infrastructure, never shipped).
// Draws come from a "tape" installed by oracle/ref_driver.cpp so that the reference
// chain and the CUDA chain can be driven by identical random numbers.
// Call sites in the reference: src/nll_kernels.cpp:45 (Gaus), :65 (Uniform),
// src/pdfz.cpp:838 (Poisson), src/mcmc.cpp:125 (GetSeed, GPU build only).
#ifndef SXMC_STUB_TRANDOM_H
#define SXMC_STUB_TRANDOM_H
#include <cstddef>
#include <cstdlib>
#include <cmath>

struct TRandom {
  const double* normals;   // standard-normal tape, consumed by Gaus
  size_t n_normals, i_normal;
  const double* uniforms;  // U(0,1) tape, consumed by Uniform
  size_t n_uniforms, i_uniform;
  unsigned long long lcg;  // fallback generator when a tape runs out / is absent
  unsigned seed;

  TRandom() : normals(0), n_normals(0), i_normal(0), uniforms(0), n_uniforms(0),
              i_uniform(0), lcg(0x9E3779B97F4A7C15ULL), seed(0) {}

  double next_u() {
    lcg = lcg * 6364136223846793005ULL + 1442695040888963407ULL;
    return ((lcg >> 11) + 0.5) * (1.0 / 9007199254740992.0);
  }
  double Uniform() {
    if (i_uniform < n_uniforms) return uniforms[i_uniform++];
    return next_u();
  }
  double Gaus(double mean, double sigma) {
    double z;
    if (i_normal < n_normals) z = normals[i_normal++];
    else z = std::sqrt(-2.0 * std::log(next_u())) * std::cos(6.283185307179586 * next_u());
    return mean + sigma * z;
  }
  int Poisson(double mean) {  // Knuth / normal approximation; not on the parity path
    if (mean <= 0) return 0;
    if (mean > 50) {
      double v = Gaus(mean, std::sqrt(mean));
      return v < 0 ? 0 : (int)(v + 0.5);
    }
    double l = std::exp(-mean), p = 1.0; int k = 0;
    do { k++; p *= next_u(); } while (p > l);
    return k - 1;
  }
  unsigned GetSeed() const { return seed; }
  void SetSeed(unsigned s) { seed = s; lcg = 0x9E3779B97F4A7C15ULL ^ s; }
};

extern TRandom* gRandom;
#endif
