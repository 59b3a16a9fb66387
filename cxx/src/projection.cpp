// sxmc::errors::Projection::get_interval (reference: src/error_estimators/projection.cpp:14-77)
// on one column of the chain, without ROOT.  Host code, as in the reference:
//   * the 1-D projection (LikelihoodSpace::get_projection, src/likelihood.cpp:76-85 =
//     TTree::Draw("name>>_hp") into an auto-binned TH1F): THLimitsFinder's 1-2-5 bin width, 100 bins
//     requested, edges at multiples of the width;
//   * TH1::Fit("gaus", "q"): chi-square fit of A exp(-(x-mu)^2 / 2 s^2) to the bins with content,
//     error sqrt(n), started from the histogram's maximum, mean and RMS -- here a Levenberg-Marquardt
//     iteration on the three parameters;
//   * the reference's bin walk, verbatim arithmetic: a one-sided limit when less than cl/2 of the
//     probability lies at or below the mode's bin, else cl/2 to each side.
// The walk is the reference's; the axis finder and the fit restate ROOT's published behaviour
// (ROOT is not on this image: parity unpinned, compared against sxmc_b200/ensemble.py's restatement).
#include <algorithm>
#include <cmath>
#include <stdexcept>
#include <vector>

#include <sxmc/likelihood.h>

namespace sxmc {
namespace errors {

// THLimitsFinder::Optimize for one axis
void root_axis_limits(double xmin, double xmax, int nbins, double& lo, double& hi, int& nb) {
  double al = xmin, ah = xmax;
  if (al == ah) ah = al + 1.0;
  const double awidth = (ah - al) / nbins;
  int jlog = (int)std::floor(std::log10(awidth));
  const double sigfig = awidth * std::pow(10.0, -jlog);
  double sig;
  if (sigfig <= 1) sig = 1.0;
  else if (sigfig <= 2) sig = 2.0;
  else if (sigfig <= 5) sig = 5.0;
  else { sig = 1.0; jlog += 1; }
  const double bwid = sig * std::pow(10.0, jlog);
  const double lwid = std::floor(al / bwid);
  const double kwid = std::floor(ah / bwid + 1.00001);
  lo = bwid * lwid;
  hi = bwid * kwid;
  nb = (int)(kwid - lwid);
}

// chi-square Gaussian fit of (x_i, y_i +- sqrt(y_i)); returns mu
static double fit_gaus_mu(const std::vector<double>& x, const std::vector<double>& y, double a0, double mu0,
                          double s0) {
  double p[3] = {a0, mu0, s0};
  const size_t n = x.size();
  auto chi2 = [&](const double* q) {
    double c = 0;
    for (size_t i = 0; i < n; ++i) {
      const double z = (x[i] - q[1]) / q[2];
      const double r = (q[0] * std::exp(-0.5 * z * z) - y[i]) / std::sqrt(y[i]);
      c += r * r;
    }
    return c;
  };
  double lambda = 1e-3, c0 = chi2(p);
  for (int it = 0; it < 200; ++it) {
    double JtJ[3][3] = {{0}}, Jtr[3] = {0};
    for (size_t i = 0; i < n; ++i) {
      const double w = 1.0 / std::sqrt(y[i]);
      const double z = (x[i] - p[1]) / p[2];
      const double e = std::exp(-0.5 * z * z);
      const double r = (p[0] * e - y[i]) * w;
      const double J[3] = {e * w, p[0] * e * z / p[2] * w, p[0] * e * z * z / p[2] * w};
      for (int a = 0; a < 3; ++a) {
        Jtr[a] += J[a] * r;
        for (int b = 0; b < 3; ++b) JtJ[a][b] += J[a] * J[b];
      }
    }
    bool improved = false;
    for (int tries = 0; tries < 30 && !improved; ++tries) {
      double A[3][4];
      for (int a = 0; a < 3; ++a) {
        for (int b = 0; b < 3; ++b) A[a][b] = JtJ[a][b] + (a == b ? lambda * JtJ[a][a] : 0.0);
        A[a][3] = -Jtr[a];
      }
      bool singular = false;
      for (int c = 0; c < 3 && !singular; ++c) {          // Gauss-Jordan with partial pivoting
        int piv = c;
        for (int r2 = c + 1; r2 < 3; ++r2)
          if (std::fabs(A[r2][c]) > std::fabs(A[piv][c])) piv = r2;
        if (std::fabs(A[piv][c]) < 1e-300) { singular = true; break; }
        if (piv != c)
          for (int k = 0; k < 4; ++k) std::swap(A[piv][k], A[c][k]);
        for (int r2 = 0; r2 < 3; ++r2) {
          if (r2 == c) continue;
          const double f = A[r2][c] / A[c][c];
          for (int k = c; k < 4; ++k) A[r2][k] -= f * A[c][k];
        }
      }
      if (singular) { lambda *= 10; continue; }
      double q[3];
      for (int a = 0; a < 3; ++a) q[a] = p[a] + A[a][3] / A[a][a];
      if (!(q[2] > 0) || !std::isfinite(q[0]) || !std::isfinite(q[1])) { lambda *= 10; continue; }
      const double c1 = chi2(q);
      if (c1 <= c0) {
        const double gain = c0 - c1;
        for (int a = 0; a < 3; ++a) p[a] = q[a];
        c0 = c1;
        lambda = std::max(lambda * 0.3, 1e-12);
        improved = true;
        if (gain <= 1e-12 * (1.0 + c0)) return p[1];
      } else {
        lambda *= 10;
      }
    }
    if (!improved) break;
  }
  return p[1];
}

Interval projection_interval(const float* values, long n, long stride, float cl, int nbins) {
  if (n <= 0) throw std::runtime_error("Projection::get_interval: empty chain");
  double vmin = values[0], vmax = values[0];
  for (long i = 0; i < n; ++i) {
    const double v = values[i * stride];
    vmin = std::min(vmin, v);
    vmax = std::max(vmax, v);
  }
  double lo, hi;
  int nb;
  root_axis_limits(vmin, vmax, nbins, lo, hi, nb);
  const double width = (hi - lo) / nb;
  std::vector<double> content(nb + 2, 0.0);          // [0] underflow, [1..nb], [nb+1] overflow
  for (long i = 0; i < n; ++i) {
    long k = (long)std::floor(((double)values[i * stride] - lo) / width) + 1;
    k = std::max(0L, std::min(k, (long)nb + 1));
    content[k] += 1.0;
  }
  // TH1::Fit("gaus"): start from the histogram's moments, bins with content only
  std::vector<double> x, y;
  double sw = 0, sx = 0, ymax = 0;
  for (int b = 1; b <= nb; ++b) {
    const double c = lo + (b - 0.5) * width;
    sw += content[b];
    sx += content[b] * c;
    ymax = std::max(ymax, content[b]);
    if (content[b] > 0) { x.push_back(c); y.push_back(content[b]); }
  }
  double mean = 0, rms = 0;
  if (sw > 0) {
    mean = sx / sw;
    double sxx = 0;
    for (int b = 1; b <= nb; ++b) {
      const double c = lo + (b - 0.5) * width;
      sxx += content[b] * (c - mean) * (c - mean);
    }
    rms = std::sqrt(sxx / sw);
  } else {
    for (long i = 0; i < n; ++i) mean += values[i * stride];
    mean /= (double)n;
  }
  double mu = mean;
  if (x.size() >= 3 && rms > 0) mu = fit_gaus_mu(x, y, ymax, mean, rms);

  double total = 0;
  for (size_t k = 0; k < content.size(); ++k) total += content[k];     // Integral(0, -1)
  auto integral = [&](int a, int b) {
    double s = 0;
    for (int k = std::max(a, 0); k <= std::min(b, nb + 1); ++k) s += content[k];
    return b >= a ? s : 0.0;
  };
  auto find_bin = [&](double v) { return v < lo ? 0 : (v >= hi ? nb + 1 : (int)((v - lo) / width) + 1); };
  auto low_edge = [&](int i) { return lo + (i - 1) * width; };

  Interval out;
  out.cl = cl;
  int imax = find_bin(mu);
  out.point_estimate = (float)mu;
  if (imax < 1) {
    imax = 1;
    out.point_estimate = (float)low_edge(1);
  }
  int ilo = 1, ihi = 0;
  if (integral(0, imax) / total < cl / 2) {            // not enough probability to the west to go central
    out.one_sided = true;
    for (int i = 0; i < nb + 1; ++i)
      if (integral(0, i) / total >= cl) { ihi = i; break; }
  } else {
    out.one_sided = false;
    for (int i = imax; i > 0; --i)
      if (integral(i, imax) / total >= cl / 2) { ilo = i; break; }
    for (int i = imax + 1; i < nb + 1; ++i)
      if (integral(imax + 1, i) / total >= cl / 2) { ihi = i; break; }
  }
  out.coverage = (float)(integral(ilo, ihi) / total);
  out.lower = (float)low_edge(ilo);
  out.upper = (float)(low_edge(ihi) + width);
  return out;
}

}  // namespace errors
}  // namespace sxmc

// test hook (tests/test_cxx_shim.py, no GPU needed): out5 = point_estimate, lower, upper, cl, coverage
extern "C" int sxmc_cxx_projection_interval(const float* values, long n, long stride, float cl, float* out5,
                                            int* one_sided) {
  try {
    const Interval iv = sxmc::errors::projection_interval(values, n, stride, cl, 100);
    out5[0] = iv.point_estimate; out5[1] = iv.lower; out5[2] = iv.upper; out5[3] = iv.cl; out5[4] = iv.coverage;
    *one_sided = iv.one_sided ? 1 : 0;
    return 0;
  } catch (...) {
    return 1;
  }
}
