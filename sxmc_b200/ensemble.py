"""Host-side mirror of the reference's ensemble loop (`ensemble()`, src/sxmc.cpp:41-137) and of
the two pieces either side of the chain inside it:

* ``make_fake_dataset`` (src/generator.cpp:10-48) -- device-side in libsxmc_b200
  (``sxmc_generator_*``): efficiencies and sampling histograms once per ensemble, one Poisson
  draw + CDF inversion launch per experiment;
* the ``Contour`` error estimator (src/error_estimators/contour.cpp:18-69) that turns a chain
  into the upper limit the ensemble collects (src/sxmc.cpp:110-126) -- device reductions over the
  resident chain rows (``sxmc_chain_contour``).

Experiments are independent: with ``world`` ranks, experiment ``i`` runs on rank ``i mod world``
with Philox stream id ``i`` (the reference's single gRandom stream is replaced by one
counter-based stream per experiment, so the result does not depend on the number of GPUs)."""
import ctypes as C
import math
import time
from dataclasses import dataclass
from statistics import NormalDist

import numpy as np

from . import sharding
from ._lib import c_dp, c_fp, c_up, c_vp, check, lib, ptr


@dataclass
class Interval:
    """src/interval.h:11-28"""
    one_sided: bool = False
    point_estimate: float = -1.0
    lower: float = -1.0
    upper: float = -1.0
    cl: float = -1.0
    coverage: float = -1.0


def chisquare_quantile_1dof(cl):
    """TMath::ChisquareQuantile(cl, 1) = (normal quantile of (1+cl)/2)^2."""
    z = NormalDist().inv_cdf(0.5 * (1.0 + float(cl)))
    return z * z


def contour_delta(cl):
    """`float delta = 0.5 * TMath::ChisquareQuantile(_cl, 1)` (contour.cpp:20)."""
    return float(np.float32(0.5 * chisquare_quantile_1dof(np.float32(cl))))


class FakeDataGenerator:
    """The per-ensemble half of make_fake_dataset: built once from the signals (their
    EvalHist handles, nexpected, n_mc), the systematics' means and the observables' limits."""

    def __init__(self, signals, systematics, observables):
        self.nsignals = len(signals)
        self._harr = (c_vp * self.nsignals)(*[s.histogram._h for s in signals])
        nexp = np.asarray([s.nexpected for s in signals], dtype=np.float64)
        n_mc = np.asarray([s.n_mc for s in signals], dtype=np.uint32)
        vals = np.asarray([m for y in systematics for m in y.means[:y.npars]] or [0.0], dtype=np.float64)
        nvals = sum(y.npars for y in systematics)
        up = np.asarray([o.upper for o in observables], dtype=np.float32)
        lo = np.asarray([o.lower for o in observables], dtype=np.float32)
        self.nobs = signals[0].histogram.nobservables
        self._g = c_vp()
        check(lib().sxmc_generator_create(self._harr, self.nsignals, ptr(nexp, c_dp), ptr(n_mc, c_up),
                                          ptr(vals, c_dp), int(nvals), ptr(up, c_fp), ptr(lo, c_fp),
                                          len(observables), C.byref(self._g)))
        self._expected = nexp
        self._cap = 0
        self._buf = None

    def efficiency(self):
        out = np.zeros(self.nsignals, dtype=np.float64)
        check(lib().sxmc_generator_efficiency(self._g, ptr(out, c_dp)))
        return out

    def __call__(self, poisson=True, seed=0, experiment=0):
        """-> (events float32 [n, nobs+1], observed uint32 [nsignals])"""
        observed = np.zeros(self.nsignals, dtype=np.uint32)
        nf = C.c_size_t(0)
        if self._buf is None:
            mean = float(np.sum(self._expected * self.efficiency()))
            self._buf = np.zeros(int(mean + 12 * np.sqrt(max(mean, 1.0)) + 256) * (self.nobs + 1), np.float32)
        while True:
            rc = lib().sxmc_generator_sample(self._g, int(bool(poisson)), int(seed), int(experiment),
                                             ptr(self._buf, c_fp), self._buf.size, ptr(observed, c_up),
                                             C.byref(nf))
            if rc != 0 and nf.value > self._buf.size:
                self._buf = np.zeros(nf.value, np.float32)     # rare: a > 12 sigma draw
                continue
            check(rc)
            break
        return self._buf[:nf.value].reshape(-1, self.nobs + 1).copy(), observed

    def close(self):
        if self._g:
            lib().sxmc_generator_destroy(self._g)
            self._g = c_vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_fake_dataset(signals, systematics, observables, poisson=True, seed=0, experiment=0):
    """std::vector<float> make_fake_dataset(signals, systematics, observables, poisson)
    (src/generator.cpp:10-48): flat float vector of rows (observables..., dataset id)."""
    g = FakeDataGenerator(signals, systematics, observables)
    try:
        events, _ = g(poisson, seed, experiment)
    finally:
        g.close()
    return events.ravel()


def contour_intervals(chain, cl):
    """{parameter name: Interval} from the chain's resident rows + the minimum NLL
    (LikelihoodSpace::extract_best_fit with ERROR_CONTOUR, src/likelihood.cpp:104-137)."""
    np_ = chain.nparameters
    lo = np.zeros(np_, np.float32)
    up = np.zeros(np_, np.float32)
    pt = np.zeros(np_, np.float32)
    lmin = C.c_float(0)
    nc = C.c_long(0)
    check(lib().sxmc_chain_contour(chain._c, contour_delta(cl), ptr(lo, c_fp), ptr(up, c_fp), ptr(pt, c_fp),
                                   C.byref(lmin), C.byref(nc)))
    out = {}
    for k, name in enumerate(chain.parameter_names[:-1]):
        out[name] = Interval(False, float(pt[k]), float(lo[k]), float(up[k]), float(np.float32(cl)), -999.0)
    return out, float(lmin.value)


def rows_contour(rows, cl):
    """The same estimator on host rows [n, nparameters+1] (uploaded to the device)."""
    rows = np.ascontiguousarray(rows, dtype=np.float32)
    n, w = rows.shape
    dev = c_vp()
    check(lib().sxmc_buf_alloc(rows.nbytes, C.byref(dev)))
    try:
        check(lib().sxmc_buf_upload(dev, rows.ctypes.data_as(c_vp), rows.nbytes))
        lo = np.zeros(w - 1, np.float32)
        up = np.zeros(w - 1, np.float32)
        pt = np.zeros(w - 1, np.float32)
        lmin = C.c_float(0)
        nc = C.c_long(0)
        check(lib().sxmc_rows_contour(dev, n, w - 1, contour_delta(cl), ptr(lo, c_fp), ptr(up, c_fp),
                                      ptr(pt, c_fp), C.byref(lmin), C.byref(nc)))
    finally:
        lib().sxmc_buf_free(dev)
    return lo, up, pt, float(lmin.value), int(nc.value)


def _root_axis_limits(xmin, xmax, nbins=100):
    """Histogram axis the way TTree::Draw("x>>h") picks it: THLimitsFinder::Optimize rounds the
    bin width up to 1, 2 or 5 x 10^k and the edges to multiples of it (restated from the published
    CERNLIB/ROOT algorithm; ROOT is not on this image, so this is unpinned)."""
    al, ah = float(xmin), float(xmax)
    if al == ah:
        ah = al + 1.0
    awidth = (ah - al) / nbins
    jlog = int(math.floor(math.log10(awidth)))
    sigfig = awidth * 10.0 ** (-jlog)
    if sigfig <= 1:
        sig = 1.0
    elif sigfig <= 2:
        sig = 2.0
    elif sigfig <= 5:
        sig = 5.0
    else:
        sig, jlog = 1.0, jlog + 1
    bwid = sig * 10.0 ** jlog
    lwid = math.floor(al / bwid)
    kwid = math.floor(ah / bwid + 1.00001)
    return bwid * lwid, bwid * kwid, int(kwid - lwid)


def _fit_gaus_mu(x, y, a0, mu0, s0):
    """Chi-square fit of A exp(-(x-mu)^2 / 2 s^2) to (x_i, y_i +- sqrt(y_i)) by Levenberg-Marquardt; the
    same iteration, step for step, as cxx/src/projection.cpp::fit_gaus_mu.  Returns mu."""
    p = np.array([a0, mu0, s0], dtype=np.float64)
    w = 1.0 / np.sqrt(y)

    def chi2(q):
        z = (x - q[1]) / q[2]
        r = (q[0] * np.exp(-0.5 * z * z) - y) * w
        return float(np.dot(r, r))

    lam, c0 = 1e-3, chi2(p)
    for _ in range(200):
        z = (x - p[1]) / p[2]
        e = np.exp(-0.5 * z * z)
        r = (p[0] * e - y) * w
        J = np.stack([e * w, p[0] * e * z / p[2] * w, p[0] * e * z * z / p[2] * w], axis=1)
        JtJ, Jtr = J.T @ J, J.T @ r
        improved = False
        for _ in range(30):
            A = JtJ + lam * np.diag(np.diag(JtJ))
            try:
                step = np.linalg.solve(A, -Jtr)
            except np.linalg.LinAlgError:
                lam *= 10
                continue
            q = p + step
            if not (q[2] > 0) or not np.isfinite(q[0]) or not np.isfinite(q[1]):
                lam *= 10
                continue
            c1 = chi2(q)
            if c1 <= c0:
                gain = c0 - c1
                p, c0 = q, c1
                lam = max(lam * 0.3, 1e-12)
                improved = True
                if gain <= 1e-12 * (1.0 + c0):
                    return float(p[1])
                break
            lam *= 10
        if not improved:
            break
    return float(p[1])


def projection_interval(values, cl, nbins=100):
    """sxmc::errors::Projection::get_interval (src/error_estimators/projection.cpp:14-77) on one
    column of the chain.  Host code: the 1-D projection (`LikelihoodSpace::get_projection`,
    src/likelihood.cpp:76-85 = TTree::Draw into an auto-binned TH1F), a chi-square Gaussian fit of
    its bins (TH1::Fit("gaus"): bins with content, error sqrt(n); Levenberg-Marquardt) for the mode, then the
    reference's bin walk: one-sided limit when less than cl/2 of the probability lies at or
    below the mode's bin, else cl/2 to each side.  The walk is the reference's arithmetic; the axis
    finder and the fit restate ROOT (unpinned)."""
    v = np.asarray(values, dtype=np.float32).astype(np.float64)
    if v.size == 0:
        raise ValueError("projection_interval: empty chain")
    lo, hi, nb = _root_axis_limits(v.min(), v.max(), nbins)
    width = (hi - lo) / nb
    idx = np.floor((v - lo) / width).astype(np.int64)
    content = np.zeros(nb + 2)                         # [0] underflow, [1..nb], [nb+1] overflow
    np.add.at(content, np.clip(idx + 1, 0, nb + 1), 1.0)
    centers = lo + (np.arange(nb) + 0.5) * width
    y = content[1:nb + 1]
    use = y > 0
    mean = float(np.average(centers, weights=np.maximum(y, 0))) if y.sum() > 0 else float(v.mean())
    rms = float(np.sqrt(np.average((centers - mean) ** 2, weights=y))) if y.sum() > 0 else 0.0
    mu = mean
    if use.sum() >= 3 and rms > 0:
        mu = _fit_gaus_mu(centers[use], y[use], float(y.max()), mean, rms)
    total = content.sum()                              # Integral(0, -1) includes under/overflow

    def integral(a, b):
        return content[max(a, 0):min(b, nb + 1) + 1].sum() if b >= a else 0.0

    def find_bin(x):
        return 0 if x < lo else (nb + 1 if x >= hi else int((x - lo) / width) + 1)

    low_edge = lambda i: lo + (i - 1) * width
    out = Interval(cl=float(np.float32(cl)))
    imax = find_bin(mu)
    out.point_estimate = mu
    if imax < 1:
        imax = 1
        out.point_estimate = low_edge(1)
    ilo, ihi = 1, 0
    if integral(0, imax) / total < cl / 2:
        out.one_sided = True
        for i in range(nb + 1):
            if integral(0, i) / total >= cl:
                ihi = i
                break
    else:
        out.one_sided = False
        for i in range(imax, 0, -1):
            if integral(i, imax) / total >= cl / 2:
                ilo = i
                break
        for i in range(imax + 1, nb + 1):
            if integral(imax + 1, i) / total >= cl / 2:
                ihi = i
                break
    out.coverage = float(integral(ilo, ihi) / total)
    out.lower = float(low_edge(ilo))
    out.upper = float(low_edge(ihi) + width)
    return out


def projection_intervals(rows, names, cl):
    """{parameter: Interval} with ERROR_PROJECTION (LikelihoodSpace::extract_best_fit,
    src/likelihood.cpp:104-137) from host rows [n, nparameters+1]."""
    rows = np.asarray(rows)
    return {name: projection_interval(rows[:, k], cl) for k, name in enumerate(names[:-1])}


def _fit_objects(fit, hists):
    from . import mcmc
    observables = [mcmc.Observable("o%d" % i, float(fit.lower[i]), float(fit.upper[i]))
                   for i in range(fit.nobs)]
    signals = [mcmc.Signal("sig%d" % j, hists[j], s.nexpected, s.n_mc, None) for j, s in enumerate(fit.signals)]
    systs = [mcmc.Systematic("y%d" % i, y.npars, list(y.means), list(y.sigmas), y.fixed)
             for i, y in enumerate(fit.systematics)]
    return signals, systs, observables


def ensemble(fit, nexperiments, signal_name="s0", confidence=0.9, seed=0, nsteps=None,
             burnin_fraction=None, rank=0, world=1, keep_chains=False, batch=148, debug_mode=False):
    """The loop of src/sxmc.cpp:62-134 over this rank's experiments.  Returns a dict with the
    experiment indices run here, their upper limits (`limits.push_back(bfi.upper)`, :126),
    per-experiment observed counts and timings.  Use ``sharding.gather_limits`` +
    ``sharding.median`` for the ensemble-wide result (:218-226).

    ``batch`` > 1: the fits of up to that many experiments advance together in one cooperative
    launch (``MCMCBatch``) when the fit is small enough for its samples to stay on chip;
    otherwise (or with batch <= 1) one chain at a time.  Both give the same chains: experiment
    ``i`` always uses Philox stream ``i``.  ``debug_mode`` is MCMC::operator()'s flag
    (`mcmc(data, nsteps, burnin, debug_mode)`, src/sxmc.cpp:97): every step accepted, no reset of the
    chain at burn-in."""
    from . import mcmc
    from ._lib import SxmcError
    nsteps = fit.nsteps if nsteps is None else int(nsteps)
    burnin = fit.burnin_fraction if burnin_fraction is None else float(burnin_fraction)
    mine = sharding.experiments_for_rank(nexperiments, rank, world)
    out = {"experiments": mine, "limits": [], "observed": [], "chains": [], "loop_ms": 0.0,
           "seconds": 0.0, "mode": "single", "launches": 0}
    delta = contour_delta(confidence)
    t_all = time.perf_counter()
    if batch > 1 and mine:
        tb = time.perf_counter()
        hists, mb = mcmc.build_fit(fit, batch=min(int(batch), 148))
        gen = FakeDataGenerator(*_fit_objects(fit, hists))
        out["phases_s"] = {"build": time.perf_counter() - tb, "generate": 0.0, "run": 0.0, "limits": 0.0}
        try:
            k = mb.parameter_names.index(signal_name) if signal_name in mb.parameter_names[:-1] else -1
            for c0 in range(0, len(mine), mb.max_chains):
                group = mine[c0:c0 + mb.max_chains]
                obs = []
                tg = time.perf_counter()
                for slot, i in enumerate(group):
                    obs.append(mb.generate_data(slot, gen, True, seed, i))   # make_fake_dataset(..., true)
                    mb.set_seed(slot, seed, i)
                tr = time.perf_counter()
                mb.run(len(group), nsteps, burnin, bool(debug_mode))                           # MCMC::operator() x len(group)
                tl = time.perf_counter()
                out["loop_ms"] += mb.loop_ms()
                out["launches"] += 1
                for slot, i in enumerate(group):
                    _, up, _, _, _ = mb.contour(slot, delta)                 # ls->get_best_fit()
                    out["limits"].append(float(up[k]) if k >= 0 else float("nan"))
                    out["observed"].append(obs[slot])
                    if keep_chains:
                        out["chains"].append(mb.rows(slot))
                te = time.perf_counter()
                out["phases_s"]["generate"] += tr - tg
                out["phases_s"]["run"] += tl - tr
                out["phases_s"]["limits"] += te - tl
            out["mode"] = "batch"
            out["seconds"] = time.perf_counter() - t_all
            return out
        except SxmcError as e:
            if e.code != 5 or out["limits"]:      # only SXMC_ENOTSMALL falls back; timeouts and aborts propagate
                raise
        finally:
            gen.close()
            mb.close()
            for h in hists:
                h.close()
    hists, chain = mcmc.build_fit(fit)
    gen = FakeDataGenerator(*_fit_objects(fit, hists))
    for i in mine:
        events, observed = gen(True, seed, i)                   # make_fake_dataset(..., true)
        chain.set_seed(seed, i)
        chain.run_resident(events, nsteps, burnin, bool(debug_mode))              # MCMC::operator(), rows stay on the device
        best_fit, _ = contour_intervals(chain, confidence)      # ls->get_best_fit()
        out["limits"].append(best_fit[signal_name].upper if signal_name in best_fit else float("nan"))
        out["observed"].append(observed)
        out["loop_ms"] += chain.loop_ms()
        out["launches"] += chain.launches()
        if keep_chains:
            out["chains"].append(chain.rows())
    gen.close()
    chain.close()
    for h in hists:
        h.close()
    out["seconds"] = time.perf_counter() - t_all
    return out
