"""Seeded fits shared by the oracle tests, the GPU parity tests and
tests/golden/make_ref_fixtures.py: the systematic arms the BASELINE workloads do not reach
(`apply_systematic`, reference src/pdfz.cpp:306-331): CTSCALE and 2-, 3- and 4-coefficient
series p = sum_i par_i * pow(x, i)."""
import hashlib

import numpy as np

from sxmc_b200 import workloads as W


def _sub(base, nsig, systs, name, nevents=5000):
    return W.FitSpec(name=name, lower=base.lower, upper=base.upper, nbins=base.nbins,
                     source_means=base.source_means[:nsig], source_sigmas=base.source_sigmas[:nsig],
                     source_fixed=base.source_fixed[:nsig], systematics=systs,
                     signals=[W.SignalSpec(samples=s.samples, nexpected=s.nexpected, source=j)
                              for j, s in enumerate(base.signals[:nsig])], data=base.data[:nevents])


def fit_ctscale():
    """3-D, ctscale(costheta) + shift(radius): single-coefficient program outside the kernel table."""
    base = W.config_D(0.002)
    systs = [W.SystSpec(type=W.CTSCALE, obs=2, means=[0.01], sigmas=[0.01]),
             W.SystSpec(type=W.SHIFT, obs=1, means=[0.0], sigmas=[0.01])]
    return _sub(base, 4, systs, "CT")


def fit_poly(ncoef):
    """3-D with one `ncoef`-coefficient series per systematic type + a 1-coefficient ctscale."""
    base = W.config_D(0.002)
    m = [0.01, 0.002, 0.0003, 0.00004][:ncoef]
    s = [0.01, 0.001, 0.0001, 0.00001][:ncoef]
    systs = [W.SystSpec(type=W.SHIFT, obs=0, means=list(m), sigmas=list(s)),
             W.SystSpec(type=W.SCALE, obs=1, means=[0.5 * v for v in m], sigmas=list(s)),
             W.SystSpec(type=W.CTSCALE, obs=2, means=[0.01], sigmas=[0.01]),
             W.SystSpec(type=W.RESOLUTION_SCALE, obs=0, extra=3, means=[2 * v for v in m], sigmas=list(s))]
    return _sub(base, 4, systs, "P%d" % ncoef)


POLY_CASES = {"CT": fit_ctscale, "P2": lambda: fit_poly(2), "P3": lambda: fit_poly(3),
              "P4": lambda: fit_poly(4)}


def off_nominal(fit, seed):
    """Seeded parameter vector a few sigma off the means (every systematic arm does work)."""
    rng = np.random.default_rng(seed)
    p = fit.parameter_means().copy()
    p[:fit.nsources] = rng.uniform(0.5, 1.5, fit.nsources)
    sig = fit.parameter_sigmas()
    p[fit.nsources:] += rng.normal(0, 1, fit.nparameters - fit.nsources) * 3 * sig[fit.nsources:]
    return p


# full-size BASELINE shapes: (name, constructor).  Their histograms are committed as SHA-256
# digests (+ norms, NLL, event sum) computed by the compiled reference.
FULL_CASES = {
    "Cfull": lambda: W.config_C(1.0),
    "Dfull": lambda: W.config_D(1.0),
    "Dstress": lambda: W.config_D(1.0, stress=True),     # 100 x 50 x 40 = 200 000 bins
}


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
