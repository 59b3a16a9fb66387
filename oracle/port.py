"""ctypes face of oracle/sxmc_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py)."""
import ctypes as C
import os

import numpy as np

from . import HERE, build, have_port

_lib = None

c_dp = C.POINTER(C.c_double)
c_fp = C.POINTER(C.c_float)
c_ip = C.POINTER(C.c_int)
c_up = C.POINTER(C.c_uint)
c_sp = C.POINTER(C.c_short)


class _Signal(C.Structure):
    _fields_ = [("samples", c_fp), ("nsamples", C.c_long), ("nfields", C.c_int),
                ("dataset", C.c_uint), ("nsyst", C.c_int), ("stype", c_sp), ("sobs", c_sp),
                ("sextra", c_sp), ("snpars", c_sp), ("spars", c_sp)]


def lib():
    global _lib
    if _lib is None:
        if not have_port():
            build()
        _lib = C.CDLL(os.path.join(HERE, "liboracle_port.so"))
        _lib.oracle_bin_volume.restype = C.c_double
        _lib.oracle_nll.restype = C.c_double
        _lib.oracle_nll_total.restype = C.c_double
        _lib.oracle_mcmc_run.restype = C.c_long
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


def syst_arrays(fit, sig):
    """Flat SystematicDescriptor arrays for one signal (global coefficient indices)."""
    ss = [fit.systematics[i] for i in sig.systematics]
    stype = np.asarray([s.type for s in ss], dtype=np.int16)
    sobs = np.asarray([s.obs for s in ss], dtype=np.int16)
    sextra = np.asarray([s.extra for s in ss], dtype=np.int16)
    snpars = np.asarray([s.npars for s in ss], dtype=np.int16)
    spars = np.asarray([p for s in ss for p in s.pidx], dtype=np.int16)
    # keep non-empty buffers so .ctypes pointers are valid
    pad = lambda a: a if a.size else np.zeros(1, dtype=np.int16)
    return len(ss), pad(stype), pad(sobs), pad(sextra), pad(snpars), pad(spars)


def _geom(fit):
    lo = np.asarray(fit.lower, dtype=np.float64)
    hi = np.asarray(fit.upper, dtype=np.float64)
    nb = np.asarray(fit.nbins, dtype=np.int32)
    return lo, hi, nb


def bin_volume(fit):
    lo, hi, nb = _geom(fit)
    return lib().oracle_bin_volume(fit.nobs, _p(lo, c_dp), _p(hi, c_dp), _p(nb, c_ip))


def bin_signal(fit, j, syst_params):
    """zero + transform + bin for signal j.  Returns (bins u32[total_nbins], norm)."""
    sig = fit.signals[j]
    lo, hi, nb = _geom(fit)
    n, stype, sobs, sextra, snpars, spars = syst_arrays(fit, sig)
    params = np.ascontiguousarray(syst_params, dtype=np.float64)
    if params.size == 0:
        params = np.zeros(1)
    bins = np.zeros(fit.total_nbins, dtype=np.uint32)
    norm = C.c_uint(0)
    noob = C.c_long(0)
    lib().oracle_bin_samples(_p(sig.samples, c_fp), C.c_long(sig.samples.shape[0]),
                             sig.samples.shape[1], fit.nobs, _p(lo, c_dp), _p(hi, c_dp),
                             _p(nb, c_ip), n, _p(stype, c_sp), _p(sobs, c_sp),
                             _p(sextra, c_sp), _p(snpars, c_sp), _p(spars, c_sp),
                             _p(params, c_dp), 1, _p(bins, c_up), C.byref(norm),
                             C.byref(noob))
    return bins, int(norm.value)


def read_bins(fit, dataset=0, data=None):
    data = fit.data if data is None else np.ascontiguousarray(data, dtype=np.float32)
    lo, hi, nb = _geom(fit)
    out = np.zeros(max(1, data.shape[0]), dtype=np.int32)
    lib().oracle_set_eval_points(_p(data, c_fp), C.c_long(data.shape[0]), fit.nobs,
                                 _p(lo, c_dp), _p(hi, c_dp), _p(nb, c_ip),
                                 C.c_uint(dataset), _p(out, c_ip))
    return out[:data.shape[0]]


def eval_pdf(fit, rb, bins, norm):
    out = np.zeros(max(1, rb.shape[0]), dtype=np.float32)
    rb = np.ascontiguousarray(rb, dtype=np.int32)
    bins = np.ascontiguousarray(bins, dtype=np.uint32)
    lib().oracle_eval_pdf(C.c_long(rb.shape[0]), _p(rb, c_ip), _p(bins, c_up),
                          C.c_uint(norm), C.c_double(bin_volume(fit)), _p(out, c_fp), 1)
    return out[:rb.shape[0]]


def eval_all(fit, pars):
    """All signals at parameter vector `pars` (length nparameters).
    Returns bins [ns, total_nbins] u32, norms [ns] u32, lut [ns, nevents] f32."""
    pars = np.asarray(pars, dtype=np.float64)
    ns, ne = fit.nsignals, fit.nevents
    bins = np.zeros((ns, fit.total_nbins), dtype=np.uint32)
    norms = np.zeros(ns, dtype=np.uint32)
    lut = np.zeros((ns, ne), dtype=np.float32)
    for j, sig in enumerate(fit.signals):
        bins[j], norms[j] = bin_signal(fit, j, pars[fit.nsources:])
        lut[j] = eval_pdf(fit, read_bins(fit, sig.dataset), bins[j], int(norms[j]))
    return bins, norms, lut


def _nll_consts(fit):
    nexp = np.asarray([s.nexpected for s in fit.signals], dtype=np.float64)
    n_mc = np.asarray([s.n_mc for s in fit.signals], dtype=np.uint32)
    sid = np.asarray([s.source for s in fit.signals], dtype=np.int16)
    return nexp, n_mc, sid


def nll(fit, lut, pars, norms):
    """Returns (nll, event_sum)."""
    lut = np.ascontiguousarray(lut, dtype=np.float32)
    pars = np.ascontiguousarray(pars, dtype=np.float64)
    norms = np.ascontiguousarray(norms, dtype=np.uint32)
    nexp, n_mc, sid = _nll_consts(fit)
    means, sigmas = fit.parameter_means(), fit.parameter_sigmas()
    ev = C.c_double(0)
    v = lib().oracle_nll(_p(lut, c_fp), C.c_size_t(fit.nevents), C.c_size_t(fit.nsignals),
                         _p(pars, c_dp), C.c_size_t(fit.nparameters),
                         C.c_size_t(fit.nsources), _p(means, c_dp), _p(sigmas, c_dp),
                         _p(nexp, c_dp), _p(n_mc, c_up), _p(sid, c_sp), _p(norms, c_up),
                         C.byref(ev))
    return float(v), float(ev.value)


def mcmc(fit, nsteps, burnin_fraction, normals, uniforms, debug_mode=False, trace=False):
    """Full chain.  Returns dict(chain [rows, np+1] f32, accepted, nll_trace, prop_trace)."""
    lo, hi, nb = _geom(fit)
    keep = []
    sigs = (_Signal * fit.nsignals)()
    for j, sig in enumerate(fit.signals):
        n, stype, sobs, sextra, snpars, spars = syst_arrays(fit, sig)
        keep += [stype, sobs, sextra, snpars, spars]
        sigs[j] = _Signal(_p(sig.samples, c_fp), sig.samples.shape[0], sig.samples.shape[1],
                          sig.dataset, n, _p(stype, c_sp), _p(sobs, c_sp), _p(sextra, c_sp),
                          _p(snpars, c_sp), _p(spars, c_sp))
    nexp, n_mc, sid = _nll_consts(fit)
    means, sigmas, fixed = fit.parameter_means(), fit.parameter_sigmas(), fit.parameter_fixed()
    normals = np.ascontiguousarray(normals, dtype=np.float64)
    uniforms = np.ascontiguousarray(uniforms, dtype=np.float64)
    npar = fit.nparameters
    chain = np.zeros((max(1, nsteps), npar + 1), dtype=np.float32)
    nll_tr = np.zeros(max(1, nsteps), dtype=np.float64)
    prop_tr = np.zeros((max(1, nsteps), npar), dtype=np.float64)
    acc = C.c_long(0)
    rows = lib().oracle_mcmc_run(
        fit.nsources, npar, _p(means, c_dp), _p(sigmas, c_dp), _p(fixed, c_ip),
        int(fit.systematics_fixed), fit.nobs, _p(lo, c_dp), _p(hi, c_dp), _p(nb, c_ip),
        fit.nsignals, sigs, _p(nexp, c_dp), _p(n_mc, c_up), _p(sid, c_sp),
        _p(fit.data, c_fp), C.c_long(fit.nevents), C.c_uint(nsteps),
        C.c_float(burnin_fraction), int(debug_mode),
        _p(normals, c_dp), C.c_size_t(normals.size), _p(uniforms, c_dp),
        C.c_size_t(uniforms.size), _p(chain, c_fp), C.c_long(chain.shape[0]),
        _p(nll_tr, c_dp) if trace else None, _p(prop_tr, c_dp) if trace else None,
        C.byref(acc))
    return dict(chain=chain[:rows].copy(), accepted=int(acc.value),
                nll_trace=nll_tr[:nsteps] if trace else None,
                prop_trace=prop_tr[:nsteps] if trace else None)


# ---------------------------------------------------------------------------------------
# Contour error estimator -- numpy restatement.  PARITY UNPINNED: the reference's code for it
# runs on ROOT's TNtuple/TTreeFormula (un-vendored) and its tests hold no vector for it.
# ---------------------------------------------------------------------------------------
def _through_ostream(v):
    """A float written with `ostream <<` (6 significant digits) and parsed back by TTreeFormula."""
    return float("%g" % float(np.float32(v)))


def contour(rows, delta):
    """rows: float32 [n, np+1], last column = NLL.  Follows LikelihoodSpace::get_contour
    (src/likelihood.cpp:88-101) and Contour::get_interval (src/error_estimators/contour.cpp:30-66).
    Returns (lower[np], upper[np], point[np], lmin, ncontour)."""
    rows = np.asarray(rows, dtype=np.float32)
    nll = rows[:, -1].astype(np.float64)
    lmin = np.float32(rows[:, -1].min())                       # GetMinimum("likelihood")
    off = _through_ostream(-lmin)
    sel = nll + off < _through_ostream(delta)                  # "likelihood+<-lmin><<delta>"
    pts = rows[sel]
    assert len(pts) > 0
    lmin2 = np.float32(pts[:, -1].min())
    off2 = _through_ostream(-lmin2)
    dnll = np.float32(0.13)
    while True:
        el = pts[pts[:, -1].astype(np.float64) + off2 < _through_ostream(dnll)]
        dnll = np.float32(dnll * np.float32(5))
        if len(el) >= 1:
            break
    npar = rows.shape[1] - 1
    lower = pts[:, :npar].min(axis=0)
    upper = pts[:, :npar].max(axis=0)
    point = ((el[:, :npar].min(axis=0).astype(np.float64) + el[:, :npar].max(axis=0).astype(np.float64)) / 2
             ).astype(np.float32)
    return lower, upper, point, float(lmin), int(len(pts))
