"""ctypes face of oracle/_ref/libsxmc_ref.so (the UNMODIFIED reference sources, CPU build)
-- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Same call shapes as oracle.port."""
import ctypes as C
import os

import numpy as np

from . import HERE, build, have_ref
from .port import c_dp, c_fp, c_ip, c_up, c_sp, _p, _nll_consts

_lib = None


class RefError(Exception):
    """A pdfz::Error thrown by the reference (message preserved)."""


def lib():
    global _lib
    if _lib is None:
        if not have_ref():
            build()
        if not have_ref():
            raise RuntimeError("oracle/_ref/libsxmc_ref.so is not built and /root/reference "
                               "is absent")
        _lib = C.CDLL(os.path.join(HERE, "_ref", "libsxmc_ref.so"))
        _lib.ref_hist_create.restype = C.c_void_p
        _lib.ref_hist_bin_volume.restype = C.c_double
        _lib.ref_hist_read_bins.restype = C.c_size_t
        _lib.ref_nll.restype = C.c_double
        _lib.ref_mcmc_run.restype = C.c_long
        _lib.ref_tape_normals_used.restype = C.c_size_t
        _lib.ref_tape_uniforms_used.restype = C.c_size_t
    return _lib


class Hist:
    """One reference pdfz::EvalHist (src/pdfz.cpp:177-236)."""

    def __init__(self, samples, nfields, nobs, lower, upper, nbins, dataset=0):
        s = np.ascontiguousarray(samples, dtype=np.float32).ravel()
        lo = np.asarray(lower, dtype=np.float64)
        hi = np.asarray(upper, dtype=np.float64)
        nb = np.asarray(nbins, dtype=np.int32)
        err = C.create_string_buffer(512)
        pad = lambda a, t: a if a.size else np.zeros(1, dtype=t)
        self.h = lib().ref_hist_create(_p(pad(s, np.float32), c_fp), C.c_size_t(s.size),
                                       int(nfields), int(nobs),
                                       _p(pad(lo, np.float64), c_dp), lo.size,
                                       _p(pad(hi, np.float64), c_dp), hi.size,
                                       _p(pad(nb, np.int32), c_ip), nb.size,
                                       C.c_uint(dataset), err, 512)
        if not self.h:
            raise RefError(err.value.decode())
        self.h = C.c_void_p(self.h)
        self.nobs = nobs

    def close(self):
        if self.h:
            lib().ref_hist_destroy(self.h)
            self.h = None

    def add_systematic(self, type_, obs, extra, pars):
        pars = np.asarray(pars, dtype=np.int16)
        rc = lib().ref_hist_add_systematic(self.h, type_, obs, extra, pars.size, _p(pars, c_sp))
        if rc:
            raise RefError("AddSystematic failed (%d)" % rc)

    def set_eval_points(self, points):
        p = np.ascontiguousarray(points, dtype=np.float32).ravel()
        err = C.create_string_buffer(512)
        pp = p if p.size else np.zeros(1, np.float32)
        if lib().ref_hist_set_eval_points(self.h, _p(pp, c_fp), C.c_size_t(p.size), err, 512):
            raise RefError(err.value.decode())

    def eval(self, params, param_offset=0, param_stride=1, do_eval_pdf=True,
             pdf_io=None, pdf_offset=0, pdf_stride=1, norm_io=None, norm_offset=0):
        """EvalAsync + EvalFinished.  Returns (pdf array, norm array) incl. untouched slots."""
        params = np.ascontiguousarray(params, dtype=np.float64)
        npts = self.n_read_bins()
        if pdf_io is None:
            pdf_io = np.zeros(max(1, pdf_offset + pdf_stride * npts), dtype=np.float32)
        if norm_io is None:
            norm_io = np.zeros(norm_offset + 1, dtype=np.uint32)
        pdf_io = np.ascontiguousarray(pdf_io, dtype=np.float32)
        norm_io = np.ascontiguousarray(norm_io, dtype=np.uint32)
        pp = params if params.size else np.zeros(1)
        lib().ref_hist_eval(self.h, _p(pp, c_dp), params.size, param_offset, param_stride,
                            int(do_eval_pdf), _p(pdf_io, c_fp), C.c_size_t(pdf_io.size),
                            pdf_offset, pdf_stride, _p(norm_io, c_up), norm_io.size,
                            norm_offset)
        return pdf_io, norm_io

    def total_nbins(self):
        return lib().ref_hist_total_nbins(self.h)

    def bin_volume(self):
        return lib().ref_hist_bin_volume(self.h)

    def bins(self):
        out = np.zeros(self.total_nbins(), dtype=np.uint32)
        lib().ref_hist_bins(self.h, _p(out, c_up))
        return out

    def n_read_bins(self):
        return int(lib().ref_hist_read_bins(self.h, None))

    def read_bins(self):
        out = np.zeros(max(1, self.n_read_bins()), dtype=np.int32)
        n = lib().ref_hist_read_bins(self.h, _p(out, c_ip))
        return out[:n]


def make_hist(fit, j):
    sig = fit.signals[j]
    h = Hist(sig.samples, sig.samples.shape[1], fit.nobs, fit.lower, fit.upper, fit.nbins,
             sig.dataset)
    for i in sig.systematics:
        s = fit.systematics[i]
        h.add_systematic(s.type, s.obs, s.extra, s.pidx)
    return h


def eval_all(fit, pars):
    """Same contract as oracle.port.eval_all, computed by the reference's own code."""
    pars = np.asarray(pars, dtype=np.float64)
    ns, ne = fit.nsignals, fit.nevents
    bins = np.zeros((ns, fit.total_nbins), dtype=np.uint32)
    norms = np.zeros(ns, dtype=np.uint32)
    lut = np.zeros((ns, ne), dtype=np.float32)
    for j in range(ns):
        h = make_hist(fit, j)
        h.set_eval_points(fit.data)
        pdf, nrm = h.eval(pars, param_offset=fit.nsources)
        bins[j], norms[j], lut[j] = h.bins(), nrm[0], pdf[:ne]
        h.close()
    return bins, norms, lut


def nll(fit, lut, pars, norms):
    lut = np.ascontiguousarray(lut, dtype=np.float32)
    pars = np.ascontiguousarray(pars, dtype=np.float64)
    norms = np.ascontiguousarray(norms, dtype=np.uint32)
    nexp, n_mc, sid = _nll_consts(fit)
    means, sigmas = fit.parameter_means(), fit.parameter_sigmas()
    ev = C.c_double(0)
    v = lib().ref_nll(_p(lut, c_fp), C.c_size_t(fit.nevents), C.c_size_t(fit.nsignals),
                      _p(pars, c_dp), C.c_size_t(fit.nparameters), C.c_size_t(fit.nsources),
                      _p(means, c_dp), _p(sigmas, c_dp), _p(nexp, c_dp), _p(n_mc, c_up),
                      _p(sid, c_sp), _p(norms, c_up), C.byref(ev))
    return float(v), float(ev.value)


def mcmc(fit, nsteps, burnin_fraction, normals, uniforms, debug_mode=False,
         sync_interval=10000, verbose=False, hists=None):
    """MCMC::operator() of the reference, random numbers taken from the tapes."""
    own = hists is None
    if own:
        hists = [make_hist(fit, j) for j in range(fit.nsignals)]
    normals = np.ascontiguousarray(normals, dtype=np.float64)
    uniforms = np.ascontiguousarray(uniforms, dtype=np.float64)
    lib().ref_set_tapes(_p(normals, c_dp), C.c_size_t(normals.size), _p(uniforms, c_dp),
                        C.c_size_t(uniforms.size), 1)
    f32 = lambda v: np.asarray(v, dtype=np.float32)
    i32 = lambda v: np.asarray(v, dtype=np.int32)
    smean, ssig, sfix = f32(fit.source_means), f32(fit.source_sigmas), i32(fit.source_fixed)
    ynp = i32([s.npars for s in fit.systematics] or [0])
    ymean = np.asarray([m for s in fit.systematics for m in s.means] or [0.0], dtype=np.float64)
    ysig = np.asarray([m for s in fit.systematics for m in s.sigmas] or [0.0], dtype=np.float64)
    yfix = i32([s.fixed for s in fit.systematics] or [0])
    nexp, n_mc, sid = _nll_consts(fit)
    sidx = i32(sid)
    harr = (C.c_void_p * fit.nsignals)(*[h.h for h in hists])
    chain = np.zeros((max(1, nsteps), fit.nparameters + 1), dtype=np.float32)
    rows = lib().ref_mcmc_run(
        fit.nsources, _p(smean, c_fp), _p(ssig, c_fp), _p(sfix, c_ip),
        len(fit.systematics), _p(ynp, c_ip), _p(ymean, c_dp), _p(ysig, c_dp), _p(yfix, c_ip),
        fit.nobs, fit.nsignals, harr, _p(nexp, c_dp), _p(n_mc, c_up), _p(sidx, c_ip),
        _p(fit.data, c_fp), C.c_size_t(fit.data.size), C.c_uint(nsteps),
        C.c_float(burnin_fraction), int(debug_mode), C.c_uint(sync_interval), int(verbose),
        _p(chain, c_fp), C.c_long(chain.shape[0]))
    used = (int(lib().ref_tape_normals_used()), int(lib().ref_tape_uniforms_used()))
    lib().ref_set_tapes(None, 0, None, 0, 1)
    if own:
        for h in hists:
            h.close()
    return dict(chain=chain[:rows].copy(), tape_used=used)


def config_dump(json_path, verbose=False):
    """The reference's own FitConfig (src/config.cpp:19-297) run on `json_path` (signals in .npy
    tables with .fields sidecars), as a dict: every ordering / rounding decision it took."""
    import json
    L = lib()
    L.ref_config_dump.restype = C.c_long
    buf = C.create_string_buffer(1 << 20)
    n = L.ref_config_dump(str(json_path).encode(), int(verbose), buf, C.c_size_t(len(buf)))
    if n < 0:
        raise RefError("FitConfig threw")
    return json.loads(buf.value.decode())
