"""Host-side mirror of the reference's ``MCMC`` and ``LikelihoodSpace`` classes
(src/mcmc.h:36-138, src/likelihood.h:25-100) over the C ABI, plus the plain-data
``Source`` / ``Signal`` / ``Systematic`` / ``Observable`` structs MCMC's constructor takes
(src/source.h:23-57, src/signal.h:29-72, src/systematic.h:26-48, src/observable.h:24-41).

The whole chain runs device-resident inside libsxmc_b200.so."""
import ctypes as C
from dataclasses import dataclass, field
from typing import List

import numpy as np

from . import pdfz
from ._lib import ChainConfig, c_dp, c_fp, c_ip, c_sp, c_up, c_vp, lib, ptr
from .pdfz import _check


@dataclass
class Source:
    name: str = ""
    index: int = 0
    mean: float = 0.0
    sigma: float = 0.0
    fixed: bool = False


@dataclass
class Systematic:
    name: str = ""
    npars: int = 1
    means: List[float] = field(default_factory=lambda: [0.0])
    sigmas: List[float] = field(default_factory=lambda: [0.0])
    fixed: bool = False


@dataclass
class Observable:
    """src/observable.h:24-41 (the fields the hot path and the generator use)"""
    name: str = ""
    lower: float = 0.0
    upper: float = 0.0


@dataclass
class Signal:
    name: str = ""
    histogram: pdfz.EvalHist = None
    nexpected: float = 0.0
    n_mc: int = 0
    source: Source = None


class LikelihoodSpace:
    """Holds the chain ntuple: rows of nparameters+1 floats, last = NLL (src/likelihood.h:30).
    Post-processing (error estimators) is out of scope for the hot path."""

    def __init__(self, samples, names):
        self.samples = samples
        self.names = names

    def get_samples(self):
        return self.samples

    def get_best_fit(self):
        """Parameters of the minimum-NLL row (src/likelihood.cpp:46-72, without the errors)."""
        i = int(np.argmin(self.samples[:, -1]))
        return dict(zip(self.names[:-1], self.samples[i, :-1])), float(self.samples[i, -1])


class MCMC:
    """MCMC(sources, signals, systematics, observables) (src/mcmc.cpp:28-130)."""

    def __init__(self, sources, signals, systematics, observables):
        self._configure(sources, signals, systematics, observables)
        self._c = c_vp()
        _check(lib().sxmc_chain_create(C.byref(self._cfg), C.byref(self._c)))
        self.nrows = 0

    def _configure(self, sources, signals, systematics, observables):
        self.nsources, self.nsignals = len(sources), len(signals)
        self.nsystematics, self.nobservables = len(systematics), len(observables)
        means, sigmas, fixed = [], [], []
        for s in sources:
            # Source.mean / sigma are float members (src/source.h:55-56)
            means.append(float(np.float32(s.mean)))
            sigmas.append(float(np.float32(s.sigma)))
            fixed.append(bool(s.fixed))
        self.systematics_fixed = True
        self.parameter_names = [s.name for s in sources]
        for y in systematics:
            if not y.fixed:
                self.systematics_fixed = False
            for j in range(y.npars):
                means.append(float(y.means[j]))
                sigmas.append(float(y.sigmas[j]))
                fixed.append(bool(y.fixed))
                self.parameter_names.append("%s_%d" % (y.name, j))
        self.parameter_names.append("likelihood")
        self.nparameters = len(means)
        self._means = np.asarray(means, dtype=np.float64)
        self._sigmas = np.asarray(sigmas, dtype=np.float64)
        self._fixed = np.asarray(fixed, dtype=np.int32)
        self._nexp = np.asarray([s.nexpected for s in signals], dtype=np.float64)
        self._n_mc = np.asarray([s.n_mc for s in signals], dtype=np.uint32)
        self._sid = np.asarray([s.source.index for s in signals], dtype=np.int16)
        self._pdfs = [s.histogram for s in signals]
        self._harr = (c_vp * self.nsignals)(*[h._h for h in self._pdfs])
        self._cfg = ChainConfig(self.nsources, self.nsignals, self.nparameters, self.nobservables,
                                ptr(self._means, c_dp), ptr(self._sigmas, c_dp), ptr(self._fixed, c_ip),
                                int(self.systematics_fixed), self.nsystematics, self._harr,
                                ptr(self._nexp, c_dp), ptr(self._n_mc, c_up), ptr(self._sid, c_sp))

    def close(self):
        if self._c:
            lib().sxmc_chain_destroy(self._c)
            self._c = c_vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration that has no reference counterpart (RNG is gRandom there)
    def set_seed(self, seed, stream_id=0):
        _check(lib().sxmc_chain_set_seed(self._c, int(seed), int(stream_id)))

    def set_tapes(self, normals, uniforms):
        if normals is None:
            _check(lib().sxmc_chain_set_tapes(self._c, None, 0, None, 0))
            return
        n = np.ascontiguousarray(normals, dtype=np.float64)
        u = np.ascontiguousarray(uniforms, dtype=np.float64)
        _check(lib().sxmc_chain_set_tapes(self._c, ptr(n, c_dp), n.size, ptr(u, c_dp), u.size))

    def set_mode(self, mode):
        _check(lib().sxmc_chain_set_mode(self._c, int(mode)))

    def set_comm(self, comm):
        _check(lib().sxmc_chain_set_comm(self._c, comm))

    def enable_trace(self, on=True):
        _check(lib().sxmc_chain_enable_trace(self._c, int(on)))

    # -- MCMC::operator() (src/mcmc.cpp:143-387)
    def __call__(self, data, nsteps, burnin_fraction, debug_mode=False, sync_interval=10000):
        d = np.ascontiguousarray(data, dtype=np.float32).ravel()
        dd = d if d.size else np.zeros(1, np.float32)
        nrows = C.c_long(0)
        _check(lib().sxmc_chain_run(self._c, ptr(dd, c_fp), d.size, int(nsteps),
                                    float(burnin_fraction), int(debug_mode), int(sync_interval),
                                    C.byref(nrows)))
        self.nrows = nrows.value
        rows = np.zeros((max(1, self.nrows), self.nparameters + 1), dtype=np.float32)
        _check(lib().sxmc_chain_read(self._c, ptr(rows, c_fp), rows.shape[0]))
        return LikelihoodSpace(rows[:self.nrows], self.parameter_names)

    def run_resident(self, data, nsteps, burnin_fraction, debug_mode=False):
        """MCMC::operator() without the device->host copy of the rows: they stay resident for
        `sxmc_chain_contour` (ensemble loop); `rows()` fetches them on demand."""
        d = np.ascontiguousarray(data, dtype=np.float32).ravel()
        dd = d if d.size else np.zeros(1, np.float32)
        nrows = C.c_long(0)
        _check(lib().sxmc_chain_run(self._c, ptr(dd, c_fp), d.size, int(nsteps),
                                    float(burnin_fraction), int(debug_mode), 10000, C.byref(nrows)))
        self.nrows = nrows.value
        self._nevents = d.size // (self.nobservables + 1)
        return self.nrows

    def rows(self):
        rows = np.zeros((max(1, self.nrows), self.nparameters + 1), dtype=np.float32)
        _check(lib().sxmc_chain_read(self._c, ptr(rows, c_fp), rows.shape[0]))
        return rows[:self.nrows]

    # -- the same call in parts (sxmc_chain_begin / advance / end), for resident chains
    def begin(self, data, nsteps, burnin_fraction, debug_mode=False):
        d = np.ascontiguousarray(data, dtype=np.float32).ravel()
        dd = d if d.size else np.zeros(1, np.float32)
        _check(lib().sxmc_chain_begin(self._c, ptr(dd, c_fp), d.size, int(nsteps),
                                      float(burnin_fraction), int(debug_mode)))
        self._nevents = d.size // (self.nobservables + 1)

    def advance(self, nsteps):
        _check(lib().sxmc_chain_advance(self._c, int(nsteps)))

    def advance_timed(self, nsteps, flush_bytes=0):
        """Returns (step_ms, bin_ms, nll_ms) summed over nsteps, from per-kernel CUDA events."""
        a, b, c = C.c_float(0), C.c_float(0), C.c_float(0)
        _check(lib().sxmc_chain_advance_timed(self._c, int(nsteps), int(flush_bytes),
                                              C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def end(self):
        nrows = C.c_long(0)
        _check(lib().sxmc_chain_end(self._c, C.byref(nrows)))
        self.nrows = nrows.value
        rows = np.zeros((max(1, self.nrows), self.nparameters + 1), dtype=np.float32)
        _check(lib().sxmc_chain_read(self._c, ptr(rows, c_fp), rows.shape[0]))
        return LikelihoodSpace(rows[:self.nrows], self.parameter_names)

    def accepted(self):
        v = C.c_long(0)
        _check(lib().sxmc_chain_accepted(self._c, C.byref(v)))
        return v.value

    def loop_ms(self):
        v = C.c_float(0)
        _check(lib().sxmc_chain_loop_ms(self._c, C.byref(v)))
        return v.value

    def launches(self):
        v = C.c_long(0)
        _check(lib().sxmc_chain_launches(self._c, C.byref(v)))
        return v.value

    def last_comm_ms(self):
        v = C.c_float(0)
        _check(lib().sxmc_chain_last_comm_ms(self._c, C.byref(v)))
        return v.value

    def bin_kernel_kind(self):
        """0 interpreter, 1 run-time dispatch, 2 compile-time program of the build, 3 NVRTC-compiled program"""
        v = C.c_int(0)
        _check(lib().sxmc_chain_bin_kernel_kind(self._c, C.byref(v)))
        return v.value

    def bin_plan(self):
        """Pipeline shape of the binning launch: warps per CTA, rows per chunk, stages per warp, histogram in
        shared memory (sxmc_chain_bin_plan)."""
        v = [C.c_int(0) for _ in range(4)]
        _check(lib().sxmc_chain_bin_plan(self._c, *[C.byref(x) for x in v]))
        return {"warps": v[0].value, "rows_per_chunk": v[1].value, "stages": v[2].value, "smem_hist": bool(v[3].value)}

    def last_path(self):
        """1 graph / stream launches, 2 persistent kernel, 4 monotone edge-search kernel"""
        v = C.c_int(0)
        _check(lib().sxmc_chain_last_path(self._c, C.byref(v)))
        return v.value

    def read_trace(self, nsteps):
        nll = np.zeros(max(1, nsteps), dtype=np.float64)
        prop = np.zeros((max(1, nsteps), self.nparameters), dtype=np.float64)
        _check(lib().sxmc_chain_read_trace(self._c, ptr(nll, c_dp), ptr(prop, c_dp), int(nsteps)))
        return nll[:nsteps], prop[:nsteps]

    # -- MCMC::nll (src/mcmc.cpp:390-415) at a host parameter vector, PDFs re-evaluated
    def set_data(self, data):
        d = np.ascontiguousarray(data, dtype=np.float32).ravel()
        dd = d if d.size else np.zeros(1, np.float32)
        _check(lib().sxmc_chain_set_data(self._c, ptr(dd, c_fp), d.size))
        self._nevents = d.size // (self.nobservables + 1)

    def nll(self, pars, rebin=True):
        p = np.ascontiguousarray(pars, dtype=np.float64)
        v, ev = C.c_double(0), C.c_double(0)
        _check(lib().sxmc_chain_nll_at(self._c, ptr(p, c_dp), int(rebin), C.byref(v), C.byref(ev)))
        return v.value, ev.value

    def nll_lut(self, lut, pars, norms):
        """MCMC::nll(lut, nevents, v, nll, nexpected, n_mc, source_id, norms, ...) of
        src/mcmc.cpp:390-415 on caller-held DEVICE buffers (``pdfz.Array``): the materialised PDF
        lookup table [nsignals][nevents], the parameter vector and the normalisations."""
        arrs = [pdfz.Array.from_host(v, t) for v, t in (
            (self._means, np.float64), (self._sigmas, np.float64), (self._nexp, np.float64),
            (self._n_mc, np.uint32), (self._sid, np.int16))]
        v, ev = C.c_double(0), C.c_double(0)
        nevents = lut.size() // max(1, self.nsignals)
        _check(lib().sxmc_nll_eval(lut.readOnlyDevicePtr(), nevents, self.nsignals, pars.readOnlyDevicePtr(),
                                   self.nparameters, self.nsources, *[a.readOnlyDevicePtr() for a in arrs],
                                   norms.readOnlyDevicePtr(), C.byref(v), C.byref(ev)))
        for a in arrs:
            a.free()
        return v.value, ev.value

    def norms(self):
        out = np.zeros(self.nsignals, dtype=np.uint32)
        _check(lib().sxmc_chain_get_norms(self._c, ptr(out, c_up)))
        return out

    def lut(self):
        out = np.zeros((self.nsignals, max(1, self._nevents)), dtype=np.float32)
        _check(lib().sxmc_chain_get_lut(self._c, ptr(out, c_fp)))
        return out[:, :self._nevents]

    def bins(self):
        tn = self._pdfs[0].TotalBins()
        out = np.zeros((self.nsignals, tn), dtype=np.uint32)
        _check(lib().sxmc_chain_get_bins(self._c, ptr(out, c_up)))
        return out


class MCMCBatch:
    """Up to 148 MCMC objects over the same signals advanced by one launch (`sxmc_batch_*`):
    the experiments of an ensemble.  Constructor arguments as for ``MCMC``."""

    def __init__(self, sources, signals, systematics, observables, max_chains=148):
        proto = MCMC.__new__(MCMC)
        proto._c = c_vp()
        MCMC._configure(proto, sources, signals, systematics, observables)
        self._proto = proto
        self.nparameters, self.nobservables = proto.nparameters, proto.nobservables
        self.parameter_names = proto.parameter_names
        self.max_chains = int(max_chains)
        self._b = c_vp()
        _check(lib().sxmc_batch_create(C.byref(proto._cfg), self.max_chains, C.byref(self._b)))

    def close(self):
        if self._b:
            lib().sxmc_batch_destroy(self._b)
            self._b = c_vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_seed(self, slot, seed, stream_id=0):
        _check(lib().sxmc_batch_set_seed(self._b, int(slot), int(seed), int(stream_id)))

    def set_tapes(self, slot, normals, uniforms):
        n = np.ascontiguousarray(normals, dtype=np.float64)
        u = np.ascontiguousarray(uniforms, dtype=np.float64)
        _check(lib().sxmc_batch_set_tapes(self._b, int(slot), ptr(n, c_dp), n.size, ptr(u, c_dp), u.size))

    def set_data(self, slot, data):
        d = np.ascontiguousarray(data, dtype=np.float32).ravel()
        dd = d if d.size else np.zeros(1, np.float32)
        _check(lib().sxmc_batch_set_data(self._b, int(slot), ptr(dd, c_fp), d.size))

    def generate_data(self, slot, generator, poisson=True, seed=0, experiment=0):
        observed = np.zeros(generator.nsignals, dtype=np.uint32)
        _check(lib().sxmc_batch_generate_data(self._b, int(slot), generator._g, int(bool(poisson)), int(seed),
                                              int(experiment), ptr(observed, c_up)))
        return observed

    def enable_trace(self, on=True):
        _check(lib().sxmc_batch_enable_trace(self._b, int(on)))

    def run(self, nchains, nsteps, burnin_fraction, debug_mode=False):
        _check(lib().sxmc_batch_run(self._b, int(nchains), int(nsteps), float(burnin_fraction), int(debug_mode)))

    def nrows(self, slot):
        a, b = C.c_long(0), C.c_long(0)
        _check(lib().sxmc_batch_nrows(self._b, int(slot), C.byref(a), C.byref(b)))
        return a.value, b.value

    def rows(self, slot):
        n, _ = self.nrows(slot)
        rows = np.zeros((max(1, n), self.nparameters + 1), dtype=np.float32)
        _check(lib().sxmc_batch_read(self._b, int(slot), ptr(rows, c_fp), rows.shape[0]))
        return rows[:n]

    def likelihood_space(self, slot):
        return LikelihoodSpace(self.rows(slot), self.parameter_names)

    def read_trace(self, slot, nsteps):
        nll = np.zeros(max(1, nsteps), dtype=np.float64)
        prop = np.zeros((max(1, nsteps), self.nparameters), dtype=np.float64)
        _check(lib().sxmc_batch_read_trace(self._b, int(slot), ptr(nll, c_dp), ptr(prop, c_dp), int(nsteps)))
        return nll[:nsteps], prop[:nsteps]

    def contour(self, slot, delta):
        np_ = self.nparameters
        lo, up, pt = (np.zeros(np_, np.float32) for _ in range(3))
        lmin, nc = C.c_float(0), C.c_long(0)
        _check(lib().sxmc_batch_contour(self._b, int(slot), float(delta), ptr(lo, c_fp), ptr(up, c_fp),
                                        ptr(pt, c_fp), C.byref(lmin), C.byref(nc)))
        return lo, up, pt, float(lmin.value), int(nc.value)

    def loop_ms(self):
        v = C.c_float(0)
        _check(lib().sxmc_batch_loop_ms(self._b, C.byref(v)))
        return v.value

    def last_path(self):
        """(1 = monotone edge-search kernel | 2 = sample-by-sample kernel, max edges per CTA)"""
        a, b = C.c_int(0), C.c_int(0)
        _check(lib().sxmc_batch_last_path(self._b, C.byref(a), C.byref(b)))
        return a.value, b.value


def build_fit(fit, sort=True, batch=0, hists=None):
    """FitSpec (sxmc_b200.workloads) -> (hists, MCMC), the way src/signal.cpp:112-169 and
    src/sxmc.cpp:96 wire the reference objects.  ``hists``: EvalHist objects already built for
    this fit (a second MCMC over the same PDFs, as `plot_fit` reuses them in the reference)."""
    reuse = hists is not None
    hists = list(hists) if reuse else []
    for sig in ([] if reuse else fit.signals):
        h = pdfz.EvalHist(sig.samples, sig.samples.shape[1], fit.nobs, fit.lower, fit.upper,
                          fit.nbins, sig.dataset)
        if not sort:
            h.SetLayout(False)
        for i in sig.systematics:
            y = fit.systematics[i]
            if y.type == pdfz.SHIFT:
                h.AddSystematic(pdfz.ShiftSystematic(y.obs, y.pidx))
            elif y.type == pdfz.SCALE:
                h.AddSystematic(pdfz.ScaleSystematic(y.obs, y.pidx))
            elif y.type == pdfz.CTSCALE:
                h.AddSystematic(pdfz.CosThetaScaleSystematic(y.obs, y.pidx))
            else:
                h.AddSystematic(pdfz.ResolutionScaleSystematic(y.obs, y.extra, y.pidx))
        hists.append(h)
    nm = getattr(fit, "names", None) or {}      # config.FitConfig.to_fitspec carries the configured names
    name = lambda kind, fmt, i: nm[kind][i] if kind in nm else fmt % i
    sources = [Source(name("sources", "s%d", i), i, fit.source_means[i], fit.source_sigmas[i], fit.source_fixed[i])
               for i in range(fit.nsources)]
    systs = [Systematic(name("systematics", "y%d", i), y.npars, list(y.means), list(y.sigmas), y.fixed)
             for i, y in enumerate(fit.systematics)]
    signals = [Signal(name("signals", "sig%d", j), hists[j], sig.nexpected, sig.n_mc, sources[sig.source])
               for j, sig in enumerate(fit.signals)]
    observables = [Observable(name("observables", "o%d", i), float(fit.lower[i]), float(fit.upper[i]))
                   for i in range(fit.nobs)]
    if batch:
        return hists, MCMCBatch(sources, signals, systs, observables, batch)
    mc = MCMC(sources, signals, systs, observables)
    return hists, mc
