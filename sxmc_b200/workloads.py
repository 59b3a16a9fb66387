"""Fit descriptions and the seeded synthetic workloads of SURVEY.md section 8.

A ``FitSpec`` is the plain-data form of what the reference builds from its JSON config
(`src/config.cpp:19-297`): observables, sources, systematics (with their GLOBAL coefficient
indices ``pidx``, `src/config.cpp:114-121`), signals (sample rows = observables, truth
fields, DATASET; `src/config.cpp:162-194`) and data events (rows = observables + dataset
id, `src/mcmc.cpp:231`).  It is consumed by the CUDA path (``sxmc_b200.api``), by the
test oracle and by ``bench.py``; it contains no compute.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

# numeric values of pdfz::Systematic::Type (src/pdfz.h:111-116)
SHIFT, SCALE, RESOLUTION_SCALE, CTSCALE = 0, 1, 2, 3


@dataclass
class SystSpec:
    type: int
    obs: int                      # field index of the observable it moves
    extra: int = 0                # field index of the truth value (RESOLUTION_SCALE)
    means: List[float] = field(default_factory=lambda: [0.0])
    sigmas: List[float] = field(default_factory=lambda: [0.0])
    fixed: bool = False
    pidx: List[int] = field(default_factory=list)   # filled by FitSpec.finalize()

    @property
    def npars(self) -> int:
        return len(self.means)


@dataclass
class SignalSpec:
    samples: np.ndarray           # float32 [nsamples, nfields], C-contiguous
    nexpected: float
    source: int
    dataset: int = 0
    systematics: Optional[List[int]] = None   # indices into FitSpec.systematics (None = all)
    n_mc: Optional[int] = None                # defaults to nsamples (src/signal.cpp:31-47)

    def __post_init__(self):
        self.samples = np.ascontiguousarray(self.samples, dtype=np.float32)
        if self.n_mc is None:
            self.n_mc = int(self.samples.shape[0])


@dataclass
class FitSpec:
    name: str
    lower: List[float]
    upper: List[float]
    nbins: List[int]
    source_means: List[float]
    source_sigmas: List[float]
    source_fixed: List[bool]
    systematics: List[SystSpec]
    signals: List[SignalSpec]
    data: np.ndarray              # float32 [nevents, nobs+1]
    nsteps: int = 1000
    burnin_fraction: float = 0.1

    def __post_init__(self):
        self.data = np.ascontiguousarray(self.data, dtype=np.float32)
        self.finalize()

    def finalize(self):
        k = 0
        for s in self.systematics:
            s.pidx = list(range(k, k + s.npars))
            k += s.npars
        for sig in self.signals:
            if sig.systematics is None:
                sig.systematics = list(range(len(self.systematics)))

    # -- derived quantities, in the reference's parameter order (src/mcmc.cpp:50-83) --
    @property
    def nobs(self) -> int:
        return len(self.nbins)

    @property
    def nsources(self) -> int:
        return len(self.source_means)

    @property
    def nsignals(self) -> int:
        return len(self.signals)

    @property
    def nparameters(self) -> int:
        return self.nsources + sum(s.npars for s in self.systematics)

    @property
    def nevents(self) -> int:
        return int(self.data.shape[0])

    @property
    def total_nbins(self) -> int:
        return int(np.prod(self.nbins))

    def parameter_means(self) -> np.ndarray:
        # Source.mean / Source.sigma are float (src/source.h:55-56), widened to double
        v = [float(np.float32(m)) for m in self.source_means]
        for s in self.systematics:
            v += list(s.means)
        return np.asarray(v, dtype=np.float64)

    def parameter_sigmas(self) -> np.ndarray:
        v = [float(np.float32(m)) for m in self.source_sigmas]
        for s in self.systematics:
            v += list(s.sigmas)
        return np.asarray(v, dtype=np.float64)

    def parameter_fixed(self) -> np.ndarray:
        v = [bool(f) for f in self.source_fixed]
        for s in self.systematics:
            v += [bool(s.fixed)] * s.npars
        return np.asarray(v, dtype=np.int32)

    @property
    def systematics_fixed(self) -> bool:
        return all(s.fixed for s in self.systematics)

    def total_samples(self) -> int:
        return int(sum(s.samples.shape[0] for s in self.signals))

    def needed_columns(self) -> int:
        """Columns the binning math reads: observables + truth fields."""
        cols = set(range(self.nobs))
        for s in self.systematics:
            cols.add(s.obs)
            if s.type == RESOLUTION_SCALE:
                cols.add(s.extra)
        return len(cols)


# ---------------------------------------------------------------------------------------
# Synthetic workloads ("Value distributions")
# ---------------------------------------------------------------------------------------

def _signal_rows(rng, n, mu, want_radius, want_cos, want_truth):
    """Rows [energy, (radius), (costheta), (mc_energy), DATASET]."""
    energy = rng.normal(mu, 1.5, n)
    cols = [energy]
    if want_radius:
        cols.append(10.0 * rng.random(n) ** (1.0 / 3.0))
    if want_cos:
        cols.append(rng.uniform(-1.0, 1.0, n))
    if want_truth:
        cols.append(energy - rng.normal(0.0, 0.3, n))
    cols.append(np.zeros(n))
    return np.stack(cols, axis=1).astype(np.float32)


def _make(name, cfg_index, nsignals, nsamples, nevents, dims, systs, nsteps, scale=1.0,
          nbins_override=None):
    """dims: 1, 2 or 3 observables (energy, radius, costheta)."""
    lower = [5.0, 0.0, -1.0][:dims]
    upper = [15.0, 10.0, 1.0][:dims]
    nbins_tbl = {1: [1000], 2: [100, 50], 3: [50, 25, 20]}
    nbins = list(nbins_override or nbins_tbl[dims])
    want_truth = any(t == RESOLUTION_SCALE for t, _ in systs)
    truth_col = dims  # first column after the observables
    mus = np.linspace(7.0, 13.0, nsignals) if nsignals > 1 else np.array([10.0])
    if isinstance(nsamples, int):
        nsamples = [nsamples] * nsignals
    nsamples = [max(1, int(n * scale)) for n in nsamples]

    signals, data_rows = [], []
    per_sig_events = max(1, nevents // nsignals)
    for j in range(nsignals):
        rng = np.random.default_rng(1000 * cfg_index + j)
        rows = _signal_rows(rng, nsamples[j], mus[j], dims >= 2, dims >= 3, want_truth)
        signals.append(SignalSpec(samples=rows, nexpected=float(per_sig_events), source=j))
        # data events: same distribution, observables only, kept when inside the domain
        drng = np.random.default_rng(1000 * cfg_index + 500 + j)
        ev = _signal_rows(drng, per_sig_events * 2, mus[j], dims >= 2, dims >= 3, False)
        ok = np.ones(len(ev), dtype=bool)
        for d in range(dims):
            ok &= (ev[:, d] >= lower[d]) & (ev[:, d] < upper[d])
        ev = ev[ok][:per_sig_events, :dims]
        data_rows.append(np.concatenate([ev, np.zeros((len(ev), 1), np.float32)], axis=1))
    data = np.concatenate(data_rows, axis=0).astype(np.float32)

    sig_tbl = {SHIFT: 0.01, SCALE: 1e-2, RESOLUTION_SCALE: 1e-3, CTSCALE: 1e-2}
    syst_specs = [SystSpec(type=t, obs=o, extra=(truth_col if t == RESOLUTION_SCALE else 0),
                           means=[0.0], sigmas=[sig_tbl[t]]) for t, o in systs]
    return FitSpec(name=name, lower=lower, upper=upper, nbins=nbins,
                   source_means=[1.0] * nsignals, source_sigmas=[0.0] * nsignals,
                   source_fixed=[False] * nsignals, systematics=syst_specs,
                   signals=signals, data=data, nsteps=nsteps)


def config_A(scale=1.0):
    """example.json shape: 1-D energy, 10 bins, 2 signals (500k, 1M), scale + resolution."""
    return _make("A", 0, 2, [500_000, 1_000_000], 2000, 1,
                 [(SCALE, 0), (RESOLUTION_SCALE, 0)], 100_000, scale, nbins_override=[10])


def config_B(scale=1.0):
    """1-D energy (1000 bins), 3 signals x 1M, one floating shift, 10k events."""
    return _make("B", 1, 3, 1_000_000, 10_000, 1, [(SHIFT, 0)], 100_000, scale)


def config_C(scale=1.0):
    """2-D energy x radius (100x50), 10 signals x 10M, scale + resolution_scale, 100k events."""
    return _make("C", 2, 10, 10_000_000, 100_000, 2,
                 [(SCALE, 0), (RESOLUTION_SCALE, 0)], 20_000, scale)


def config_D(scale=1.0, stress=False):
    """3-D energy x radius x costheta (50x25x20), 30 signals, 100M samples in total."""
    return _make("D", 3, 30, 3_333_333, 100_000, 3,
                 [(SCALE, 0), (RESOLUTION_SCALE, 0), (SHIFT, 1)], 2_000, scale,
                 nbins_override=[100, 50, 40] if stress else None)


def bench_pdfz_shape(scale=1.0):
    """The reference's own benchmark: 10M N(0,1) samples, 1-D, 1000 bins on [-3,3),
    100k clamped-Gaussian eval points, one shift systematic at 0
    (bench/bench_sxmc.cpp:34-102)."""
    n = max(1, int(10_000_000 * scale))
    rng = np.random.default_rng(424242)
    rows = np.stack([rng.normal(0, 1, n), np.zeros(n)], axis=1).astype(np.float32)
    ev = np.clip(rng.normal(0, 1, 100_000), -2.999, 2.999)
    data = np.stack([ev, np.zeros_like(ev)], axis=1).astype(np.float32)
    return FitSpec(name="bench_pdfz", lower=[-3.0], upper=[3.0], nbins=[1000],
                   source_means=[1.0], source_sigmas=[0.0], source_fixed=[False],
                   systematics=[SystSpec(type=SHIFT, obs=0, means=[0.0], sigmas=[0.01])],
                   signals=[SignalSpec(samples=rows, nexpected=100_000.0, source=0)],
                   data=data, nsteps=100)


PDFZ_GROUP_NSAMPLES = [1e3, 2e5, 1e4, 1e3, 1e3, 3e6, 5e5, 1e6, 8e4, 2e4] + [1e3] * 19


def pdfz_group_shape(scale=1.0):
    """The reference's second benchmark (bench/bench_sxmc.cpp:105-225): 29 PDFs of very
    different sizes (10^3 ... 3x10^6 N(0,1) samples), 1-D, 1000 bins on [-3,3), one shift
    systematic at 0, 100k clamped-Gaussian eval points, all evaluated together."""
    rng = np.random.default_rng(434343)
    signals = []
    for j, n in enumerate(PDFZ_GROUP_NSAMPLES):
        n = max(1, int(n * scale))
        rows = np.stack([rng.normal(0, 1, n), np.zeros(n)], axis=1).astype(np.float32)
        signals.append(SignalSpec(samples=rows, nexpected=1000.0, source=j))
    ev = np.clip(rng.normal(0, 1, max(1, int(100_000 * min(1.0, scale * 10)))), -2.999, 2.999)
    data = np.stack([ev, np.zeros_like(ev)], axis=1).astype(np.float32)
    ns = len(signals)
    return FitSpec(name="pdfz_group", lower=[-3.0], upper=[3.0], nbins=[1000],
                   source_means=[1.0] * ns, source_sigmas=[0.0] * ns, source_fixed=[False] * ns,
                   systematics=[SystSpec(type=SHIFT, obs=0, means=[0.0], sigmas=[0.01])],
                   signals=signals, data=data, nsteps=100)


WORKLOADS = {"A": config_A, "B": config_B, "C": config_C, "D": config_D,
             "bench_pdfz": bench_pdfz_shape, "pdfz_group": pdfz_group_shape}
