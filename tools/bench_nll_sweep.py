#!/usr/bin/env python
"""Roofline of the NLL pass (PDF lookup + event likelihood: eval_pdf + nll_event_chunks,
reference src/pdfz.cpp:411-436 and src/nll_kernels.cpp:89-116, launched at src/mcmc.cpp:314-323)
as a function of the number of data events.

A fit with FIXED systematics takes NLL-only steps (src/mcmc.cpp:264: the PDFs are not
re-evaluated), so a timed step is exactly one launch of lookup_nll_kernel.  Algorithmic bytes
per SURVEY 8(d): 4 B per (event, signal) -- the float lut element the reference's kernel reads;
this kernel reads 4 B of read_bins per (event, signal) instead and gathers the bin count from L2.
Two forms:
  per_event  every event looked up on its own (the only form a multi-dataset fit can use)
  grouped    events counted per bin once per fit, the kernel loops over bins (all signals read
             one dataset); its traffic does not depend on nevents at all, so its "GB/s" is an
             EQUIVALENT rate, not bytes moved

  python tools/bench_nll_sweep.py [--max-events 1e8] > profiles/rNN_nll_sweep.json"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def make_fit(nsignals, nevents, dims=1, seed=5):
    """Small Monte-Carlo sets (the NLL pass does not touch them), `nevents` data events."""
    from sxmc_b200 import workloads as W
    rng = np.random.default_rng(seed)
    lower, upper = [5.0, 0.0, -1.0][:dims], [15.0, 10.0, 1.0][:dims]
    nbins = {1: [1000], 2: [100, 50], 3: [50, 25, 20]}[dims]
    mus = np.linspace(7.0, 13.0, nsignals)
    signals = []
    for j in range(nsignals):
        rows = W._signal_rows(rng, 20000, mus[j], dims >= 2, dims >= 3, False)
        signals.append(W.SignalSpec(samples=rows, nexpected=nevents / nsignals, source=j))
    ev = W._signal_rows(rng, nevents, 10.0, dims >= 2, dims >= 3, False)[:, :dims]
    for d in range(dims):
        ev[:, d] = np.clip(ev[:, d], lower[d], np.nextafter(np.float32(upper[d]), np.float32(-np.inf)))
    data = np.concatenate([ev, np.zeros((nevents, 1), np.float32)], axis=1)
    systs = [W.SystSpec(type=W.SHIFT, obs=0, means=[0.0], sigmas=[0.01], fixed=True)]
    return W.FitSpec(name="nll_sweep", lower=lower, upper=upper, nbins=nbins,
                     source_means=[1.0] * nsignals, source_sigmas=[0.0] * nsignals,
                     source_fixed=[False] * nsignals, systematics=systs, signals=signals, data=data)


def measure(nsignals, nevents, grouped, steps, dims=1):
    from sxmc_b200 import mcmc, pdfz
    fit = make_fit(nsignals, nevents, dims)
    pdfz.set_bin_tuning(no_grouped_lookup=not grouped)
    try:
        hists, chain = mcmc.build_fit(fit)
        chain.set_seed(3, 0)
        chain.begin(fit.data, 2 * steps + 8, 0.1)
        flush = (256 << 20) if 4 * nsignals * nevents < (512 << 20) else 0
        chain.advance_timed(4, flush)
        _, _, nll_ms = chain.advance_timed(steps, flush)
        chain.end()
        chain.close()
        for h in hists:
            h.close()
    finally:
        pdfz.set_bin_tuning()
    return nll_ms / steps


def sweep(peak, quick=False, max_events=1e8):
    events = [10**4, 10**5, 10**6, 10**7] + ([] if quick else [10**8])
    events = [n for n in events if n <= max_events]
    out = {"what": "lookup_nll_kernel on NLL-only steps (fixed systematics): ms per launch (CUDA events), "
                   "algorithmic bytes = 4 B x nevents x nsignals (SURVEY 8(d)), frac = GB/s / peak; "
                   "'grouped' loops over bins, not events: an equivalent rate, not bytes moved; L2 flushed "
                   "before each launch while the lookup table (4 B x nevents x nsignals) is under 512 MB",
           "peak_gbs": peak, "rows": []}
    for ns in (3, 10, 30):
        for ne in events:
            if quick and ns == 30 and ne > 10**6:
                continue
            row = {"nsignals": ns, "nevents": ne, "algorithmic_bytes": 4 * ns * ne}
            for form in ("per_event", "grouped"):
                steps = 20 if ne <= 10**6 else (6 if ne <= 10**7 else 3)
                try:
                    ms = measure(ns, ne, form == "grouped", steps)
                    gbs = 4.0 * ns * ne / (ms * 1e-3) / 1e9
                    row[form] = {"ms": round(ms, 5), "gbs": round(gbs, 1), "frac": round(gbs / peak, 4)}
                except Exception as e:
                    row[form] = {"error": repr(e)}
            out["rows"].append(row)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-events", type=float, default=1e8)
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    import bench
    from sxmc_b200 import _lib
    _lib.check(_lib.lib().sxmc_init(0))
    peak, _ = bench.measured_peak()
    print(json.dumps(sweep(peak, quick=a.quick, max_events=a.max_events)), flush=True)


if __name__ == "__main__":
    main()
