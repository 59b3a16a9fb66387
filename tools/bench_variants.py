#!/usr/bin/env python
"""The binning kernel outside the benchmark's own systematic programs: event-timed
launch of ONE chain per variant, steps back to back, on the shapes of workloads C and D:
  table     the compile-time program of the workload (SXMC_FAST_TABLE)
  dynamic   the same fit through the run-time dispatching instantiation (what a program outside the table gets)
  generic   the per-sample interpreter (what multi-coefficient fits get)
  other     D's shape with shift(radius) + ctscale(costheta): a program that is not in the table
  stress    D with 100 x 50 x 40 = 200 000 bins (histograms in global memory, L2 atomics)
One JSON object per line.
  python tools/bench_variants.py > profiles/rNN_bin_variants.jsonl"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(fit, label, steps=20, rounds=5, **tuning):
    from sxmc_b200 import mcmc, pdfz
    import bench
    pdfz.set_bin_tuning(**tuning)
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(1, 0)
    chain.begin(fit.data, (rounds + 1) * steps + 8, 0.1)
    chain.advance_timed(4, 0)
    kind = {0: "interpreter", 1: "run-time dispatch", 2: "compile-time program (build)", 3: "compile-time program (NVRTC)"}[chain.bin_kernel_kind()]
    v = []
    for _ in range(rounds):
        s, b, n = chain.advance_timed(steps, 0)
        v.append(1e3 * b / steps)
    chain.end()
    chain.close()
    for h in hists:
        h.close()
    pdfz.set_bin_tuning()
    nbytes = 4 * fit.needed_columns() * fit.total_samples()
    peak, _ = bench.measured_peak()
    med = float(np.median(v))
    out = {"case": label, "kernel": kind, "bytes": nbytes, "bin_us": round(med, 2), "gbs": round(nbytes / med / 1e3, 1),
           "frac": round(nbytes / med / 1e3 / peak, 4), "peak_gbs": peak}
    print(json.dumps(out), flush=True)


def main():
    from sxmc_b200 import _lib, workloads as W
    _lib.check(_lib.lib().sxmc_init(0))
    for name in ("C", "D"):
        fit = W.WORKLOADS[name](1.0)
        run(fit, name + " table")
        run(fit, name + " dynamic", force_dynamic=True)
        run(fit, name + " generic", force_generic=True, rounds=3, steps=6)
        if name == "D":
            other = W.config_D(1.0)
            other.systematics = [W.SystSpec(type=W.SHIFT, obs=1, means=[0.0], sigmas=[0.01]),
                                 W.SystSpec(type=W.CTSCALE, obs=2, means=[0.0], sigmas=[0.01])]
            other.finalize()
            for sig in other.signals:
                sig.systematics = [0, 1]
            run(other, "D shape, shift(radius)+ctscale(costheta) [not in the table]")
            run(other, "D shape, shift(radius)+ctscale(costheta), JIT off", no_jit=True)
        del fit
    run(W.config_D(1.0, stress=True), "D stress 200000 bins (global histograms)")


if __name__ == "__main__":
    main()
