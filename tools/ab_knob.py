#!/usr/bin/env python
"""A/B inside one process: chain steps issued back to back with and without one kernel-selection
knob of pdfz.set_bin_tuning (device time of K-step blocks, the two chains alternating):
  no_pdl     steps without programmatic dependent launch between the binning and lookup kernels
  cta_split  global-histogram binning by per-CTA segments instead of per-warp ranges
  python tools/ab_knob.py --knob cta_split --workload D [--shard 8] [--stress] [--steps 20] [--rounds 7]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="D")
    ap.add_argument("--steps", type=int, default=20)      # < 32: plain stream launches, no captured graph
    ap.add_argument("--rounds", type=int, default=7)
    ap.add_argument("--knob", default="no_pdl")
    ap.add_argument("--values", default="1")              # comma-separated values of the knob, each its own chain
    ap.add_argument("--shard", type=int, default=1)       # rank 0's 1/G slice of every signal
    ap.add_argument("--stress", action="store_true")
    a = ap.parse_args()
    from sxmc_b200 import _lib, mcmc, pdfz, sharding, workloads
    _lib.check(_lib.lib().sxmc_init(0))
    fit = workloads.WORKLOADS[a.workload](1.0, stress=True) if a.stress else workloads.WORKLOADS[a.workload](1.0)
    if a.shard > 1:
        fit = sharding.shard_fit_samples(fit, 0, a.shard)
    # two chains over the same PDFs: a handle keeps the knobs it was created with
    hists, chain_pdl = mcmc.build_fit(fit)
    chains = {"default": chain_pdl}
    for v in a.values.split(","):
        pdfz.set_bin_tuning(**{a.knob: int(v)})
        _, chains["%s=%s" % (a.knob, v)] = mcmc.build_fit(fit, hists=hists)
    pdfz.set_bin_tuning()
    for c in chains.values():
        c.set_seed(1, 0)
        c.begin(fit.data, (a.rounds + 2) * a.steps + 8, 0.1)
        c.advance(a.steps)
    res = {k: [] for k in chains}
    for _ in range(a.rounds):
        for name, c in chains.items():
            ms0 = c.loop_ms()
            c.advance(a.steps)
            res[name].append(1e3 * (c.loop_ms() - ms0) / a.steps)
    for c in chains.values():
        c.end()
    for name, v in res.items():
        print("%s %-12s us per step: median %7.2f  min %7.2f  max %7.2f" % (fit.name, name, np.median(v), min(v), max(v)))


if __name__ == "__main__":
    main()
