#!/usr/bin/env python
"""A/B inside one process: chain steps issued back to back with and without programmatic
dependent launch between the binning and the lookup kernels (device time of K-step blocks).
  python tools/ab_pdl.py --workload D [--steps 20] [--rounds 7]"""
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
    a = ap.parse_args()
    from sxmc_b200 import _lib, mcmc, pdfz, workloads
    _lib.check(_lib.lib().sxmc_init(0))
    fit = workloads.WORKLOADS[a.workload](1.0)
    # two chains over the same PDFs: a handle keeps the knobs it was created with
    hists, chain_pdl = mcmc.build_fit(fit)
    pdfz.set_bin_tuning(no_pdl=True)
    _, chain_off = mcmc.build_fit(fit, hists=hists)
    pdfz.set_bin_tuning()
    chains = {"pdl": chain_pdl, "no pdl": chain_off}
    for c in chains.values():
        c.set_seed(1, 0)
        c.begin(fit.data, (a.rounds + 2) * a.steps + 8, 0.1)
        c.advance(a.steps)
    res = {"pdl": [], "no pdl": []}
    for _ in range(a.rounds):
        for name, c in chains.items():
            ms0 = c.loop_ms()
            c.advance(a.steps)
            res[name].append(1e3 * (c.loop_ms() - ms0) / a.steps)
    for c in chains.values():
        c.end()
    for name, v in res.items():
        print("%s %-7s us per step: median %7.2f  min %7.2f  max %7.2f" % (fit.name, name, np.median(v), min(v), max(v)))


if __name__ == "__main__":
    main()
