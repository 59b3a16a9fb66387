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
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(1, 0)
    chain.begin(fit.data, (2 * a.rounds + 2) * a.steps + 8, 0.1)
    chain.advance(a.steps)
    res = {"pdl": [], "no pdl": []}
    for _ in range(a.rounds):
        for name, off in (("pdl", False), ("no pdl", True)):
            pdfz.set_bin_tuning(no_pdl=off)
            ms0 = chain.loop_ms()
            chain.advance(a.steps)
            res[name].append(1e3 * (chain.loop_ms() - ms0) / a.steps)
    pdfz.set_bin_tuning()
    chain.end()
    for name, v in res.items():
        print("%s %-7s us per step: median %7.2f  min %7.2f  max %7.2f" % (fit.name, name, np.median(v), min(v), max(v)))


if __name__ == "__main__":
    main()
