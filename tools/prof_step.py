#!/usr/bin/env python
"""Small driver for ncu: builds one workload and runs a few plain-stream MCMC steps.
  python tools/prof_step.py --workload C --steps 3 [--scale 1.0] [--no-flush]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--stress", action="store_true", help="workload D with 100 x 50 x 40 = 200 000 bins")
    a = ap.parse_args()
    from sxmc_b200 import _lib, mcmc, workloads
    _lib.check(_lib.lib().sxmc_init(0))
    fit = workloads.config_D(a.scale, stress=True) if a.stress else workloads.WORKLOADS[a.workload](a.scale)
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(1, 0)
    chain.begin(fit.data, a.steps, 0.1)
    s, b, n = chain.advance_timed(a.steps, 0 if a.no_flush else (256 << 20))
    chain.end()
    print("workload %s: %d steps, bin %.4f ms/step, nll %.4f ms/step" %
          (a.workload, a.steps, b / a.steps, n / a.steps))


if __name__ == "__main__":
    main()
