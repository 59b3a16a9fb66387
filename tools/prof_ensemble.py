#!/usr/bin/env python
"""Ensemble throughput (config E shape: workload B, Poisson fake data, nsteps per experiment):
  python tools/prof_ensemble.py --experiments 148 --steps 20000 [--batch 148] [--scale 1.0]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="B")
    ap.add_argument("--experiments", type=int, default=148)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--batch", type=int, default=148)
    ap.add_argument("--scale", type=float, default=1.0)
    a = ap.parse_args()
    from sxmc_b200 import _lib, ensemble, sharding, workloads
    _lib.check(_lib.lib().sxmc_init(0))
    fit = workloads.WORKLOADS[a.workload](a.scale)
    ensemble.ensemble(fit, min(a.experiments, 4), nsteps=200, burnin_fraction=0.1, batch=a.batch)   # warm-up
    t0 = time.perf_counter()
    r = ensemble.ensemble(fit, a.experiments, "s0", 0.9, seed=1, nsteps=a.steps, burnin_fraction=0.1,
                          batch=a.batch)
    wall = time.perf_counter() - t0
    steps = a.experiments * a.steps
    print(json.dumps({"workload": a.workload, "mode": r["mode"], "experiments": a.experiments,
                      "steps_per_experiment": a.steps, "wall_s": wall, "loop_ms": r["loop_ms"],
                      "experiments_per_s": a.experiments / wall, "chain_steps_per_s": steps / wall,
                      "us_per_chain_step_device": 1e3 * r["loop_ms"] / steps,
                      "median_limit": sharding.median(r["limits"]), "launches": r["launches"]}))


if __name__ == "__main__":
    main()
