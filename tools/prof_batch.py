#!/usr/bin/env python
"""Phase breakdown of the chain batch kernel (every CTA's %globaltimer stamps of super-step 20).
  python tools/prof_batch.py [--workload B] [--chains 148] [--steps 400]"""
import argparse, os, sys
import ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="B")
    ap.add_argument("--chains", type=int, default=148)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--reverse", action="store_true", help="signals in reverse order (does a slow CTA follow its samples?)")
    a = ap.parse_args()
    from sxmc_b200 import _lib, mcmc, workloads
    L = _lib.lib()
    _lib.check(L.sxmc_init(0))
    fit = workloads.WORKLOADS[a.workload](a.scale)
    if a.reverse:
        fit.signals.reverse()
    hists, mb = mcmc.build_fit(fit, batch=a.chains)
    for k in range(a.chains):
        mb.set_seed(k, 5, k)
        mb.set_data(k, fit.data)
    _lib.check(L.sxmc_batch_phase_stamps(mb._b, 1, None))
    mb.run(a.chains, a.steps, 0.1)
    mb.run(a.chains, a.steps, 0.1)
    raw = np.zeros(148 * 8 + 16, dtype=np.uint64)
    _lib.check(L.sxmc_batch_phase_stamps(mb._b, 1, raw.ctypes.data_as(C.POINTER(C.c_uint64))))
    t = raw[:148 * 8].reshape(148, 8).astype(np.int64)
    c = raw[148 * 8:148 * 8 + 5].astype(np.int64)
    print("  finish_step of chain 0 [clk]: nll total %d, decide %d, row + stats %d, propose %d" %
          tuple(int(c[i + 1] - c[i]) for i in range(4)))
    base = t[:, 0].min()
    rel = (t[:, :8] - base) / 1e3
    per = 1e3 * mb.loop_ms() / (a.steps + 1)
    print("workload %s, %d chains: %.1f us/super-step = %.3f us/chain-step (path, max edges = %s)" %
          (a.workload, a.chains, per, per / a.chains, mb.last_path()))
    names = ["start", "bin done", "barrier1 out", "pdf table", "tail done", "barrier2 out", "event sum", "hist cleared"]
    for k, n in enumerate(names):
        if k in (3, 6, 7):
            own = rel[:a.chains, k]        # only CTAs that own a chain stamp these
            print("  %-13s min/mean/max %8.2f %8.2f %8.2f us" % (n, own.min(), own.mean(), own.max()))
            continue
        print("  %-13s min/mean/max %8.2f %8.2f %8.2f us" % (n, rel[:, k].min(), rel[:, k].mean(), rel[:, k].max()))
    d = rel[:, 1] - rel[:, 0]
    print("  bin phase per CTA: min/mean/max %.2f %.2f %.2f us -> %.3f us per chain (mean)" %
          (d.min(), d.mean(), d.max(), d.mean() / a.chains))
    print("  bin phase per CTA [us]:")
    for r0 in range(0, 148, 16):
        print("   %3d: " % r0 + " ".join("%5.0f" % v for v in d[r0:r0 + 16]))


if __name__ == "__main__":
    main()
