#!/usr/bin/env python
"""Phase breakdown of the persistent chain kernel (CTA 0 %globaltimer stamps)."""
import argparse, os, sys
import ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="B")
    ap.add_argument("--steps", type=int, default=512)
    ap.add_argument("--scale", type=float, default=1.0)
    a = ap.parse_args()
    from sxmc_b200 import _lib, mcmc, workloads
    L = _lib.lib()
    _lib.check(L.sxmc_init(0))
    fit = workloads.WORKLOADS[a.workload](a.scale)
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(1, 0)
    chain.set_mode(2)
    _lib.check(L.sxmc_chain_phase_stamps(chain._c, 1, None))
    chain.begin(fit.data, 2 * a.steps, 0.1)
    chain.advance(a.steps)
    ms0 = chain.loop_ms()
    chain.advance(a.steps)
    ms = chain.loop_ms() - ms0
    out = np.zeros((64, 8), dtype=np.uint64)
    _lib.check(L.sxmc_chain_phase_stamps(chain._c, 1, out.ctypes.data_as(C.POINTER(C.c_uint64))))
    chain.end()
    t = out.astype(np.int64)
    d = np.diff(t, axis=1)[8:56]            # steady-state steps
    nxt = t[9:57, 0] - t[8:56, 7]
    names = ["bin pass", "bin barrier", "lookup+reduce", "arrive+wait all", "partials+terms", "finish", "publish", "-> next step"]
    print("workload %s persistent: %.2f us/step over %d steps" % (a.workload, 1e3 * ms / a.steps, a.steps))
    for k in range(7):
        print("  %-16s %7.2f us" % (names[k], d[:, k].mean() / 1e3))
    print("  %-16s %7.2f us" % (names[7], nxt.mean() / 1e3))

if __name__ == "__main__":
    main()
