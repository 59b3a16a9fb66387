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
    raw = np.zeros(2048, dtype=np.uint64)
    _lib.check(L.sxmc_chain_phase_stamps(chain._c, 1, raw.ctypes.data_as(C.POINTER(C.c_uint64))))
    out = raw[:512].reshape(64, 8)
    cta = raw[512:512 + 148 * 4].reshape(148, 4).astype(np.int64)
    chain.end()
    t = out.astype(np.int64)
    d = np.diff(t, axis=1)[8:56]            # steady-state steps
    nxt = t[9:57, 0] - t[8:56, 7]
    names = ["bin pass", "bin barrier", "lookup+reduce", "arrive+wait all", "partials+terms", "finish", "publish", "-> next step"]
    print("workload %s persistent: %.2f us/step over %d steps" % (a.workload, 1e3 * ms / a.steps, a.steps))
    for k in range(7):
        print("  %-16s %7.2f us" % (names[k], d[:, k].mean() / 1e3))
    print("  %-16s %7.2f us" % (names[7], nxt.mean() / 1e3))

    dbg = raw[1200:1200 + 148 * 4].reshape(148, 4).astype(np.int64)
    dd = np.diff(dbg, axis=1) / 1e3
    print("  inside the bin pass (step 20; last segment of each CTA): setup  chunks  merge  [us] mean/max")
    print("   %.2f/%.2f  %.2f/%.2f  %.2f/%.2f" % (dd[:, 0].mean(), dd[:, 0].max(), dd[:, 1].mean(), dd[:, 1].max(), dd[:, 2].mean(), dd[:, 2].max()))
    base = cta[:, 0].min()
    rel = (cta - base) / 1e3
    print("  per-CTA (step 20), us since first CTA started: start  bin-end  barrier-out  lookup-end")
    for q in (0, 1, 37, 73, 74, 110, 147):
        print("   cta %3d: %6.2f %6.2f %6.2f %6.2f" % (q, *rel[q]))
    print("   start   min/mean/max: %.2f %.2f %.2f" % (rel[:, 0].min(), rel[:, 0].mean(), rel[:, 0].max()))
    print("   bin dur min/mean/max: %.2f %.2f %.2f" % tuple(f((rel[:, 1] - rel[:, 0])) for f in (np.min, np.mean, np.max)))
    print("   bin-end min/mean/max: %.2f %.2f %.2f" % (rel[:, 1].min(), rel[:, 1].mean(), rel[:, 1].max()))
    print("   bar-out min/mean/max: %.2f %.2f %.2f" % (rel[:, 2].min(), rel[:, 2].mean(), rel[:, 2].max()))
    print("   lookup dur min/mean/max: %.2f %.2f %.2f" % tuple(f((rel[:, 3] - rel[:, 2])) for f in (np.min, np.mean, np.max)))


if __name__ == "__main__":
    main()
