#!/usr/bin/env python
"""Phase split of lookup_nll_kernel (PDF lookup + event likelihood + NLL + Metropolis + proposal)
from %globaltimer stamps (sxmc_nll_debug_stamps).

  python tools/prof_nll.py --workload D [--steps 6] [--block 1]"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="D")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--block", type=int, default=8, help="steps back to back before the stamps of the last launch are read")
    a = ap.parse_args()
    from sxmc_b200 import _lib, mcmc, workloads
    L = _lib.lib()
    _lib.check(L.sxmc_init(0))
    fit = workloads.WORKLOADS[a.workload](a.scale)
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(1, 0)
    chain.begin(fit.data, a.steps * a.block + 8, 0.1)
    chain.advance_timed(4, 0)
    _lib.check(L.sxmc_nll_debug_stamps(1, None))
    out = np.zeros((1202, 8), dtype=np.uint64)
    rows, evs = [], []
    for _ in range(a.steps):
        s, b, n = chain.advance_timed(a.block, 0)
        _lib.check(L.sxmc_nll_debug_stamps(1, out.ctypes.data_as(C.POINTER(C.c_uint64))))
        st = out.astype(np.int64)
        live = st[:1200, 0] > 0
        t0 = st[:1200][live][:, 0].min()
        fin = (st[1200] - t0) / 1e3
        blocks = (st[:1200][live] - t0) / 1e3
        rows.append((fin, blocks, int(st[1201, 0]), int(live.sum())))
        evs.append(1e3 * n / a.block)
        out[:] = 0
    _lib.check(L.sxmc_nll_debug_stamps(0, None))
    chain.end()
    fin, blocks, who, nb = rows[-1]
    print("workload %s: %d signals, %d bins, %d events; lookup_nll_kernel %d blocks; event-timed %.1f us (median of %d)" %
          (fit.name, fit.nsignals, fit.total_nbins, fit.nevents, nb, float(np.median(evs)), len(evs)))
    q = lambda v: "min %6.2f med %6.2f max %6.2f" % (np.min(v), np.median(v), np.max(v))
    names = ["entry", "state prefetched + next buffer cleared", "per-signal constants ready", "lookup loop done",
             "block sum done", "ticket taken"]
    print("over blocks, us since the first block's entry:")
    for k, nm in enumerate(names):
        col = blocks[:, k]
        col = col[col > -1e6]
        print("  %-42s %s" % (nm, q(col)))
    print("block that finished the step (block %d):" % who)
    for k, nm in enumerate(names + ["event total known (partials summed)", "step finished (NLL, decision, row, proposal)"]):
        print("  %-42s %8.2f" % (nm, fin[k]))
    med = np.median(np.array([r[0] for r in rows]), axis=0)
    print("median over %d launches of the finishing block's stamps: %s" % (len(rows), " ".join("%.1f" % v for v in med)))


if __name__ == "__main__":
    main()
