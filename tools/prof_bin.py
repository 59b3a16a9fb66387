#!/usr/bin/env python
"""Where a launch of the fast binning kernel spends its time: %globaltimer stamps of every CTA
(sxmc_bin_debug_stamps) reduced to prologue / steady state / epilogue and the tail imbalance.

  python tools/prof_bin.py --workload D [--scale 1.0] [--shard 8] [--steps 6] [--no-flush]

--shard G bins rank 0's 1/G slice of every signal (the sample-sharded step's binning launch)."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def q(a):
    a = np.asarray(a, dtype=np.float64)
    return "min %7.1f  med %7.1f  max %7.1f" % (a.min(), np.median(a), a.max())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="D")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--shard", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--stress", action="store_true")
    ap.add_argument("--json", default=None)
    ap.add_argument("--block", type=int, default=1,
                    help="steps back to back before the stamps of the LAST launch are read (1: isolated launches)")
    ap.add_argument("--global-hist", action="store_true", help="histograms in global memory (L2 atomics only)")
    a = ap.parse_args()
    from sxmc_b200 import _lib, mcmc, sharding, workloads
    L = _lib.lib()
    _lib.check(L.sxmc_init(0))
    fit = workloads.config_D(a.scale, stress=True) if a.stress else workloads.WORKLOADS[a.workload](a.scale)
    if a.shard > 1:
        fit = sharding.shard_fit_samples(fit, 0, a.shard)
    nbytes = 4 * fit.needed_columns() * fit.total_samples()
    from sxmc_b200 import pdfz
    if a.global_hist:
        pdfz.set_bin_tuning(smem_hist_max_bins=0)
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(1, 0)
    chain.begin(fit.data, a.steps * a.block + 4, 0.1)
    flush = 0 if a.no_flush else (256 << 20)
    chain.advance_timed(3, flush)
    _lib.check(L.sxmc_bin_debug_stamps(1, None))
    out = np.zeros((256, 32), dtype=np.uint64)
    rows = []
    for _ in range(a.steps):
        s, b, n = chain.advance_timed(a.block, flush)
        b /= a.block
        _lib.check(L.sxmc_bin_debug_stamps(1, out.ctypes.data_as(C.POINTER(C.c_uint64))))
        st = out.astype(np.int64)
        endc = 31
        live = st[:, endc] > 0
        g = st[live]
        t0 = g[:, 0].min()
        us = lambda col: (g[:, col] - t0) / 1e3
        end = us(endc)
        warps = (g[:, 16:32 - 0][:, :16] - t0) / 1e3      # per-warp loop exit of the last segment
        warps = np.where(g[:, 16:32] > 0, warps, np.nan)[:, :16]
        rows.append({
            "event_ms": b, "grid": int(live.sum()), "kernel_us": float(end.max()),
            "entry": us(0), "init": us(1), "first_data": us(2), "seg0_params": us(4),
            "nseg": g[:, 3], "end": end, "warps_last": warps,
            "seg0_loop": us(5), "seg0_done": us(6), "seg1_params": us(7), "seg1_loop": us(8), "seg1_done": us(9),
            "smid": g[:, 11] - 1,
        })
        out[:] = 0
    _lib.check(L.sxmc_bin_debug_stamps(0, None))
    chain.end()
    r = rows[-1]
    ev = np.array([x["event_ms"] for x in rows]) * 1e3
    ks = np.array([x["kernel_us"] for x in rows])
    print("workload %s%s scale %g shard 1/%d: %.1f MB per launch, grid %d" %
          (fit.name, " (stress)" if a.stress else "", a.scale, a.shard, nbytes / 1e6, r["grid"]))
    print("event-timed bin launch (us):      %s   -> %.0f GB/s" % (q(ev), nbytes / np.median(ev) / 1e3))
    print("first CTA entry -> last CTA exit: %s   -> %.0f GB/s" % (q(ks), nbytes / np.median(ks) / 1e3))
    print("outside the kernel (launch + drain): %.1f us (median)" % (np.median(ev) - np.median(ks)))
    print("last launch, us since the first CTA's entry, over CTAs:")
    print("  CTA entry                 %s" % q(r["entry"]))
    print("  pipeline initialised      %s" % q(r["init"]))
    print("  first data in smem        %s" % q(r["first_data"]))
    if (r["seg0_params"] > -1e6).all():      # (the per-warp pass of global-histogram launches has no segments: stamp unset)
        print("  seg 0 parameters resolved %s" % q(r["seg0_params"]))
    print("  segments per CTA          %s" % q(r["nseg"]))
    two = r["nseg"] >= 2
    if two.any():
        gap = (r["seg1_params"] - r["seg0_loop"])[two]
        print("  segment switch (warp 0 out of loop -> next segment's parameters): %s  (%d CTAs)" % (q(gap), two.sum()))
        print("    of which merge + clear  %s" % q((r["seg0_done"] - r["seg0_loop"])[two]))
    w = r["warps_last"]
    wmax, wmin = np.nanmax(w, axis=1), np.nanmin(w, axis=1)
    print("  last segment: first warp out %s" % q(wmin))
    print("                last warp out  %s" % q(wmax))
    print("                spread in CTA  %s" % q(wmax - wmin))
    print("  CTA exit                  %s" % q(r["end"]))
    print("  epilogue (last warp out -> exit) %s" % q(r["end"] - wmax))
    busy = r["end"] - r["entry"]
    print("  CTA lifetime              %s" % q(busy))
    print("  idle tail (kernel end - CTA exit), mean %.1f us = %.1f %% of the kernel" %
          (float((r["end"].max() - r["end"]).mean()), 100 * float((r["end"].max() - r["end"]).mean()) / r["end"].max()))
    # are the same SMs slow launch after launch?  (lifetime of single-segment CTAs, by %smid)
    life = {}
    for x in rows:
        one = x["nseg"] == 1
        for sm, t in zip(x["smid"][one], (x["end"] - x["entry"])[one]):
            life.setdefault(int(sm), []).append(float(t))
    sms = sorted(k for k, v in life.items() if len(v) == len(rows))
    if len(sms) > 8 and len(rows) >= 3:
        m = np.array([life[k] for k in sms])            # [sm, launch]
        rel = m / m.mean(axis=0, keepdims=True)
        avg = rel.mean(axis=1)
        cc = np.corrcoef(rel.T)
        print("per-SM relative lifetime (single-segment CTAs, %d SMs, %d launches): launch-to-launch correlation "
              "%.2f; persistent spread min %.3f max %.3f std %.3f" %
              (len(sms), len(rows), float(cc[np.triu_indices(len(rows), 1)].mean()), avg.min(), avg.max(), avg.std()))
        order = np.argsort(avg)
        print("  fastest SMs (smid: rel): " + " ".join("%d:%.3f" % (sms[i], avg[i]) for i in order[:10]))
        print("  slowest SMs (smid: rel): " + " ".join("%d:%.3f" % (sms[i], avg[i]) for i in order[-10:]))
        print("  by smid: " + " ".join("%.3f" % avg[i] for i in range(len(sms))))
    if a.json:
        with open(a.json, "w") as f:
            json.dump({"workload": fit.name, "scale": a.scale, "shard": a.shard, "bytes": nbytes,
                       "event_us": ev.tolist(), "kernel_us": ks.tolist(),
                       "cta_exit_us": r["end"].tolist(), "cta_entry_us": r["entry"].tolist()}, f)


if __name__ == "__main__":
    main()
