#!/usr/bin/env python
"""The reference's own two benchmarks, with its metric (bench/bench_sxmc.cpp:34-225):
  pdfz        10M N(0,1) samples, 1-D, 1000 bins on [-3,3), 100k eval points, shift = 0
  pdfz_group  29 PDFs of 10^3 ... 3x10^6 samples, same binning, evaluated together
metric = nsamples * nreps / wall clock of  nreps x (EvalAsync on every PDF; EvalFinished on every
PDF)  through the pdfz::EvalHist mirror (host calls, launches and the stream sync inside).
Also the device-only rate from CUDA events around the batched launch.
  python tools/bench_pdfz.py [--reps 100]"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
K40 = 2.99546e9      # README.md:322, the best published number (Tesla K40)


def run(name, reps):
    from sxmc_b200 import _lib, pdfz, workloads
    L = _lib.lib()
    fit = workloads.WORKLOADS[name](1.0)
    ns, ne = fit.nsignals, fit.nevents
    params = pdfz.Array.from_host([0.0], np.float64)
    norms = pdfz.Array(ns, np.uint32)
    lut = pdfz.Array(ns * ne, np.float32)
    hists = []
    for j, sig in enumerate(fit.signals):
        h = pdfz.EvalHist(sig.samples, sig.samples.shape[1], 1, fit.lower, fit.upper, fit.nbins, sig.dataset)
        h.AddSystematic(pdfz.ShiftSystematic(0, pdfz.Array.from_host([0], np.int16)))
        h.SetEvalPoints(fit.data)
        h.SetPDFValueBuffer(lut, j * ne, 1)
        h.SetNormalizationBuffer(norms, j)
        h.SetParameterBuffer(params, 0)
        h.EvalAsync()          # warm-up (builds the layout), as the reference does
        h.EvalFinished()
        hists.append(h)
    total = fit.total_samples()
    harr = (_lib.c_vp * ns)(*[h._h for h in hists])

    def batched():            # the for-all EvalAsync / for-all EvalFinished idiom as one launch
        _lib.check(L.sxmc_hists_eval_async(harr, ns, 1))
        _lib.check(L.sxmc_hist_eval_finished(hists[0]._h))

    for _ in range(5):
        batched()
    _lib.check(L.sxmc_enable_kernel_timing(1))
    t0 = time.perf_counter()
    for _ in range(reps):
        batched()
    wall = time.perf_counter() - t0
    _lib.check(L.sxmc_enable_kernel_timing(0))
    # the reference's literal loop: one EvalAsync per PDF, then one EvalFinished per PDF
    t0 = time.perf_counter()
    for _ in range(reps):
        for h in hists:
            h.EvalAsync()
        for h in hists:
            h.EvalFinished()
    wall_each = time.perf_counter() - t0
    out = {"benchmark": name, "nsignals": ns, "samples_total": total, "eval_points": ne, "reps": reps,
           "samples_per_s_batched_call": total * reps / wall, "us_per_rep_batched_call": 1e6 * wall / reps,
           "samples_per_s_per_pdf_calls": total * reps / wall_each,
           "published_best": {"value": K40, "hardware": "Tesla K40", "source": "README.md:322"}}
    if name == "bench_pdfz":
        out["vs_published_best"] = out["samples_per_s_batched_call"] / K40
    for h in hists:
        h.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=100)
    a = ap.parse_args()
    from sxmc_b200 import _lib
    _lib.check(_lib.lib().sxmc_init(0))
    for name in ("bench_pdfz", "pdfz_group"):
        print(json.dumps(run(name, a.reps)), flush=True)


if __name__ == "__main__":
    main()
