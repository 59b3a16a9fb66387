#!/usr/bin/env python
"""SURVEY 8(d) figures beyond the headline line, one JSON object per workload:
steps/s with floating and with fixed systematics, MC samples binned/s, roofline fraction of the
binning kernel (cold L2, CUDA events per kernel), production-path steps/s.
  python tools/bench_extra.py [--workloads A,B,C,D,bench_pdfz] > profiles/rNN_extra.jsonl"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
STEPS = {"A": (400, 2000), "B": (400, 2000), "bench_pdfz": (100, 400), "C": (30, 200), "D": (20, 100)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="A,B,bench_pdfz,C,D")
    a = ap.parse_args()
    import bench
    from sxmc_b200 import _lib, mcmc, workloads
    _lib.check(_lib.lib().sxmc_init(0))
    peak, peak_src = bench.measured_peak()
    for name in a.workloads.split(","):
        fit = workloads.WORKLOADS[name](1.0)
        kt, kp = STEPS[name]
        out = {"workload": name, "mc_samples_per_step": fit.total_samples(), "nevents": fit.nevents,
               "nsignals": fit.nsignals, "total_nbins": fit.total_nbins}
        for fixed in (False, True):
            for y in fit.systematics:
                y.fixed = fixed
            hists, chain = mcmc.build_fit(fit)
            for h in hists:
                h.Prepare()
            chain.set_seed(11, 0)
            chain.begin(fit.data, 2 * (kt + kp) + 64, 0.1)
            chain.advance_timed(8, bench.L2_FLUSH_BYTES)
            step_ms, bin_ms, nll_ms = chain.advance_timed(kt, bench.L2_FLUSH_BYTES)
            chain.advance(32)
            ms0 = chain.loop_ms()
            chain.advance(kp)
            prod_ms = chain.loop_ms() - ms0
            path = chain.last_path()
            chain.end()
            key = "fixed" if fixed else "floating"
            out[key] = {"steps_per_s_cold_l2": kt / (step_ms * 1e-3), "bin_ms": bin_ms / kt, "nll_ms": nll_ms / kt,
                        "steps_per_s_production": kp / (prod_ms * 1e-3),
                        "production_path": {1: "graph", 2: "persistent", 4: "monotone"}.get(path, path)}
            if not fixed:
                ab = bench.algorithmic_bytes_per_step(fit)
                ach = ab / (bin_ms / kt * 1e-3) / 1e9
                out["samples_binned_per_s_cold_l2"] = fit.total_samples() / (bin_ms / kt * 1e-3)
                out["samples_binned_per_s_production"] = fit.total_samples() * kp / (prod_ms * 1e-3)
                out["roofline"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                   "algorithmic_bytes_per_launch": ab, "peak_source": peak_src}
            chain.close()
            for h in hists:
                h.close()
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
