#!/usr/bin/env python
"""bench.py -- the driver-facing benchmark of the sxmc per-step likelihood hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload B]

One "step" = one MCMC step of the named workload with floating systematics: re-bin every
signal's Monte-Carlo samples under the proposed systematic parameters (zero + transform +
bin), look up the PDFs at the data events, reduce the event likelihood, total the NLL with
the Gaussian constraints, Metropolis accept/reject, propose the next point
(reference src/mcmc.cpp:261-378).  Workload names are SURVEY.md section 8's A-D; the
default is B = BASELINE.json configs[1].

Own arm (`--impl b200`, default), per rank = per GPU:
  value  MCMC steps/s, inputs resident in HBM, each step timed with CUDA events on the
         chain's stream with an L2 flush (256 MiB memset) before it, outside the interval;
         max over ranks of the summed step time; N ranks run N independent chains
         (ensemble sharding, no data-path collective) => weak scaling.
  e2e    the same metric through the public API call a user makes -- MCMC.__call__
         (sxmc_chain_run + sxmc_chain_read over the C ABI) with HOST buffers: upload of the
         data events, SetEvalPoints, initial evaluation, graph capture, the K-step loop and
         the device->host read of the chain rows are all inside the timed region.
  roofline / cpu_baseline / clocks / gpu_launches as the contract asks.

Reference arm (`--impl reference`): the UNMODIFIED reference sources compiled for the CPU
(oracle/_ref/libsxmc_ref.so, the hemi host-loop build; falls back to the C restatement
oracle/liboracle_port.so when that library is absent), same workload, a bounded number of
full-size steps.  oracle/ is only ever the checker / the timed CPU baseline here.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "mcmc_steps_per_sec"
UNIT = "steps/s"
L2_FLUSH_BYTES = 256 << 20          # > 126 MB L2

WORKLOAD_TEXT = {
    "A": "A: config/example.json shape -- 1-D energy (10 bins), 2 signals (500k, 1M MC samples), "
         "scale + resolution_scale systematics, 2k data events",
    "B": "B: 1-D energy PDF (1000 bins), 3 signals, 1M MC samples/signal, shift systematic, "
         "10k data events",
    "C": "C: 2-D energy x radius (100x50 bins), 10 signals, 10M MC samples/signal, scale + "
         "resolution_scale systematics, 100k data events",
    "D": "D: 3-D energy x radius x costheta (50x25x20 bins), 30 signals, 100M MC samples in "
         "total, scale + resolution_scale + shift systematics, 100k data events",
    "E": "E: ensemble of fake experiments on the workload-B fit (Poisson fake data, mean 10k events; "
         "20000 MCMC steps, burn-in 0.1, 90 % contour limit per experiment), experiments sharded "
         "round-robin over the GPUs",
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="B", choices=sorted(WORKLOAD_TEXT))
    ap.add_argument("--scale", type=float, default=1.0, help="scale MC sample counts (debug)")
    ap.add_argument("--cpu-budget-s", type=float, default=20.0,
                    help="CPU seconds the cpu_baseline / reference arm may spend stepping")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra workload-C / ensemble objects")
    ap.add_argument("--shard-samples", action="store_true",
                    help="N>1: ONE chain whose Monte-Carlo samples are cut over the GPUs, histograms "
                         "summed with NCCL every step (strong scaling; BASELINE.json configs[3])")
    ap.add_argument("--p2p", action="store_true",
                    help="with --shard-samples: exchange the histograms with the peer-memory kernel "
                         "(p2p_kernels.cu) instead of ncclAllReduce")
    ap.add_argument("--p2p-form", type=int, default=0,
                    help="with --p2p: 0 automatic (reduce-scatter + all-gather above 2 ranks), 1 one-shot, 2 two-shot")
    ap.add_argument("--experiments", type=int, default=296,
                    help="workload E: fake experiments per GPU (whole batches of 148)")
    ap.add_argument("--esteps", type=int, default=20000, help="workload E: MCMC steps per experiment")
    return ap.parse_args()


def measured_peak():
    """HBM GB/s: MEASURED_PEAKS.json (driver-written) else the profiling guide's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(name):
    """dram bytes per launch of the binning kernel from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01b_ncu_traffic.json")) as f:
            return json.load(f).get(name)
    except Exception:
        return None


def make_fit(name, scale):
    from sxmc_b200 import workloads
    return workloads.WORKLOADS["B" if name == "E" else name](scale)


def algorithmic_bytes_per_step(fit):
    """4 B x (observables + truth fields the math reads) per MC sample."""
    return 4 * fit.needed_columns() * fit.total_samples()


# ---------------------------------------------------------------------------------------
# clocks during the timed region (NVML, sampled from a thread)
# ---------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.002)

    def start(self):
        if self.nv is not None:
            self._stop.clear()
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        if self._t is not None:
            self._stop.set()
            self._t.join()
            self._t = None

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------
# CPU arms: the reference's own CPU build (oracle/_ref) or the C restatement (oracle/port)
# ---------------------------------------------------------------------------------------
def cpu_steps_per_s(fit, budget_s, want_steps=None):
    """Times full-size MCMC steps of `fit` on ONE host core (the reference is strictly
    single-threaded).  Returns (steps/s, kind, nsteps timed, description)."""
    import oracle
    from oracle import port
    kind = "port"
    hists = None
    if oracle.have_ref() or os.path.isdir(os.path.join(oracle.REFERENCE_ROOT, "src")):
        from oracle import ref
        kind = "reference"
        hists = [ref.make_hist(fit, j) for j in range(fit.nsignals)]

    def run(n):
        rng = np.random.default_rng(7)
        normals = rng.normal(size=(n + 1) * fit.nparameters)
        uniforms = rng.random(max(1, n))
        t0 = time.perf_counter()
        if kind == "reference":
            ref.mcmc(fit, n, 0.1, normals, uniforms, hists=hists)
        else:
            port.mcmc(fit, n, 0.1, normals, uniforms)
        return time.perf_counter() - t0

    # MCMC::operator() = setup + initial evaluation + n steps: time n1 and n2, use the slope
    n1 = 1
    t1 = run(n1)
    per_step = max(t1 / 2.0, 1e-6)          # setup+init ~ one step
    n2 = int(max(3, min(budget_s / per_step, want_steps or 1e9, 5000)))
    t2 = run(n1 + n2)
    rate = n2 / max(t2 - t1, 1e-9)
    if hists is not None:
        for h in hists:
            h.close()
    what = ("%d full-size steps of workload %s (MCMC::operator() timed at %d and %d steps, slope), "
            "1 host thread" % (n2, fit.name, n1, n1 + n2))
    return rate, kind, n2, what


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    fit = make_fit(args.workload, args.scale)
    rate, kind, n, what = cpu_steps_per_s(fit, args.cpu_budget_s, want_steps=args.steps)
    if args.workload == "E":
        # one experiment = esteps full-size steps (fake-data generation and the limit are
        # negligible beside them on the CPU): experiments/s from the measured step rate
        erate = rate / args.esteps
        what_e = what + "; experiments/s = steps/s / %d steps per experiment" % args.esteps
        print(json.dumps({
            "impl": "reference", "metric": "ensemble_experiments_per_sec", "value": erate,
            "unit": "experiments/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 / rate, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic (seeded numpy, SURVEY 8 distributions)",
            "config": {"workload": WORKLOAD_TEXT["E"], "steps_per_experiment": args.esteps},
            "chain_steps_per_s": rate,
            "cpu_baseline": {"value": erate, "unit": "experiments/s", "cores": 1, "kind": kind,
                             "sample": what_e, "host_cores_available": os.cpu_count()},
            "e2e": {"value": erate, "unit": "experiments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}), flush=True)
        return
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / rate,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (seeded numpy, SURVEY 8 distributions)",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "scale": args.scale,
                   "mc_samples_per_step": fit.total_samples(), "nevents": fit.nevents},
        "samples_binned_per_s": rate * fit.total_samples(),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": 1, "kind": kind, "sample": what,
                         "host_cores_available": os.cpu_count()},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------
# own arm
# ---------------------------------------------------------------------------------------
def dist_setup(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def measure_fit(fit, args, local, rank, sampler, do_e2e=True):
    """Builds the fit on this rank's GPU and measures it.  Returns a dict of raw numbers."""
    import torch
    from sxmc_b200 import mcmc
    K, Wm = args.steps, args.warmup
    t_build0 = time.perf_counter()
    comm = None
    if getattr(args, "shard_samples", False) and torch.distributed.is_initialized():
        from sxmc_b200 import sharding
        world = torch.distributed.get_world_size()
        fit_local = sharding.shard_fit_samples(fit, rank, world)
        hists, chain = mcmc.build_fit(fit_local)
        if getattr(args, "p2p", False):
            comm = "p2p"
            from sxmc_b200 import pdfz as _pdfz
            _pdfz.set_bin_tuning(p2p_form=getattr(args, "p2p_form", 0))
            sharding.connect_p2p(chain, rank, world)
        else:
            comm = sharding.Comm(device=local)
            chain.set_comm(comm.handle)
    else:
        hists, chain = mcmc.build_fit(fit)
    for h in hists:
        h.Prepare()                      # sorted, tiled SoA layout (first-call Optimize analogue)
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t_build0
    # independent chain per rank (ensemble sharding) / the same chain on every rank (sample sharding)
    chain.set_seed(20261104, 0 if comm is not None else rank)

    out = {"build_s": build_s}
    total = 2 * Wm + 2 * K
    chain.begin(fit.data, total, 0.1)    # re-tunes at 10 % and 20 % of all steps, as in a real fit

    def barrier():
        torch.cuda.synchronize()
        if torch.distributed.is_initialized():
            torch.distributed.barrier()
        torch.cuda.synchronize()

    # (1) value: per-step CUDA events, L2 flushed before every step (outside the intervals)
    chain.advance_timed(Wm, L2_FLUSH_BYTES)
    barrier()
    sampler.start()
    l0 = chain.launches()
    if comm == "p2p":
        clk0 = sharding.p2p_clocks(chain)
    step_ms, bin_ms, nll_ms = chain.advance_timed(K, L2_FLUSH_BYTES)
    out["comm_ms"] = chain.last_comm_ms()
    if comm == "p2p":
        clk1 = sharding.p2p_clocks(chain)
        out["p2p_us"] = [round((b - a) / K * 1e-3, 3) for a, b in zip(clk0, clk1)]
    out["gpu_launches"] = chain.launches() - l0
    barrier()
    out.update(step_ms=step_ms, bin_ms=bin_ms, nll_ms=nll_ms)

    # (2) the production form: sxmc_chain_advance picks the kernel (monotone / persistent / graph)
    chain.advance(Wm)
    barrier()
    ms0 = chain.loop_ms()
    if comm == "p2p":
        clk0 = sharding.p2p_clocks(chain)
    chain.advance(K)
    out["warm_ms"] = chain.loop_ms() - ms0
    if comm == "p2p":
        clk1 = sharding.p2p_clocks(chain)
        out["p2p_us_warm"] = [round((b - a) / K * 1e-3, 3) for a, b in zip(clk0, clk1)]
    barrier()
    ls = chain.end()
    out["accepted"] = chain.accepted()
    out["rows"] = int(ls.get_samples().shape[0])

    # (3) e2e through the public call with host buffers
    if do_e2e:
        chain(fit.data, max(32, Wm), 0.1)          # warm-up of the whole call
        barrier()
        t0 = time.perf_counter()
        ls = chain(fit.data, K, 0.1)
        rows = ls.get_samples()
        torch.cuda.synchronize()
        out["e2e_s"] = time.perf_counter() - t0
        out["e2e_h2d"] = fit.data.nbytes / K
        out["e2e_d2h"] = (rows.nbytes + 4096) / K     # rows + the chain-state readback
        barrier()
    sampler.stop()
    chain.close()
    for h in hists:
        h.close()
    if comm is not None:
        torch.cuda.synchronize()
        torch.distributed.barrier()
        if comm != "p2p":
            comm.close()
    return out


def measure_ensemble(fit, nexperiments, esteps, rank, world, seed=20261104):
    """Config E on this rank: its share of `nexperiments * world` experiments through
    sxmc_b200.ensemble.ensemble() -- fake data drawn on the device, chains advanced 148 at a time by
    the batch kernel, contour limit per experiment.  Wall clock of the whole call (host logic,
    data generation, launches, limit extraction)."""
    import torch
    from sxmc_b200 import ensemble
    ensemble.ensemble(fit, 4 * world, nsteps=200, burnin_fraction=0.1, rank=rank, world=world)   # warm-up
    torch.cuda.synchronize()
    if torch.distributed.is_initialized():
        torch.distributed.barrier()
    t0 = time.perf_counter()
    r = ensemble.ensemble(fit, nexperiments * world, "s0", 0.9, seed=seed, nsteps=esteps,
                          burnin_fraction=0.1, rank=rank, world=world)
    torch.cuda.synchronize()
    r["wall_s"] = time.perf_counter() - t0
    return r


def run_ensemble_arm(args, rank, world, local):
    import torch
    from sxmc_b200 import _lib, sharding
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.check(_lib.lib().sxmc_init(local))
    sampler = ClockSampler(local)
    fit = make_fit("E", args.scale)
    sampler.start()
    r = measure_ensemble(fit, args.experiments, args.esteps, rank, world)
    sampler.stop()
    wall = allreduce_max(r["wall_s"])
    loop_ms = allreduce_max(r["loop_ms"])
    limits = sharding.gather_limits(r["experiments"], r["limits"], args.experiments * world)
    total = args.experiments * world
    line = {
        "metric": "ensemble_experiments_per_sec", "value": total / wall, "unit": "experiments/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * wall / (total * args.esteps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic (seeded numpy, SURVEY 8 distributions; "
        "Poisson fake data drawn on the device)",
        "config": {"workload": WORKLOAD_TEXT["E"], "experiments": total, "steps_per_experiment": args.esteps,
                   "parallelism": "experiment i on GPU i mod N, Philox stream i; one gather of the limits",
                   "l2": "not applicable: the samples are resident in shared memory for the whole launch"},
        "chain_steps_per_s": total * args.esteps / wall,
        "us_per_chain_step_device": 1e3 * loop_ms / (args.experiments * args.esteps),
        "time_to_4096_experiments_s": 4096.0 / (total / wall),
        "median_upper_limit": sharding.median(limits), "mode": r["mode"],
        "gpu_launches": r["launches"],
        "e2e": {"value": total / wall, "unit": "experiments/s", "h2d_bytes_per_step": 0.0,
                "d2h_bytes_per_step": 64.0 / args.esteps,
                "what": "ensemble.ensemble(): the value above IS the end-to-end wall clock (fake data are "
                        "generated on the device; per experiment the limit comes back)"},
        "clocks": sampler.summary(),
    }
    if rank == 0:
        print(json.dumps(line), flush=True)
    if torch.distributed.is_initialized():
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def allreduce_max(x):
    import torch
    if not torch.distributed.is_initialized():
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def run_own_arm(args, rank, world, local):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
    from sxmc_b200 import _lib
    _lib.check(_lib.lib().sxmc_init(local))
    peak, peak_src = measured_peak()
    sampler = ClockSampler(local)

    fit = make_fit(args.workload, args.scale)
    K = args.steps
    r = measure_fit(fit, args, local, rank, sampler)
    p2p_phases = None
    if "p2p_us" in r:
        box = [None] * world
        torch.distributed.all_gather_object(box, (r["p2p_us"], r.get("p2p_us_warm")))
        p2p_phases = {"what": "per rank [wait for every peer's partial, sum, close], device globaltimer",
                      "cold": [b[0] for b in box], "back_to_back": [b[1] for b in box]}
    step_ms = allreduce_max(r["step_ms"])
    warm_ms = allreduce_max(r["warm_ms"])
    e2e_s = allreduce_max(r["e2e_s"])
    sharded = args.shard_samples and world > 1
    nchains = 1 if sharded else world
    value = nchains * K / (step_ms * 1e-3)
    alg_bytes = algorithmic_bytes_per_step(fit) // (world if sharded else 1)
    bin_avg_ms = r["bin_ms"] / K
    achieved = alg_bytes / (bin_avg_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": step_ms / K, "higher_is_better": True,
        "scaling": "strong" if sharded else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (seeded numpy, SURVEY 8 distributions)",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "scale": args.scale,
                   "mc_samples_per_step": fit.total_samples(), "nevents": fit.nevents,
                   "nparameters": fit.nparameters,
                   "l2": "flushed before every timed step (256 MiB memset outside the event "
                         "interval); the working set (%.1f MB) would otherwise be L2-resident"
                         % (alg_bytes / 1e6),
                   "parallelism": ("sample sharding: each GPU bins 1/%d of every signal's samples, uint32 "
                                   "histograms summed with %s every step, one chain"
                                   % (world, "one peer-memory kernel (NVLink P2P loads)" if args.p2p
                                      else "ncclAllReduce"))
                   if sharded else "1 independent chain per GPU (ensemble sharding), no collective"},
        "samples_binned_per_s": value * fit.total_samples(),
        "allreduce_ms_per_step": r["comm_ms"] / K,
        "p2p_kernel_us_per_step": p2p_phases,
        "e2e": {"value": nchains * K / e2e_s, "unit": UNIT,
                "h2d_bytes_per_step": r["e2e_h2d"], "d2h_bytes_per_step": r["e2e_d2h"],
                "what": "MCMC.__call__(host data, K steps, burn-in 0.1) -> host chain rows: "
                        "sxmc_chain_run + sxmc_chain_read"},
        "gpu_launches": r["gpu_launches"],
        "roofline": {"bound": "hbm", "kernel": "bin_fast_kernel (transform + binning, all signals "
                     "in one launch)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak,
                     "traffic": ncu_traffic(args.workload) if args.scale == 1.0 else None,
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "avg_launch_ms": bin_avg_ms, "share_of_step": r["bin_ms"] / r["step_ms"]},
        "kernels_ms_per_step": {"bin": bin_avg_ms, "lookup_nll_metropolis": r["nll_ms"] / K},
        "chain_back_to_back": {"value": nchains * K / (warm_ms * 1e-3), "unit": UNIT,
                               "ms_per_step": warm_ms / K,
                               "note": "production form: chain.advance(K), steps back to back (monotone / persistent kernel for small fits, CUDA graph otherwise)"},
        "accept_rate": r["accepted"] / float(2 * K + 2 * args.warmup),
        "build_s": r["build_s"],
    }
    return line, fit, sampler, peak, peak_src


def main():
    args = parse_args()
    rank, world, local = dist_setup(args)
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if args.workload == "E":
        run_ensemble_arm(args, rank, world, local)
        return
    line, fit, sampler, peak, peak_src = run_own_arm(args, rank, world, local)
    import torch

    # extra: the HBM-bound workload C on the same box (inputs >> L2), N=1 only
    if world == 1 and not args.no_extra and args.workload == "B" and args.scale == 1.0:
        try:
            import copy
            a2 = copy.copy(args)
            a2.steps, a2.warmup = 20, 3
            fitc = make_fit("C", 1.0)
            rc = measure_fit(fitc, a2, local, rank, ClockSampler(local), do_e2e=False)
            ab = algorithmic_bytes_per_step(fitc)
            ach = ab / (rc["bin_ms"] / a2.steps * 1e-3) / 1e9
            line["workload_C"] = {
                "workload": WORKLOAD_TEXT["C"], "steps": a2.steps,
                "steps_per_s": a2.steps / (rc["step_ms"] * 1e-3),
                "samples_binned_per_s": a2.steps / (rc["step_ms"] * 1e-3) * fitc.total_samples(),
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                             "frac": ach / peak, "traffic": ncu_traffic("C"),
                             "algorithmic_bytes_per_launch": ab,
                             "avg_launch_ms": rc["bin_ms"] / a2.steps},
                "nll_ms_per_step": rc["nll_ms"] / a2.steps,
                "build_s": rc["build_s"]}
            del fitc
        except Exception as e:      # the headline must still print
            line["workload_C"] = {"error": repr(e)}

    # extra: config E (ensemble) on the same fit, one batch of 148 experiments, N=1 only
    if world == 1 and not args.no_extra and args.workload == "B" and args.scale == 1.0:
        try:
            re_ = measure_ensemble(fit, 148, 20000, 0, 1)
            line["ensemble_E"] = {
                "workload": WORKLOAD_TEXT["E"], "experiments": 148, "steps_per_experiment": 20000,
                "experiments_per_s": 148 / re_["wall_s"], "chain_steps_per_s": 148 * 20000 / re_["wall_s"],
                "us_per_chain_step_device": 1e3 * re_["loop_ms"] / (148 * 20000), "mode": re_["mode"],
                "gpu_launches": re_["launches"],
                "note": "148 chains advance in lockstep in one cooperative launch (chain batch kernel)"}
        except Exception as e:
            line["ensemble_E"] = {"error": repr(e)}

    # extra: the reference's own published benchmark (bench/bench_sxmc.cpp:34-102)
    if world == 1 and not args.no_extra and args.workload == "B" and args.scale == 1.0:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_pdfz
            bp = bench_pdfz.run("bench_pdfz", 50)
            line["bench_pdfz"] = {"samples_per_s": bp["samples_per_s_batched_call"], "reps": 50,
                                  "what": "10M samples, 1-D, 1000 bins, 100k eval points: reps x (EvalAsync + "
                                          "EvalFinished) wall clock, the reference's metric",
                                  "published_best": bp["published_best"],
                                  "vs_published_best": bp["vs_published_best"]}
        except Exception as e:
            line["bench_pdfz"] = {"error": repr(e)}

    line["clocks"] = sampler.summary()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rate, kind, n, what = cpu_steps_per_s(fit, args.cpu_budget_s)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": 1, "kind": kind,
                                "sample": what, "host_cores_available": os.cpu_count()}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if torch.distributed.is_initialized():
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def _only_json_on_stdout():
    """Libraries under us print to stdout now and then (NCCL's version banner, the reference's
    progress lines in the CPU arm).  The contract is ONE JSON line there: everything written to
    fd 1 while the run lasts goes to stderr, and print() is pointed at the real stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real


if __name__ == "__main__":
    _only_json_on_stdout()
    main()
