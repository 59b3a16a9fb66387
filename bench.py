#!/usr/bin/env python
"""bench.py -- the driver-facing benchmark of the sxmc per-step likelihood hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload D]

One "step" = one MCMC step of the named workload with floating systematics: re-bin every
signal's Monte-Carlo samples under the proposed systematic parameters (zero + transform +
bin), look up the PDFs at the data events, reduce the event likelihood, total the NLL with
the Gaussian constraints, Metropolis accept/reject, propose the next point
(reference src/mcmc.cpp:261-378).  Workload names are SURVEY.md section 8's A-E.

Default workload: **D** (BASELINE.json configs[3]: 3-D, 30 signals, 100M MC samples, 1.6 GB of
sample columns streamed per step) -- the largest BASELINE configuration that fits one GPU.
  N = 1   one chain on one GPU.
  N > 1   the SAME chain, sample-sharded: every GPU bins 1/N of every signal's samples, the
          uint32 histograms are summed over the GPUs every step by one peer-memory kernel over
          NVLink (p2p_kernels.cu; `--exchange nccl` for ncclAllReduce), every rank then takes
          the same Metropolis decision => "scaling": "strong".  `--replicas` runs N independent
          chains instead (ensemble sharding, no collective, weak scaling).

Own arm (`--impl b200`, default):
  value  MCMC steps/s with everything resident in HBM.  The K-step block is repeated
         `repeats` times (a 20-step block of D lasts 6 ms: one block is at the mercy of a single
         host hiccup); value = K / (max over ranks of the MEDIAN block); the per-rank min /
         median / max block is printed so a straggler is visible.  C and D stream 10x the L2 per
         step (no flush applies, none done): a block is K steps issued back to back the way a
         fit runs them (sxmc_chain_advance), one CUDA-event interval on the chain's stream
         around it; kernels_ms_per_step and the roofline come from a second pass with an event
         around every kernel (bin | exchange | lookup+NLL+Metropolis).  A/B/E-sized fits are
         timed step by step with a 256 MiB memset before each step, outside the intervals.
  e2e    the same metric through the public call a user makes -- MCMC.__call__
         (sxmc_chain_run + sxmc_chain_read over the C ABI) with HOST buffers: upload of the
         data events, SetEvalPoints, initial evaluation, graph capture, the K-step loop and
         the device->host read of the chain rows are all inside the timed region (wall clock,
         median of `repeats_e2e` calls, max over ranks).
  roofline / cpu_baseline / clocks / gpu_launches as the contract asks; nested objects carry
  the other BASELINE workloads (C, B, ensemble E, the reference's bench_pdfz shape, an
  nevents sweep of the NLL pass) measured in the same process.

Reference arm (`--impl reference`): the UNMODIFIED reference sources compiled for the CPU
(oracle/_ref/libsxmc_ref.so, the hemi host-loop build; falls back to the C restatement
oracle/liboracle_port.so when that library is absent), same workload, a bounded number of
full-size steps on one host thread (the reference is single-threaded).  oracle/ is only ever
the checker / the timed CPU baseline here.
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
L2_BYTES = 126 << 20
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


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="D", choices=sorted(WORKLOAD_TEXT))
    ap.add_argument("--scale", type=float, default=1.0, help="scale MC sample counts (debug)")
    ap.add_argument("--repeats", type=int, default=0,
                    help="timed K-step blocks (0: as many as make ~600 timed steps, 3..50)")
    ap.add_argument("--cpu-budget-s", type=float, default=20.0,
                    help="CPU seconds the cpu_baseline / reference arm may spend stepping")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the nested extra workloads")
    ap.add_argument("--replicas", action="store_true",
                    help="N>1: N independent chains (ensemble sharding, no collective) instead of one "
                         "sample-sharded chain")
    ap.add_argument("--shard-samples", action="store_true", help="(default for N>1; kept for old command lines)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N>1 sample sharding: histogram exchange by the peer-memory kernel or ncclAllReduce")
    ap.add_argument("--p2p", action="store_true", help="(default; kept for old command lines)")
    ap.add_argument("--p2p-form", type=int, default=0,
                    help="peer-memory exchange: 0 automatic (fused into a rank-sharded lookup kernel where the fit "
                         "allows it), 1 one-shot all-reduce kernel, 2 reduce-scatter + all-gather kernel")
    ap.add_argument("--flush", default="auto", choices=["auto", "always", "never"],
                    help="L2 flush (256 MiB memset) before every timed step; auto = only when a step "
                         "streams less than 4x the L2")
    ap.add_argument("--experiments", type=int, default=296,
                    help="workload E: fake experiments per GPU (whole batches of 148)")
    ap.add_argument("--esteps", type=int, default=20000, help="workload E: MCMC steps per experiment")
    return ap.parse_args(argv)


def measured_peak():
    """HBM GB/s: MEASURED_PEAKS.json (driver-written) else the profiling guide's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(name):
    """dram bytes per launch of the binning kernel from the committed ncu --set full captures."""
    for fn in ("r02_ncu_traffic.json", "r01b_ncu_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as f:
                v = json.load(f).get(name)
            if v is not None:
                return v
        except Exception:
            pass
    return None


def make_fit(name, scale):
    from sxmc_b200 import workloads
    return workloads.WORKLOADS["B" if name == "E" else name](scale)


def algorithmic_bytes_per_step(fit):
    """4 B x (observables + truth fields the math reads) per MC sample."""
    return 4 * fit.needed_columns() * fit.total_samples()


def wants_flush(args, fit):
    if args.flush == "always":
        return True
    if args.flush == "never":
        return False
    return algorithmic_bytes_per_step(fit) < 4 * L2_BYTES


def config_for(args, fit, world):
    """The `config` object, identical in both arms (it names the workload and how the b200 arm
    runs it; the reference arm always runs the same chain on one host thread)."""
    step_mb = algorithmic_bytes_per_step(fit) / 1e6
    sharded = world > 1 and not args.replicas and args.workload != "E"
    if sharded:
        par = ("sample sharding: each of the %d GPUs bins 1/%d of every signal's samples, uint32 histograms "
               "summed every step by %s, one chain" %
               (world, world, ("the lookup kernel itself, which is dealt over the GPUs: it reads its bins' partial counts "
                               "from every GPU's memory over NVLink and exchanges partial sums, not histograms"
                               if args.p2p_form in (0, 3) else "one peer-memory all-reduce kernel (NVLink P2P loads/stores)")
                if args.exchange == "p2p" else "ncclAllReduce"))
    elif world > 1:
        par = "%d independent chains, one per GPU (ensemble sharding), no collective" % world
    else:
        par = "one chain on one GPU"
    flush = wants_flush(args, fit)
    return {"workload": WORKLOAD_TEXT[args.workload], "scale": args.scale,
            "mc_samples_per_step": fit.total_samples(), "nevents": fit.nevents,
            "nparameters": fit.nparameters, "parallelism": par,
            "l2": ("flushed before every timed step (256 MiB memset outside the event interval); the "
                   "%.1f MB a step streams would otherwise be L2-resident" % step_mb) if flush else
                  ("inputs larger than L2: a step streams %.0f MB of sample columns through a 126 MB L2, "
                   "no flush" % step_mb)}


# ---------------------------------------------------------------------------------------
# clocks during the timed region (NVML, sampled from a thread)
# ---------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index, period_s=0.025):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.power_w, self.mem_mhz = [], []
        self.period_s = period_s
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
                self.mem_mhz.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_MEM)))
                self.power_w.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def start(self):
        if self.nv is not None and self._t is None:
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
                "reasons": sorted(self.reasons), "samples": len(self.samples),
                "sample_period_ms": 1e3 * self.period_s,
                "sm_mhz_min": float(np.min(self.samples)),
                "mem_mhz": float(np.median(self.mem_mhz)) if self.mem_mhz else None,
                "power_w_median": float(np.median(self.power_w)) if self.power_w else None,
                "power_w_max": float(np.max(self.power_w)) if self.power_w else None}


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


def cpu_baseline_object(rate, unit, kind, what):
    nproc = os.cpu_count() or 1
    return {"value": rate, "unit": unit, "cores": 1, "kind": kind, "sample": what,
            "host_cores_available": nproc,
            # independent chains (an ensemble) could run one per core; ONE chain cannot
            "cpu_ideal_parallel": {"value": rate * nproc, "unit": unit,
                                   "what": "1-core rate x %d host cores: the ideal process-parallel rate for "
                                           "INDEPENDENT chains (ensembles); a single chain is serial in the "
                                           "reference" % nproc}}


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
            "config": ensemble_config(args, world),
            "chain_steps_per_s": rate,
            "cpu_baseline": cpu_baseline_object(erate, "experiments/s", kind, what_e),
            "e2e": {"value": erate, "unit": "experiments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}), flush=True)
        return
    sharded = world > 1 and not args.replicas
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / rate,
        "higher_is_better": True, "scaling": "strong" if sharded else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic (seeded numpy, SURVEY 8 distributions)",
        "config": config_for(args, fit, world),
        "samples_binned_per_s": rate * fit.total_samples(),
        "cpu_baseline": cpu_baseline_object(rate, UNIT, kind, what),
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------
# own arm
# ---------------------------------------------------------------------------------------
def dist_setup():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def dist_on():
    import torch
    return torch.distributed.is_available() and torch.distributed.is_initialized()


def barrier():
    import torch
    torch.cuda.synchronize()
    if dist_on():
        torch.distributed.barrier()
    torch.cuda.synchronize()


def allreduce_max(x):
    import torch
    if not dist_on():
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def gather_objects(obj):
    import torch
    if not dist_on():
        return [obj]
    box = [None] * torch.distributed.get_world_size()
    torch.distributed.all_gather_object(box, obj)
    return box


def auto_repeats(args, K):
    if args.repeats > 0:
        return args.repeats
    return int(min(50, max(3, -(-600 // max(1, K)))))


def measure_chain(fit, fit_local, hists, chain, args, rank, comm, flush, do_e2e=True, repeats=None,
                  e2e_repeats=3):
    """Measures one built chain.  `comm` is None, "p2p" or a sharding.Comm.  Returns raw numbers."""
    import torch
    from sxmc_b200 import sharding
    K, Wm = args.steps, args.warmup
    R = repeats or auto_repeats(args, K)
    flush_bytes = L2_FLUSH_BYTES if flush else 0
    # independent chain per rank (replicas) / the same chain on every rank (sample sharding)
    chain.set_seed(20261104, 0 if comm is not None else rank)
    out = {"repeats": R}
    total = 2 * Wm + 2 * R * K + 64
    chain.begin(fit.data, total, 0.1)    # re-tunes at 10 % and 20 % of all steps, as in a real fit

    # (1) value: R blocks of K steps, every kernel of every step between CUDA events
    chain.advance_timed(Wm, flush_bytes)
    barrier()
    l0 = chain.launches()
    clk0 = sharding.p2p_clocks(chain) if comm == "p2p" else None
    blocks, bins, nlls, comms = [], [], [], []
    for _ in range(R):
        step_ms, bin_ms, nll_ms = chain.advance_timed(K, flush_bytes)
        blocks.append(step_ms); bins.append(bin_ms); nlls.append(nll_ms); comms.append(chain.last_comm_ms())
    out["gpu_launches"] = (chain.launches() - l0) // R
    if comm == "p2p":
        clk1 = sharding.p2p_clocks(chain)
        out["p2p_us"] = [round((b - a) / (R * K) * 1e-3, 3) for a, b in zip(clk0, clk1)]
    barrier()
    out.update(block_ms=blocks, bin_ms=float(np.median(bins)), nll_ms=float(np.median(nlls)),
               comm_ms=float(np.median(comms)) if comm is not None else None,
               step_ms=float(np.median(blocks)))

    # (2) the production form: sxmc_chain_advance picks the kernel (monotone / persistent / graph)
    chain.advance(Wm)
    barrier()
    warm = []
    clk0 = sharding.p2p_clocks(chain) if comm == "p2p" else None
    for _ in range(R):
        ms0 = chain.loop_ms()
        chain.advance(K)
        warm.append(chain.loop_ms() - ms0)
    out["warm_ms"] = float(np.median(warm))
    out["warm_blocks"] = warm
    # what `value` is: the production form when a step streams far more than the L2 holds (no flush
    # to apply); the flushed, kernel-by-kernel pass for the small fits
    out["value_ms"] = out["step_ms"] if flush else out["warm_ms"]
    out["value_blocks"] = blocks if flush else warm
    out["value_from"] = "flushed" if flush else "production"
    if comm == "p2p":
        clk1 = sharding.p2p_clocks(chain)
        out["p2p_us_warm"] = [round((b - a) / (R * K) * 1e-3, 3) for a, b in zip(clk0, clk1)]
    barrier()
    ls = chain.end()
    out["accepted"] = chain.accepted()
    out["steps_run"] = 2 * Wm + 2 * R * K
    out["rows"] = int(ls.get_samples().shape[0])
    out["path"] = {1: "graph", 2: "persistent", 4: "monotone"}.get(chain.last_path(), chain.last_path())
    out["bin_plan"] = dict(chain.bin_plan(), kernel={0: "interpreter", 1: "run-time dispatch", 2: "program of the build",
                                                      3: "NVRTC-compiled program"}[chain.bin_kernel_kind()])

    # (3) e2e through the public call with host buffers
    if do_e2e:
        chain(fit.data, max(32, Wm), 0.1)          # warm-up of the whole call
        ts = []
        for _ in range(e2e_repeats):
            barrier()
            t0 = time.perf_counter()
            ls = chain(fit.data, K, 0.1)
            rows = ls.get_samples()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        out["e2e_s"] = float(np.median(ts))
        out["e2e_all_s"] = ts
        out["e2e_h2d"] = fit.data.nbytes / K
        out["e2e_d2h"] = (rows.nbytes + 4096) / K     # rows + the chain-state readback
        barrier()
    return out


def build_chain(fit, args, rank, world, local, exchange):
    """(fit_local, hists, chain, comm) for this rank."""
    import torch
    from sxmc_b200 import mcmc, sharding
    t0 = time.perf_counter()
    comm = None
    fit_local = fit
    if exchange is not None and world > 1:
        fit_local = sharding.shard_fit_samples(fit, rank, world)
    if exchange == "p2p":                # handles copy the knobs when they are created
        from sxmc_b200 import pdfz
        pdfz.set_bin_tuning(p2p_form=args.p2p_form)
    hists, chain = mcmc.build_fit(fit_local)
    if exchange is not None and world > 1:
        comm = attach_exchange(chain, exchange, args, rank, world, local)
    for h in hists:
        h.Prepare()                      # sorted, tiled SoA layout (first-call Optimize analogue)
    torch.cuda.synchronize()
    return fit_local, hists, chain, comm, time.perf_counter() - t0


def attach_exchange(chain, exchange, args, rank, world, local):
    from sxmc_b200 import pdfz, sharding
    if exchange == "p2p":
        sharding.connect_p2p(chain, rank, world)
        return "p2p"
    comm = sharding.Comm(device=local)
    chain.set_comm(comm.handle)
    return comm


def close_chain(chain, comm):
    import torch
    chain.close()
    if comm is not None:
        torch.cuda.synchronize()
        if dist_on():
            torch.distributed.barrier()
        if comm != "p2p":
            comm.close()


def roofline_object(fit, r, K, peak, peak_src, world_share, traffic):
    alg_bytes = algorithmic_bytes_per_step(fit) // world_share
    bin_avg_ms = r["bin_ms"] / K
    achieved = alg_bytes / (bin_avg_ms * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": "bin_fast_kernel (transform + binning, all signals in one launch)",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
            "avg_launch_ms": bin_avg_ms, "share_of_step": r["bin_ms"] / r["step_ms"]}


def rank_spread(blocks, K):
    b = np.asarray(blocks) / K
    return [round(float(b.min()), 5), round(float(np.median(b)), 5), round(float(b.max()), 5)]


def ensemble_config(args, world):
    return {"workload": WORKLOAD_TEXT["E"], "experiments": args.experiments * world,
            "steps_per_experiment": args.esteps,
            "parallelism": "experiment i on GPU i mod N, Philox stream i; one gather of the limits",
            "l2": "not applicable: the samples are resident in shared memory for the whole launch"}


def measure_ensemble(fit, nexperiments, esteps, rank, world, seed=20261104):
    """Config E on this rank: its share of `nexperiments * world` experiments through
    sxmc_b200.ensemble.ensemble() -- fake data drawn on the device, chains advanced 148 at a time by
    the batch kernel, contour limit per experiment.  Wall clock of the whole call (host logic,
    data generation, launches, limit extraction)."""
    import torch
    from sxmc_b200 import ensemble
    ensemble.ensemble(fit, 4 * world, nsteps=200, burnin_fraction=0.1, rank=rank, world=world)   # warm-up
    best = None
    walls = []
    for _ in range(3):                    # the call is ~0.5 s of which half is host work: median of three
        barrier()
        t0 = time.perf_counter()
        r = ensemble.ensemble(fit, nexperiments * world, "s0", 0.9, seed=seed, nsteps=esteps,
                              burnin_fraction=0.1, rank=rank, world=world)
        torch.cuda.synchronize()
        r["wall_s"] = time.perf_counter() - t0
        walls.append(r["wall_s"])
        best = r
    med = float(np.median(walls))
    best["wall_s"] = med
    best["walls_s"] = walls
    return best


def ensemble_line(args, rank, world, sampler, fit=None):
    from sxmc_b200 import sharding
    fit = fit or make_fit("E", args.scale)
    r = measure_ensemble(fit, args.experiments, args.esteps, rank, world)
    wall = allreduce_max(r["wall_s"])
    loop_ms = allreduce_max(r["loop_ms"])
    limits = sharding.gather_limits(r["experiments"], r["limits"], args.experiments * world)
    total = args.experiments * world
    return {
        "metric": "ensemble_experiments_per_sec", "value": total / wall, "unit": "experiments/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * wall / (total * args.esteps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic (seeded numpy, SURVEY 8 distributions; "
        "Poisson fake data drawn on the device)",
        "config": ensemble_config(args, world),
        "chain_steps_per_s": total * args.esteps / wall,
        "us_per_chain_step_device": 1e3 * loop_ms / (args.experiments * args.esteps),
        "time_to_4096_experiments_s": 4096.0 / (total / wall),
        "median_upper_limit": sharding.median(limits), "mode": r["mode"],
        "wall_s_of_3_calls": r.get("walls_s"), "host_phases_s_last_call": r.get("phases_s"),
        "gpu_launches": r["launches"],
        "e2e": {"value": total / wall, "unit": "experiments/s", "h2d_bytes_per_step": 0.0,
                "d2h_bytes_per_step": 64.0 / args.esteps,
                "what": "ensemble.ensemble(): the value above IS the end-to-end wall clock (fake data are "
                        "generated on the device; per experiment the limit comes back)"},
    }


def init_device(world, local):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
    from sxmc_b200 import _lib
    _lib.check(_lib.lib().sxmc_init(local))


def run_ensemble_arm(args, rank, world, local):
    import torch
    init_device(world, local)
    sampler = ClockSampler(local)
    sampler.start()
    line = ensemble_line(args, rank, world, sampler)
    sampler.stop()
    line["clocks"] = sampler.summary()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist_on():
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def extra_workload(name, args, local, peak, peak_src, steps, warmup, repeats):
    """One of the other BASELINE workloads on this GPU, as a nested object."""
    import copy
    a2 = copy.copy(args)
    a2.steps, a2.warmup, a2.workload = steps, warmup, name
    fit = make_fit(name, 1.0)
    flush = wants_flush(a2, fit)
    _, hists, chain, comm, build_s = build_chain(fit, a2, 0, 1, local, None)
    r = measure_chain(fit, fit, hists, chain, a2, 0, None, flush, do_e2e=True, repeats=repeats, e2e_repeats=3)
    close_chain(chain, None)
    for h in hists:
        h.close()
    K = a2.steps
    out = {"workload": WORKLOAD_TEXT[name], "steps": K, "repeats": r["repeats"],
           "steps_per_s": K / (r["value_ms"] * 1e-3), "value_from": r["value_from"],
           "samples_binned_per_s": K / (r["value_ms"] * 1e-3) * fit.total_samples(),
           "l2": config_for(a2, fit, 1)["l2"],
           "roofline": roofline_object(fit, r, K, peak, peak_src, 1, ncu_traffic(name)),
           "kernels_ms_per_step": {"bin": r["bin_ms"] / K, "lookup_nll_metropolis": r["nll_ms"] / K},
           "chain_back_to_back": {"value": K / (r["warm_ms"] * 1e-3), "unit": UNIT, "path": r["path"]},
           "e2e": {"value": K / r["e2e_s"], "unit": UNIT}, "build_s": build_s}
    if r["path"] == "monotone":
        out["chain_back_to_back"]["note"] = (
            "monotone edge-search kernel: ~1000 bin-edge searches per step instead of a pass over the "
            "samples; 'samples binned/s' derived from it would be an EQUIVALENT rate, not bytes moved")
    return out


def run_own_arm(args, rank, world, local):
    import torch
    init_device(world, local)
    peak, peak_src = measured_peak()
    sampler = ClockSampler(local)
    K = args.steps
    sharded = world > 1 and not args.replicas
    fit = make_fit(args.workload, args.scale)
    flush = wants_flush(args, fit)
    exchange = args.exchange if sharded else None
    fit_local, hists, chain, comm, build_s = build_chain(fit, args, rank, world, local, exchange)
    sampler.start()
    r = measure_chain(fit, fit_local, hists, chain, args, rank, comm, flush)
    sampler.stop()

    spreads = gather_objects(rank_spread(r["value_blocks"], K))
    p2p_phases = None
    if "p2p_us" in r:
        box = gather_objects((r["p2p_us"], r.get("p2p_us_warm")))
        p2p_phases = {"what": "per rank, device globaltimer, us per step: [wait for every peer's partial histogram, "
                              "sum (fused form: the rank-sharded lookup, until the rank's last block is done), close "
                              "(second flag round), then block 0 of the fused kernel: peers' counts gathered + constants ready, "
                              "bins evaluated, block sum, partial sums stored + fenced + ticket]",
                      "timed": [b[0] for b in box], "back_to_back": [b[1] for b in box]}
    step_ms = allreduce_max(r["value_ms"])
    split_ms = allreduce_max(r["step_ms"])
    warm_ms = allreduce_max(r["warm_ms"])
    e2e_s = allreduce_max(r["e2e_s"])
    nchains = 1 if (sharded or world == 1) else world
    value = nchains * K / (step_ms * 1e-3)
    share = world if sharded else 1
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": step_ms / K, "higher_is_better": True,
        "scaling": "strong" if sharded else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (seeded numpy, SURVEY 8 distributions)",
        "config": config_for(args, fit, world),
        "repeats": r["repeats"],
        "timing": {"what": ("value = K / (max over ranks of the median K-step block); a block = K steps issued back to "
                            "back the way a fit runs them (sxmc_chain_advance: binning and lookup kernels chained by "
                            "programmatic dependent launch), one CUDA-event interval on the chain's stream around the "
                            "block; kernels_ms_per_step / roofline come from a second pass of the same blocks with an "
                            "event around every kernel (which serialises the two kernels: its step is "
                            "split_pass_ms_per_step)") if r["value_from"] == "production" else
                           ("value = K / (max over ranks of the median K-step block); a block = sum of its steps' "
                            "CUDA-event intervals on the chain's stream, L2 flushed before every step"),
                   "split_pass_ms_per_step": split_ms / K,
                   "ms_per_step_by_rank_min_median_max_over_blocks": spreads},
        "samples_binned_per_s": value * fit.total_samples(),
        "allreduce_ms_per_step": (r["comm_ms"] / K) if r["comm_ms"] is not None else None,
        "p2p_kernel_us_per_step": p2p_phases,
        "e2e": {"value": nchains * K / e2e_s, "unit": UNIT,
                "h2d_bytes_per_step": r["e2e_h2d"], "d2h_bytes_per_step": r["e2e_d2h"],
                "what": "MCMC.__call__(host data, K steps, burn-in 0.1) -> host chain rows: "
                        "sxmc_chain_run + sxmc_chain_read; wall clock, median of %d calls" % len(r["e2e_all_s"]),
                "calls_s": r["e2e_all_s"]},
        "gpu_launches": r["gpu_launches"],
        "roofline": roofline_object(fit, r, K, peak, peak_src, share,
                                    ncu_traffic(args.workload) if (args.scale == 1.0 and share == 1) else None),
        "kernels_ms_per_step": {"bin": r["bin_ms"] / K, "lookup_nll_metropolis": r["nll_ms"] / K},
        "bin_plan": r["bin_plan"],
        "chain_back_to_back": {"value": nchains * K / (warm_ms * 1e-3), "unit": UNIT,
                               "ms_per_step": warm_ms / K, "path": r["path"],
                               "note": "production form: chain.advance(K), steps back to back (monotone / "
                                       "persistent kernel for small fits, CUDA graph otherwise)"},
        "accept_rate": r["accepted"] / float(r["steps_run"]),
        "build_s": build_s,
    }

    # N > 1: the other exchange on the same sharded samples, nested
    # (p2p_allreduce: the peer-memory all-reduce kernel + the replicated lookup, i.e. the step without
    # the fusion; nccl: ncclAllReduce in its place)
    if sharded and not args.no_extra:
        from sxmc_b200 import mcmc, pdfz
        others = ["nccl", "p2p_allreduce"] if (args.exchange == "p2p" and args.p2p_form in (0, 3)) else \
                 (["nccl"] if args.exchange == "p2p" else ["p2p"])
        for other in others:
            try:
                close_chain(chain, comm)
                pdfz.set_bin_tuning(p2p_form=(2 if world > 2 else 1) if other == "p2p_allreduce" else args.p2p_form)
                _, chain = mcmc.build_fit(fit_local, hists=hists)
                pdfz.set_bin_tuning()
                comm = attach_exchange(chain, "p2p" if other == "p2p_allreduce" else other, args, rank, world, local)
                r2 = measure_chain(fit, fit_local, hists, chain, args, rank, comm, flush, do_e2e=False,
                                   repeats=max(3, r["repeats"] // 3))
                s2 = allreduce_max(r2["value_ms"])
                line["exchange_" + other] = {
                    "steps_per_s": K / (s2 * 1e-3), "ms_per_step": s2 / K,
                    "allreduce_ms_per_step": r2["comm_ms"] / K,
                    "kernels_ms_per_step": {"bin": r2["bin_ms"] / K, "lookup_nll_metropolis": r2["nll_ms"] / K},
                    "chain_back_to_back_steps_per_s": K / (allreduce_max(r2["warm_ms"]) * 1e-3)}
            except Exception as e:
                line["exchange_" + other] = {"error": repr(e)}
    close_chain(chain, comm)
    for h in hists:
        h.close()
    del hists, chain, fit_local
    return line, fit, sampler, peak, peak_src


def main():
    args = parse_args()
    rank, world, local = dist_setup()
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if args.workload == "E":
        run_ensemble_arm(args, rank, world, local)
        return
    line, fit, sampler, peak, peak_src = run_own_arm(args, rank, world, local)
    import torch
    extras = not args.no_extra and args.scale == 1.0 and args.workload == "D"

    # nested: config E (ensemble) at this N -- 148 experiments per GPU through the batch kernel
    if extras:
        try:
            a2 = parse_args([])
            a2.experiments, a2.esteps, a2.steps, a2.warmup = 148, 20000, args.steps, args.warmup
            el = ensemble_line(a2, rank, world, sampler)
            line["ensemble_E"] = {k: el[k] for k in ("value", "unit", "config", "chain_steps_per_s",
                                                     "us_per_chain_step_device", "time_to_4096_experiments_s",
                                                     "median_upper_limit", "mode", "gpu_launches",
                                                     "wall_s_of_3_calls", "host_phases_s_last_call")}
        except Exception as e:      # the headline must still print
            line["ensemble_E"] = {"error": repr(e)}

    # nested, N = 1 only: the other BASELINE workloads and the reference's own benchmark shape
    if extras and world == 1:
        del fit
        fit = None
        for name, (st, wu, rp) in (("C", (20, 5, 10)), ("B", (200, 20, 5))):
            try:
                line["workload_" + name] = extra_workload(name, args, local, peak, peak_src, st, wu, rp)
            except Exception as e:
                line["workload_" + name] = {"error": repr(e)}
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
        try:
            import bench_nll_sweep
            line["nll_sweep"] = bench_nll_sweep.sweep(peak, quick=True)
        except Exception as e:
            line["nll_sweep"] = {"error": repr(e)}

    line["clocks"] = sampler.summary()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        if fit is None:
            fit = make_fit(args.workload, args.scale)
        rate, kind, n, what = cpu_steps_per_s(fit, args.cpu_budget_s)
        line["cpu_baseline"] = cpu_baseline_object(rate, UNIT, kind, what)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist_on():
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
