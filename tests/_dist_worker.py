"""Worker launched by tests/test_sharding.py under torch.distributed.run (one process per
rank).  argv: <what> <outdir>."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def gloo_limits(outdir):
    import torch.distributed as dist
    from sxmc_b200 import sharding
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n = 11
    mine = sharding.experiments_for_rank(n, rank, world)
    limits = [np.float32(0.5 * i * i - 3 * i) for i in mine]      # stands in for a fit's limit
    full = sharding.gather_limits(mine, limits, n)
    np.save(os.path.join(outdir, "limits_rank%d.npy" % rank), full)
    dist.barrier()
    dist.destroy_process_group()


def nccl_sample_sharded(outdir):
    import torch
    import torch.distributed as dist
    from sxmc_b200 import _lib, mcmc, pdfz, sharding, workloads as W
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    _lib.check(_lib.lib().sxmc_init(local))
    out = {}

    def single(fit, nsteps, pre):
        """the unsharded chain, on every rank (same seed => same chain)"""
        hists, chain = mcmc.build_fit(fit)
        chain.set_seed(99, 0)
        chain.set_mode(3)
        ls = chain(fit.data, nsteps, 0.1)
        out.update({pre + "rows": ls.get_samples().copy(), pre + "bins": chain.bins(),
                    pre + "norms": chain.norms(), pre + "accepted": chain.accepted()})
        chain.close()
        for h in hists:
            h.close()

    def sharded(fit, nsteps, pre, p2p_form=None):
        """this rank's slice of every signal; histograms summed over NCCL (p2p_form None) or over
        NVLink peer memory: 0 automatic (the exchange fused into a rank-sharded lookup where the
        fit allows it), 1 one-shot all-reduce kernel, 2 reduce-scatter + all-gather kernel"""
        part = sharding.shard_fit_samples(fit, rank, world)
        comm = None
        if p2p_form is not None:
            pdfz.set_bin_tuning(p2p_form=p2p_form)
        hists, chain = mcmc.build_fit(part)
        pdfz.set_bin_tuning()
        chain.set_seed(99, 0)
        if p2p_form is None:
            comm = sharding.Comm(device=local)
            chain.set_comm(comm.handle)
        else:
            sharding.connect_p2p(chain, rank, world)
        ls = chain(fit.data, nsteps, 0.1)
        out.update({pre + "rows": ls.get_samples().copy(), pre + "bins": chain.bins(),
                    pre + "norms": chain.norms(), pre + "accepted": chain.accepted()})
        if p2p_form is not None:
            nex, aborted = sharding.p2p_status(chain)
            out.update({pre + "exchanges": nex, pre + "aborted": aborted})
        torch.cuda.synchronize()
        dist.barrier()
        chain.close()
        for h in hists:
            h.close()
        if comm is not None:
            comm.close()

    fit = W.config_C(0.002)                      # 5000 bins, 10 signals, one dataset: the fused form applies
    single(fit, 300, "single_")
    sharded(fit, 300, "")                        # NCCL
    sharded(fit, 300, "p2p_", 0)                 # fused exchange + rank-sharded lookup
    sharded(fit, 300, "p2p1_", 1)                # one-shot all-reduce kernel
    sharded(fit, 300, "p2p2_", 2)                # reduce-scatter + all-gather kernel
    # 200 000 bins, 30 signals: the fused lookup's logical grid is capped (two bins per thread)
    fit = W.config_D(0.002, stress=True)
    single(fit, 60, "D_single_")
    sharded(fit, 60, "D_p2p_", 0)
    np.savez(os.path.join(outdir, "sharded_rank%d.npz" % rank), **out)
    dist.destroy_process_group()


if __name__ == "__main__":
    {"gloo_limits": gloo_limits, "nccl_sample_sharded": nccl_sample_sharded}[sys.argv[1]](sys.argv[2])
