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
    from sxmc_b200 import _lib, mcmc, sharding, workloads as W
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    _lib.check(_lib.lib().sxmc_init(local))
    fit = W.config_C(0.002)
    nsteps = 300
    out = {}

    # the unsharded chain, on every rank (same seed => same chain)
    hists, chain = mcmc.build_fit(fit)
    chain.set_seed(99, 0)
    chain.set_mode(3)
    ls = chain(fit.data, nsteps, 0.1)
    out.update(single_rows=ls.get_samples().copy(), single_bins=chain.bins(),
               single_norms=chain.norms(), single_accepted=chain.accepted())
    chain.close()
    for h in hists:
        h.close()

    # sample-sharded: this rank's slice of every signal, histograms summed over NCCL
    comm = sharding.Comm(device=local)
    part = sharding.shard_fit_samples(fit, rank, world)
    hists, chain = mcmc.build_fit(part)
    chain.set_seed(99, 0)
    chain.set_comm(comm.handle)
    ls = chain(fit.data, nsteps, 0.1)
    out.update(rows=ls.get_samples().copy(), bins=chain.bins(), norms=chain.norms(),
               accepted=chain.accepted())
    chain.close()
    for h in hists:
        h.close()
    torch.cuda.synchronize()
    dist.barrier()
    comm.close()

    # the same over NVLink peer memory: one exchange kernel per step instead of ncclAllReduce
    hists, chain = mcmc.build_fit(part)
    chain.set_seed(99, 0)
    sharding.connect_p2p(chain, rank, world)
    ls = chain(fit.data, nsteps, 0.1)
    nex, aborted = sharding.p2p_status(chain)
    out.update(p2p_rows=ls.get_samples().copy(), p2p_bins=chain.bins(), p2p_norms=chain.norms(),
               p2p_accepted=chain.accepted(), p2p_exchanges=nex, p2p_aborted=aborted)
    torch.cuda.synchronize()
    dist.barrier()
    chain.close()
    for h in hists:
        h.close()

    # ... and its reduce-scatter + all-gather form (what more than two ranks use), forced here
    from sxmc_b200 import pdfz
    pdfz.set_bin_tuning(p2p_form=2)
    hists, chain = mcmc.build_fit(part)
    chain.set_seed(99, 0)
    sharding.connect_p2p(chain, rank, world)
    ls = chain(fit.data, nsteps, 0.1)
    nex, aborted = sharding.p2p_status(chain)
    out.update(p2p2_rows=ls.get_samples().copy(), p2p2_bins=chain.bins(), p2p2_norms=chain.norms(),
               p2p2_accepted=chain.accepted(), p2p2_exchanges=nex, p2p2_aborted=aborted)
    torch.cuda.synchronize()
    dist.barrier()
    chain.close()
    for h in hists:
        h.close()
    pdfz.set_bin_tuning(p2p_form=0)
    np.savez(os.path.join(outdir, "sharded_rank%d.npz" % rank), **out)
    dist.destroy_process_group()


if __name__ == "__main__":
    {"gloo_limits": gloo_limits, "nccl_sample_sharded": nccl_sample_sharded}[sys.argv[1]](sys.argv[2])
