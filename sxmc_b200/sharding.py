"""Multi-GPU host logic: one process per GPU, `torch.distributed` for the
plumbing.  Two independent ways the path shards:

* ensemble sharding -- experiment ``i`` of ``nexperiments`` (reference loop
  `src/sxmc.cpp:62-134`) runs on rank ``i mod world``; no data-path collective, one final
  gather of the per-experiment limits;
* sample sharding -- every signal's Monte-Carlo sample array is cut into ``world``
  contiguous slices; per step each rank bins its slice and the concatenated
  ``uint32 [nsignals][total_nbins] + norms`` buffer is summed over ranks on the chain's
  stream (``sxmc_chain_set_comm``).  Integer sums: histograms are bit-identical for any
  world size, and every rank takes the same Metropolis decision from the same Philox
  counters.

Nothing here computes on samples; the functions only decide who owns what and move small
host objects between ranks (works with the ``gloo`` backend on CPU and ``nccl`` on GPUs).
"""
import copy
import ctypes as C

import numpy as np

from ._lib import c_vp, check, lib


def slice_bounds(n, rank, world):
    """Contiguous slice [lo, hi) of n items owned by `rank`; slices differ by at most 1."""
    n, rank, world = int(n), int(rank), int(world)
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return (n * rank) // world, (n * (rank + 1)) // world


def shard_fit_samples(fit, rank, world):
    """FitSpec whose signals hold only this rank's slice of the samples.  ``n_mc`` stays the
    GLOBAL sample count (the efficiency `norm / n_mc`, `src/nll_kernels.cpp:105`, is taken
    after the histograms are summed over ranks); data events are replicated."""
    out = copy.copy(fit)
    out.signals = []
    for sig in fit.signals:
        lo, hi = slice_bounds(sig.samples.shape[0], rank, world)
        s = copy.copy(sig)
        s.samples = np.ascontiguousarray(sig.samples[lo:hi])
        s.n_mc = int(sig.n_mc)
        out.signals.append(s)
    return out


def experiments_for_rank(nexperiments, rank, world):
    """Experiment indices rank `rank` runs: i mod world == rank (round robin keeps the
    ranks within one experiment of each other for any nexperiments)."""
    return list(range(int(rank), int(nexperiments), int(world)))


def gather_limits(local_indices, local_limits, nexperiments):
    """All ranks' (experiment index, limit) pairs -> one float32 array in experiment order
    on every rank (NaN where an experiment produced no limit).  The reference appends to one
    vector in loop order (`src/sxmc.cpp:126`) and takes its median (`:218-226`)."""
    import torch.distributed as dist
    out = np.full(int(nexperiments), np.nan, dtype=np.float32)
    pairs = [(int(i), float(v)) for i, v in zip(local_indices, local_limits)]
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        gathered = [None] * dist.get_world_size()
        dist.all_gather_object(gathered, pairs)
    else:
        gathered = [pairs]
    for part in gathered:
        for i, v in part:
            out[i] = v
    return out


def median(values):
    """`median()` of `src/utils.cpp` semantics: middle of the sorted list (mean of the two
    middle elements for even counts)."""
    v = np.sort(np.asarray(values, dtype=np.float32))
    v = v[~np.isnan(v)]
    if v.size == 0:
        return -1.0
    m = v.size // 2
    return float(v[m]) if v.size % 2 else float((v[m - 1] + v[m]) / 2)


class Comm:
    """NCCL communicator of libsxmc_b200 (`sxmc_comm_*`), one rank per process.  The 128-byte
    unique id is made on rank 0 and broadcast through the already initialised
    torch.distributed group (any backend)."""

    def __init__(self, rank=None, world=None, device=None):
        import torch
        import torch.distributed as dist
        self.rank = dist.get_rank() if rank is None else int(rank)
        self.world = dist.get_world_size() if world is None else int(world)
        uid = (C.c_ubyte * 128)()
        if self.rank == 0:
            check(lib().sxmc_comm_unique_id(C.cast(uid, c_vp)))
        box = [bytes(uid)]
        dist.broadcast_object_list(box, src=0,
                                   device=torch.device("cuda", device) if device is not None else None)
        buf = (C.c_ubyte * 128).from_buffer_copy(box[0])
        self._c = c_vp()
        check(lib().sxmc_comm_create(C.cast(buf, c_vp), self.world, self.rank, C.byref(self._c)))

    @property
    def handle(self):
        return self._c

    def allreduce_u32(self, dev_ptr, count):
        check(lib().sxmc_comm_allreduce_u32(self._c, dev_ptr, int(count)))

    def close(self):
        if self._c:
            lib().sxmc_comm_destroy(self._c)
            self._c = c_vp()


def connect_p2p(chain, rank=None, world=None):
    """Sample sharding over NVLink peer memory (``sxmc_chain_p2p_*``): export this rank's
    histogram window, all-gather the 64-byte IPC handles through torch.distributed, open the
    peers'.  Call on every rank, after ``chain.set_data`` or before the first run."""
    import torch.distributed as dist
    rank = dist.get_rank() if rank is None else int(rank)
    world = dist.get_world_size() if world is None else int(world)
    h = (C.c_ubyte * 64)()
    check(lib().sxmc_chain_p2p_export(chain._c, C.cast(h, c_vp)))
    gathered = [None] * world
    dist.all_gather_object(gathered, bytes(h))
    blob = (C.c_ubyte * (64 * world)).from_buffer_copy(b"".join(gathered))
    check(lib().sxmc_chain_p2p_connect(chain._c, C.cast(blob, c_vp), world, rank))
    dist.barrier()


def p2p_status(chain):
    n, a = C.c_uint(0), C.c_int(0)
    check(lib().sxmc_chain_p2p_status(chain._c, C.byref(n), C.byref(a)))
    return n.value, bool(a.value)


def p2p_clocks(chain):
    """(wait, sum, close) nanoseconds the exchange has spent so far, summed over steps -- in the fused form
    "sum" is the rank-sharded lookup; then four more, block 0 of the fused kernel: peers' counts gathered and
    constants ready / bins evaluated / block sum / partial sums stored, fenced and ticket taken."""
    out = (C.c_double * 7)()
    check(lib().sxmc_chain_p2p_clocks(chain._c, out))
    return tuple(out)
