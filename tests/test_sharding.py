"""N>1 host logic: who owns which samples / experiments and how the results
come back.  CPU part: two `gloo` processes.  GPU part (`-m gpu`, needs >= 2 devices): a
sample-sharded chain over NCCL must produce the same histograms and the same chain rows as
the unsharded chain on one GPU."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "_dist_worker.py")


def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _torchrun(nproc, what, outdir, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), WORKER, what, str(outdir)]
    # a rank that waits for a peer gives up after 20 s (default 60): a protocol bug fails the test, not the box
    env = dict(os.environ, SXMC_P2P_TIMEOUT_S=os.environ.get("SXMC_P2P_TIMEOUT_S", "20"))
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_slice_bounds_cover_every_sample_once():
    from sxmc_b200 import sharding
    for n in (0, 1, 7, 1000, 3_333_333):
        for world in (1, 2, 3, 8):
            cuts = [sharding.slice_bounds(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[r][1] == cuts[r + 1][0] for r in range(world - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.slice_bounds(10, 2, 2)


def test_shard_fit_samples_partitions_rows_and_keeps_n_mc():
    from sxmc_b200 import sharding, workloads as W
    fit = W.config_C(0.0005)
    parts = [sharding.shard_fit_samples(fit, r, 3) for r in range(3)]
    for j, sig in enumerate(fit.signals):
        glued = np.concatenate([p.signals[j].samples for p in parts], axis=0)
        assert np.array_equal(glued, sig.samples)
        assert all(p.signals[j].n_mc == sig.n_mc for p in parts)
    assert all(np.array_equal(p.data, fit.data) for p in parts)


def test_experiment_round_robin_and_median():
    from sxmc_b200 import sharding
    for n in (0, 1, 5, 4096):
        for world in (1, 2, 8):
            owned = [sharding.experiments_for_rank(n, r, world) for r in range(world)]
            assert sorted(i for o in owned for i in o) == list(range(n))
            assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
    assert sharding.median([3.0, 1.0, 2.0]) == 2.0
    assert sharding.median([4.0, 1.0, 2.0, 3.0]) == 2.5        # src/utils.h:77-89
    assert sharding.median([]) == -1.0                          # src/sxmc.cpp:226


def test_gloo_world2_gathers_limits_in_experiment_order(tmp_path):
    """Two CPU processes each "run" their experiments and gather; both must end with the
    same full vector, equal to the single-process result."""
    _torchrun(2, "gloo_limits", tmp_path)
    a = np.load(tmp_path / "limits_rank0.npy")
    b = np.load(tmp_path / "limits_rank1.npy")
    want = np.asarray([np.float32(0.5 * i * i - 3 * i) for i in range(11)], dtype=np.float32)
    assert np.array_equal(a, want) and np.array_equal(b, want)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_sample_sharded_chain_equals_single_gpu_chain(tmp_path, world):
    """NCCL and the three peer-memory forms (the exchange fused into a rank-sharded lookup, the one-shot
    and the reduce-scatter + all-gather all-reduce kernels) at 2, 4 and 8 ranks: bins, norms, rows and
    the accepted count of the sharded chain equal the unsharded chain bit for bit on EVERY rank.  Each
    world size runs where the box has that many GPUs."""
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs (this box has %d)" % (world, torch.cuda.device_count() if torch.cuda.is_available() else 0))
    _torchrun(world, "nccl_sample_sharded", tmp_path, timeout=900)
    zs = [np.load(tmp_path / ("sharded_rank%d.npz" % r)) for r in range(world)]
    z0 = zs[0]
    assert z0["rows"].shape[0] > 0
    for z in zs:
        for pre in ("", "p2p_", "p2p1_", "p2p2_"):
            for k in ("bins", "norms", "rows"):
                assert np.array_equal(z[pre + k], z0["single_" + k]), (pre + k)
            assert z[pre + "accepted"] == z0["single_accepted"]
        for pre in ("p2p_", "p2p1_", "p2p2_"):
            assert not z[pre + "aborted"] and z[pre + "exchanges"] == 301     # initial evaluation + 300 steps
        for k in ("bins", "norms", "rows"):
            assert np.array_equal(z["D_p2p_" + k], z0["D_single_" + k]), ("D_p2p_" + k)
        assert z["D_p2p_accepted"] == z0["D_single_accepted"]
        assert not z["D_p2p_aborted"] and z["D_p2p_exchanges"] == 61
