"""Compare the files tests/_dist_worker.py nccl_sample_sharded wrote for any world size:
every rank's NCCL / one-shot / two-shot chain must equal the unsharded chain bit for bit."""
import glob
import sys

import numpy as np

files = sorted(glob.glob(sys.argv[1] + "/sharded_rank*.npz"))
z0 = np.load(files[0])
for f in files:
    z = np.load(f)
    for pre in ("", "p2p_", "p2p2_"):
        for k in ("bins", "norms", "rows"):
            assert np.array_equal(z[pre + k], z0["single_" + k]), (f, pre + k)
        assert z[pre + "accepted"] == z0["single_accepted"], (f, pre)
    assert not z["p2p_aborted"] and not z["p2p2_aborted"]
print("ok: %d ranks, %d rows, all three exchanges equal the unsharded chain" % (len(files), z0["rows"].shape[0]))
