"""The N > 1 path: world_size-2 gloo test of the sharding logic on CPU, and (on a box with >= 2 GPUs) the sharded solve on
every GPU the box offers (2, 4 or 8 ranks) against the oracle."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

WORKER = os.path.join(os.path.dirname(__file__), "multi_rank_worker.py")


def _launch(args, nproc, port, extra_env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), WORKER] + list(args)
    # a rank that waits in a peer exchange gives up after 30 s (instead of the default 60) so that a broken job fails fast
    env = dict(os.environ, OMP_NUM_THREADS="4", MCC_PEER_TIMEOUT_S="30", **(extra_env or {}))
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "MULTI_RANK_OK " + " ".join(args) in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def test_sharded_blocks_sum_to_whole_gloo(built):
    _launch(["gloo-oracle"], 2, 29611)


def test_shard_bounds_cover_and_balance(built):
    """rig.shard_bounds: contiguous, covering, balanced by observation count, empty shards allowed"""
    import numpy as np
    from multicam_calibration_b200 import synth
    from multicam_calibration_b200.rig import shard_bounds
    rig = synth.config_c4(nframes=300)
    for world in (1, 2, 4, 8):
        b = shard_bounds(rig.npts, world)
        assert b[0] == 0 and b[-1] == rig.num_frames and np.all(np.diff(b) >= 0)
        per = [int(rig.npts[b[r]:b[r + 1]].sum()) for r in range(world)]
        assert sum(per) == rig.num_obs and max(per) - min(per) <= 2 * int(rig.npts.sum(axis=1).max())
    b = shard_bounds(synth.make_synthetic_rig(2, 1, 77, detect_prob=1.0).npts, 4)
    assert list(np.diff(b)).count(0) == 3


def _gpu_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.gpu
@pytest.mark.parametrize("p2p", ["1", "0"])
@pytest.mark.parametrize("case", ["small", "rejected", "c4", "empty"])
def test_sharded_solve(built, case, p2p):
    """all GPUs of the box (2, 4 or 8 ranks), frames sharded by observation count, against the oracle's solve of the
    whole rig: the per-iteration exchanges go through the peer-memory all-reduce (MCC_P2P=1, the default on one node) or
    through NCCL (MCC_P2P=0).  See tests/multi_rank_worker.py for the cases."""
    n = _gpu_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run under gpurun --gpus 2)")
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    out = _launch(["nccl-solve", case], n, 29620 + 2 * ["small", "rejected", "c4", "empty"].index(case) + int(p2p), {"MCC_P2P": p2p})
    print(out[-400:])
