"""The N > 1 path: world_size-2 gloo test of the sharding logic on CPU, and (on a box with >= 2 GPUs) the NCCL solve."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

WORKER = os.path.join(os.path.dirname(__file__), "multi_rank_worker.py")


def _launch(mode, nproc, port, extra_env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), WORKER, mode]
    env = dict(os.environ, OMP_NUM_THREADS="2", **(extra_env or {}))
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "MULTI_RANK_OK " + mode in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_sharded_blocks_sum_to_whole_gloo(built):
    _launch("gloo-oracle", 2, 29611)


@pytest.mark.gpu
@pytest.mark.parametrize("p2p", ["1", "0"])
def test_sharded_solve_nccl(built, p2p):
    """two ranks, each with half of the frames, against the single-rank solve: the per-iteration exchanges go through
    the peer-memory all-reduce (MCC_P2P=1, the default on one node) or through NCCL (MCC_P2P=0)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (run under gpurun --gpus 2)")
    _launch("nccl-solve", 2, 29613 + int(p2p), {"MCC_P2P": p2p})
