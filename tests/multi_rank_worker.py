"""Worker for the multi-rank tests (launched with torch.distributed.run, one process per rank).

mode "gloo-oracle": CPU, backend gloo.  Each rank evaluates the oracle's normal-equation blocks on ITS frame shard;
    the camera blocks, camera gradient and cost are all-reduced (sum) and must equal the blocks of the whole rig:
    the additivity the sharded solve relies on (calibrator.cpp:466-506: residual blocks couple one frame's pose
    with the shared camera parameters only).
mode "nccl-solve": GPU, one rank per GPU.  Each rank feeds its shard to libmcc through the C ABI, the library
    all-reduces the reduced camera system over NCCL; camera parameters must agree on every rank and, with the
    frame poses gathered, match the oracle's solve of the whole rig.
"""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)


def main():
    mode = sys.argv[1]
    import torch
    import torch.distributed as dist
    from multicam_calibration_b200 import synth
    from multicam_calibration_b200.rig import shard_bounds
    import oracle
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    rig = synth.make_synthetic_rig(3, 40, 31, detect_prob=0.8)
    b = shard_bounds(rig.npts, world)
    shard = rig.slice_frames(int(b[rank]), int(b[rank + 1]))
    if mode == "gloo-oracle":
        dist.init_process_group("gloo")
        nb = oracle.Oracle(shard).normal_blocks(True, num_threads=2)
        C = nb["Hcc"].shape[0]
        buf = torch.tensor(np.concatenate([nb["Hcc"].reshape(-1), nb["gc"], [nb["cost"]]]))
        dist.all_reduce(buf)
        if rank == 0:
            full = oracle.Oracle(rig).normal_blocks(True, num_threads=2)
            got = buf.numpy()
            want = np.concatenate([full["Hcc"].reshape(-1), full["gc"], [full["cost"]]])
            scale = np.abs(want).max()
            assert np.abs(got[:C * C] - want[:C * C]).max() <= 1e-12 * np.abs(want[:C * C]).max()
            assert np.abs(got[C * C:-1] - want[C * C:-1]).max() <= 1e-12 * np.abs(want[C * C:-1]).max()
            assert abs(got[-1] - want[-1]) <= 1e-13 * want[-1]
            # per-frame blocks of a shard are the whole rig's blocks of those frames
            lo, hi = int(b[0]), int(b[1])
            np.testing.assert_allclose(nb["Hff"], full["Hff"][lo:hi], rtol=1e-13, atol=0)
            print("MULTI_RANK_OK gloo-oracle world", world, "scale", scale)
        dist.barrier()
        dist.destroy_process_group()
        return
    if mode == "nccl-solve":
        from multicam_calibration_b200 import capi
        local = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        cal = capi.calibrator_from_rig(shard, device=local)
        obj = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        cal.commInit(world, rank, obj[0])
        s = cal.runCalibration()
        p = cal.getParams()
        Cc = rig.num_cam_params
        cam = torch.tensor(p[:Cc], device="cuda")
        gathered = [torch.zeros_like(cam) for _ in range(world)]
        dist.all_gather(gathered, cam)
        poses = [None] * world
        dist.all_gather_object(poses, p[Cc:])
        if rank == 0:
            for g in gathered[1:]:
                assert torch.equal(g, gathered[0]), "camera parameters differ between ranks"
            o = oracle.Oracle(rig)
            so = o.solve()
            po = o.get_params()
            pg = np.concatenate([p[:Cc]] + poses)
            rel = np.max(np.abs(pg - po) / np.maximum(np.abs(po), 1e-3))
            assert abs(s.final_cost - so.final_cost) <= 1e-9 * so.final_cost, (s.final_cost, so.final_cost)
            assert rel <= 1e-7, rel
            assert s.num_iterations == so.num_iterations, (s.num_iterations, so.num_iterations)
            print("MULTI_RANK_OK nccl-solve world", world, "cost", s.final_cost, "rel", rel)
        dist.barrier()
        cal.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
