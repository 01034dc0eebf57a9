"""Worker for the multi-rank tests (launched with torch.distributed.run, one process per rank).

mode "gloo-oracle": CPU, backend gloo.  Each rank evaluates the oracle's normal-equation blocks on ITS frame shard;
    the camera blocks, camera gradient and cost are all-reduced (sum) and must equal the blocks of the whole rig:
    the additivity the sharded solve relies on (calibrator.cpp:466-506: residual blocks couple one frame's pose
    with the shared camera parameters only).
mode "nccl-solve <case>": GPU, one rank per GPU.  Each rank feeds its shard to libmcc through the C ABI, the library
    sums the reduced camera system over the ranks (peer windows or NCCL); cost and camera parameters must be
    bit-identical on every rank and, with the frame poses gathered, match the oracle's solve of the WHOLE rig at the
    north-star tolerances.  Cases:
      small        3 cameras, 40 frames (all steps accepted)
      rejected     a rig started far from the solution and solved with min_relative_decrease = 0.7 (tests/fixtures_util.py):
                   five REJECTED steps, which take the sharded solve through the non-speculative branch (small
                   all-reduce after the cost-only pass) and through the stop / resynchronisation of the queued iterations
      c4           config c4's rig, 400 frames (the bench workload at a size the oracle solves in seconds)
      empty        one frame over all ranks (intrinsics fixed): every rank but one has an EMPTY shard and must still take part
"""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)


def make_case(case):
    from multicam_calibration_b200 import synth
    if case == "small":
        return synth.make_synthetic_rig(3, 40, 31, detect_prob=0.8), None
    if case == "rejected":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from fixtures_util import rejected_steps_rig
        rig, mrd = rejected_steps_rig()
        return rig, mrd
    if case == "c4":
        return synth.config_c4(nframes=400), None
    if case == "empty":
        rig = synth.make_synthetic_rig(2, 1, 77, detect_prob=1.0)
        rig.fix_all_intrinsics = True       # one frame: pose + extrinsics only, so that the problem is well posed
        return rig, None
    raise SystemExit("unknown case " + case)


def main():
    mode = sys.argv[1]
    import torch
    import torch.distributed as dist
    from multicam_calibration_b200 import synth
    from multicam_calibration_b200.rig import shard_bounds
    import oracle
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if mode == "gloo-oracle":
        rig = synth.make_synthetic_rig(3, 40, 31, detect_prob=0.8)
        b = shard_bounds(rig.npts, world)
        shard = rig.slice_frames(int(b[rank]), int(b[rank + 1]))
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
        case = sys.argv[2]
        from multicam_calibration_b200 import capi
        local = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        rig, mrd = make_case(case)
        b = shard_bounds(rig.npts, world)
        shard = rig.slice_frames(int(b[rank]), int(b[rank + 1]))
        if case == "empty":
            assert sum(int(b[r + 1] - b[r]) == 0 for r in range(world)) == world - 1
        cal = capi.Calibrator(local)
        cal.setCameras(shard.cameras)
        cal.setFixAllIntrinsics(rig.fix_all_intrinsics)
        if shard.num_frames:
            cal.addFrames(shard)
        obj = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        cal.commInit(world, rank, obj[0])
        opt = capi.default_options()
        if mrd is not None:
            opt.min_relative_decrease = mrd
        s = cal.runCalibration(opt)
        p = cal.getParams()
        Cc = rig.num_cam_params
        # residual-only pass over the shards: the cost all ranks agree on is the cost of the returned parameters
        c_eval = cal.evalCost(use_loss=True)
        mine = (float(s.final_cost).hex(), p[:Cc].tobytes(), s.num_iterations, s.num_successful, s.num_unsuccessful, s.why, float(c_eval).hex())
        allr = [None] * world
        dist.all_gather_object(allr, mine)
        poses = [None] * world
        dist.all_gather_object(poses, p[Cc:])
        transport = cal.commTransport()
        if rank == 0:
            for r in range(1, world):
                assert allr[r] == allr[0], "rank %d disagrees with rank 0 on cost / camera parameters / iteration counts" % r
            want_transport = 2 if os.environ.get("MCC_P2P") == "0" else 1
            assert transport == want_transport, (transport, want_transport)
            assert abs(c_eval - s.final_cost) <= 1e-12 * s.final_cost
            o = oracle.Oracle(rig)
            oopt = oracle.default_options()
            if mrd is not None:
                oopt.min_relative_decrease = mrd
            so = o.solve(oopt)
            po = o.get_params()
            pg = np.concatenate([p[:Cc]] + poses)
            rel = np.max(np.abs(pg - po) / np.maximum(np.abs(po), 1e-3))
            assert abs(s.final_cost - so.final_cost) <= 1e-9 * so.final_cost, (s.final_cost, so.final_cost)
            assert rel <= 1e-7, rel
            assert s.num_iterations == so.num_iterations, (s.num_iterations, so.num_iterations)
            assert (s.num_successful, s.num_unsuccessful, s.why) == (so.num_successful, so.num_unsuccessful, so.why)
            if case == "rejected":
                assert s.num_unsuccessful >= 3, "this fixture is here for its rejected steps"
            print("MULTI_RANK_OK nccl-solve", case, "world", world, "transport", transport, "cost", s.final_cost, "rel", rel,
                  "iterations", s.num_iterations, "rejected", s.num_unsuccessful)
        dist.barrier()
        cal.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    try:
        main()
    except BaseException:
        # leave at once: with an exception pending, interpreter shutdown would wait in the communicators' destructors for
        # the other ranks, which sit in a barrier until the process-group watchdog fires ten minutes later
        import traceback
        traceback.print_exc()
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(1)
