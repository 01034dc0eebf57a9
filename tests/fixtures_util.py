"""Test fixtures shared by the GPU parity tests and the multi-rank worker."""
import numpy as np


def rejected_steps_rig():
    """3 cameras, 40 frames, started far from the solution (poses off by ~0.3 rad / 10 cm, focal lengths by 20 %).  Solved with
    ``min_relative_decrease = 0.7`` its second step is not good enough, the trust region shrinks through five REJECTED steps and
    recovers: 21 iterations, 16 successful, 5 unsuccessful in the oracle.  Exercises the non-speculative branch of the solver
    (cost-only pass after a rejection, small all-reduce with several ranks) and the stop / resume of the queued iterations.
    Returns (rig, min_relative_decrease)."""
    from multicam_calibration_b200 import synth
    seed = 31
    rig = synth.make_synthetic_rig(3, 40, seed, detect_prob=0.8)
    rng = np.random.default_rng(seed + 1000)
    for f in range(rig.num_frames):
        dR = synth.rodrigues(rng.normal(0, 0.3, 3))
        rig.poses[f, :3, :3] = dR @ rig.poses[f, :3, :3]
        rig.poses[f, :3, 3] += rng.normal(0, 0.1, 3)
    for c in rig.cameras:
        c.intrinsics[:2] *= 0.8
    return rig, 0.7
