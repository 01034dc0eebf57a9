#!/usr/bin/env python
"""Regenerates the committed golden fixtures under tests/golden/.

Run in the build container (needs cv2 for the homography initial poses):

    python tests/golden/make_golden.py

Writes, per reference test (``radtan``, ``equidistant``; /root/reference/test/<type>/):
  * ``<type>_init_poses.npy``   [900,4,4] float64 -- initial cam0 poses, restating
        /root/reference/src/get_init_pose.cpp:81-140 with Python cv2 on the regenerated corners
  * ``<type>_corners_digest.json`` -- counts and a sha256 of the regenerated corners
        (the corners themselves are rebuilt deterministically at test time by
        ``multicam_calibration_b200.synth.make_test_corners``; the digest pins them)
and ``reference_logs.json``: the numbers transcribed from the reference's own logs
(/root/reference/test/radtan/correct_output.txt, test/equidistant/correct_output.txt) that the oracle is
pinned against.  No file under /root/reference is read at test time.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from multicam_calibration_b200 import synth  # noqa: E402


def digest(npts, w, im):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(npts, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(w, dtype=np.float32).tobytes())
    h.update(np.ascontiguousarray(im, dtype=np.float32).tobytes())
    return h.hexdigest()


def main():
    for model in ("radtan", "equidistant"):
        npts, w, im = synth.make_test_corners(model)
        cam = synth.KAT_CAMERAS[model]
        off = np.concatenate([[0], np.cumsum(npts)])
        poses = np.zeros((len(npts), 4, 4))
        for f in range(len(npts)):
            poses[f] = synth.init_pose_homography(w[off[f]:off[f + 1]], im[off[f]:off[f + 1]], cam["intrinsics"], cam["distortion_coeffs"], model)
        np.save(os.path.join(HERE, f"{model}_init_poses.npy"), poses)
        with open(os.path.join(HERE, f"{model}_corners_digest.json"), "w") as fh:
            json.dump(dict(frames=int(len(npts)), points=int(npts.sum()), residuals=int(2 * npts.sum()), sha256=digest(npts, w, im)), fh, indent=1)
        print(model, "frames", len(npts), "points", int(npts.sum()))

    logs = {
        # /root/reference/test/radtan/correct_output.txt:13-27, 49-55, 68, 75, 79-81
        "radtan": dict(
            num_params=5412, num_residuals=715662,
            cost=[3.753430e-01, 6.909148e-06, 6.808999e-06, 6.741513e-06, 6.701133e-06, 6.694080e-06, 6.693676e-06, 6.693400e-06,
                  6.693219e-06, 6.693186e-06, 6.693185e-06, 6.693185e-06],
            cost_change=[0.0, 3.75e-01, 1.00e-07, 6.75e-08, 4.04e-08, 7.05e-09, 4.04e-10, 2.76e-10, 1.81e-10, 3.25e-11, 1.07e-12, 5.01e-15],
            gradient=[2.29e+04, 1.78e+00, 1.28e-02, 7.61e-03, 2.52e-03, 4.78e-04, 1.44e-04, 8.59e-05, 3.44e-05, 6.12e-06, 4.10e-07, 6.94e-08],
            step=[0.0, 1.54e-03, 2.22e-04, 4.05e-04, 4.62e-04, 2.42e-04, 4.80e-05, 3.11e-06, 1.76e-06, 9.62e-07, 1.96e-07, 1.39e-08],
            tr_ratio=[0.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.01],
            tr_radius=[1.00e+04, 3.00e+04, 9.00e+04, 2.70e+05, 8.10e+05, 2.43e+06, 7.29e+06, 2.19e+07, 6.56e+07, 1.97e+08, 5.90e+08, 1.77e+09],
            initial_cost=3.753430e-01, final_cost=6.693185e-06, iterations=12, successful=12, unsuccessful=0,
            termination="parameter", termination_value=9.110088e-13,
            distortion_coeffs=[0.155003, -0.180002, -0.005, -0.000300001, 0.0440008, 0.00500274, 0.00399778, -0.000399239],
            intrinsics=[151, 152, 173, 144], avg_reproj_err=6.11635e-06, max_reproj_err=2.37361e-05, max_err_frame=84),
        # /root/reference/test/equidistant/correct_output.txt:13-24, 46-52, 65, 72, 76-78
        "equidistant": dict(
            num_params=5408, num_residuals=720000,
            initial_cost=8.718698e-04, final_cost=6.403777e-06, iterations=9, successful=8, unsuccessful=1,
            termination="function",
            distortion_coeffs=[0.155, -0.180001, -0.00499917, -0.000300337],
            intrinsics=[151, 152, 173, 144], avg_reproj_err=5.96461e-06, max_reproj_err=1.99193e-05, max_err_frame=9),
    }
    with open(os.path.join(HERE, "reference_logs.json"), "w") as fh:
        json.dump(logs, fh, indent=1)


if __name__ == "__main__":
    main()
