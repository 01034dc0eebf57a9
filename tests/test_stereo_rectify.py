"""CPU: rectification_matrix / projection_matrix against golden vectors made with Python cv2
(tests/golden/make_golden_rectify.py: cv2.fisheye.stereoRectify / cv2.stereoRectify with CALIB_ZERO_DISPARITY, the calls
of the reference's compute_RPQ, calibration_nodelet.cpp:373-410).  Tolerances: rotations 5e-9; focal length and
baseline term 1e-6 relative; principal point 2e-3 px (OpenCV averages float32 points there)."""
import json
import os
import subprocess

import numpy as np

from conftest import GOLDEN, ROOT


def test_compute_rpq_matches_cv2(tmp_path):
    exe = tmp_path / "rectify_cli"
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([gxx, "-O1", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", str(exe),
                           os.path.join(ROOT, "tests", "cpp", "rectify_cli.cpp")])
    cases = json.load(open(os.path.join(GOLDEN, "rectify_golden.json")))["cases"]
    lines = []
    for c in cases:
        vals = [c["model"], c["resolution"][0], c["resolution"][1]] + c["intr1"] + c["intr2"] + [len(c["dist1"])] + c["dist1"] + c["dist2"]
        vals += [float(x) for x in np.array(c["T_cn_cnm1"]).reshape(-1)]
        lines.append(" ".join(repr(float(v)) if isinstance(v, float) else str(v) for v in vals))
    out = subprocess.run([str(exe)], input="\n".join(lines) + "\n", capture_output=True, text=True, check=True).stdout.strip().splitlines()
    assert len(out) == len(cases)
    for c, line in zip(cases, out):
        v = np.array([float(x) for x in line.split()])
        R1, R2, P1, P2 = v[:9].reshape(3, 3), v[9:18].reshape(3, 3), v[18:30].reshape(3, 4), v[30:42].reshape(3, 4)
        np.testing.assert_allclose(R1, c["R1"], atol=5e-9)   # cv2.Rodrigues re-orthonormalises the 11-digit README matrix by SVD first
        np.testing.assert_allclose(R2, c["R2"], atol=5e-9)
        for P, want in ((P1, np.array(c["P1"])), (P2, np.array(c["P2"]))):
            assert abs(P[0, 0] - want[0, 0]) <= 1e-6 * want[0, 0] and P[1, 1] == P[0, 0]
            assert abs(P[0, 2] - want[0, 2]) <= 2e-3 and abs(P[1, 2] - want[1, 2]) <= 2e-3
            assert abs(P[0, 3] - want[0, 3]) <= 1e-6 * max(1.0, abs(want[0, 3]))
            assert P[2, 2] == 1.0 and P[0, 1] == 0 and P[1, 0] == 0 and P[1, 3] == 0
