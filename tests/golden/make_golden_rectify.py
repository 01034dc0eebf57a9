#!/usr/bin/env python
"""Golden vectors for the stereo rectification entries of the output YAML.

The reference computes them with OpenCV in compute_RPQ (/root/reference/src/calibration_nodelet.cpp:373-410):
cv::fisheye::stereoRectify for equidistant cameras (rvec / tvec narrowed to float through cv::Affine3f::Vec3,
get_init_pose.cpp:33-44) and cv::stereoRectify for radtan ones (R, T as double), both with CALIB_ZERO_DISPARITY.
This script runs Python cv2 on a few camera pairs and stores inputs and outputs in rectify_golden.json.

    python tests/golden/make_golden_rectify.py
"""
import json
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def euler(ax, ay, az):
    cx, sx, cy, sy, cz, sz = np.cos(ax), np.sin(ax), np.cos(ay), np.sin(ay), np.cos(az), np.sin(az)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def K(i):
    return np.array([[i[0], 0, i[2]], [0, i[1], i[3]], [0, 0, 1.0]])


def main():
    cases = []
    readme_T = np.array([[0.99999965648, -0.00013331925, 0.00081808159, -0.19946344647],
                         [0.00013388601, 0.99999975107, -0.00069277676, -0.00005674605],
                         [-0.00081798903, 0.00069288605, 0.99999942540, 0.00010022941], [0, 0, 0, 1.0]])
    T2 = np.eye(4); T2[:3, :3] = euler(0.02, -0.015, 0.01); T2[:3, 3] = [-0.12, 0.004, -0.003]
    T3 = np.eye(4); T3[:3, :3] = euler(-0.01, 0.03, -0.02); T3[:3, 3] = [0.08, -0.002, 0.005]     # second camera on the left
    i1, i2 = [605.054, 604.66, 641.791, 508.728], [605.097, 604.321, 698.772, 573.558]
    eq = [([-0.0146915, 0.000370396, -0.00425216, 0.0015107], [-0.0146155, -0.00291494, -0.000681981, 0.000221855])]
    rt = [([-0.05, 0.02, 0.001, -0.0005], [-0.04, 0.025, 0.0008, -0.0002]), ([0.03, -0.01, 0.0, 0.0], [0.02, 0.0, 0.001, 0.0])]
    size = (1280, 1024)
    for T in (readme_T, T2, T3):
        for d1, d2 in eq:
            rvec = cv2.Rodrigues(T[:3, :3])[0].reshape(3).astype(np.float32).astype(np.float64)   # Affine3f::Vec3
            tvec = T[:3, 3].astype(np.float32).astype(np.float64)
            R1, R2, P1, P2, Q = cv2.fisheye.stereoRectify(K(i1), np.array(d1), K(i2), np.array(d2), size, rvec, tvec, flags=cv2.CALIB_ZERO_DISPARITY)
            cases.append(dict(model="equidistant", intr1=i1, intr2=i2, dist1=d1, dist2=d2, resolution=size, T_cn_cnm1=T.tolist(),
                              R1=R1.tolist(), R2=R2.tolist(), P1=P1.tolist(), P2=P2.tolist()))
        for d1, d2 in rt:
            R1, R2, P1, P2, Q, _, _ = cv2.stereoRectify(K(i1), np.array(d1), K(i2), np.array(d2), size, T[:3, :3].copy(), T[:3, 3].copy(),
                                                        flags=cv2.CALIB_ZERO_DISPARITY)
            cases.append(dict(model="radtan", intr1=i1, intr2=i2, dist1=d1, dist2=d2, resolution=size, T_cn_cnm1=T.tolist(),
                              R1=R1.tolist(), R2=R2.tolist(), P1=P1.tolist(), P2=P2.tolist()))
    with open(os.path.join(HERE, "rectify_golden.json"), "w") as f:
        json.dump(dict(cv2=cv2.__version__, cases=cases), f, indent=1)
    print(len(cases), "cases")


if __name__ == "__main__":
    main()
