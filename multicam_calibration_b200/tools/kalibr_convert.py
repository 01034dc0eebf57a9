#!/usr/bin/env python3
"""Wire-format neighbours of the calibration YAML, Python 3, no ROS.

    python -m multicam_calibration_b200.tools.kalibr_convert camerainfo -i calib.yaml -o cam0.yaml [-c cam0] [-n name]
    python -m multicam_calibration_b200.tools.kalibr_convert urdf -f calib.yaml -o rig.urdf -p base -c cam0 cam1 --frame_ids left right

`camerainfo` does what /root/reference/src/kalibr_to_camerainfo.py does (one camera of a Kalibr file as an
OpenCV / ROS camera_info YAML: K from the four intrinsics, radtan renamed plumb_bob, identity rectification,
P = [K | 0]).  `urdf` does what /root/reference/src/kalibr_to_urdf.py does (the T_cn_cnm1 chain as fixed joints: each
joint carries the inverse transform, i.e. the pose of camera n in the frame of camera n-1, as xyz + static-axes
roll/pitch/yaw); the reference takes the angles from tf.transformations.euler_from_matrix, restated here in numpy.
"""
import argparse
import math
import sys

import numpy as np
import yaml


def camera_info_from_kalibr(cam: dict, camera_name=None) -> dict:
    fx, fy, cx, cy = (float(v) for v in cam["intrinsics"])
    K = [fx, 0.0, cx, 0.0, fy, cy, 0.0, 0.0, 1.0]
    model = cam["distortion_model"]
    coeffs = [float(v) for v in cam["distortion_coeffs"]]
    return {
        "image_width": int(cam["resolution"][0]),
        "image_height": int(cam["resolution"][1]),
        "camera_name": cam["rostopic"] if camera_name is None else camera_name,
        "camera_matrix": {"rows": 3, "cols": 3, "data": K},
        "distortion_model": "plumb_bob" if model == "radtan" else model,
        "distortion_coefficients": {"rows": 1, "cols": len(coeffs), "data": coeffs},
        "rectification_matrix": {"rows": 3, "cols": 3, "data": [1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0]},
        "projection_matrix": {"rows": 3, "cols": 4, "data": [fx, 0.0, cx, 0.0, 0.0, fy, cy, 0.0, 0.0, 0.0, 1.0, 0.0]},
    }


def rpy_from_matrix(R) -> tuple:
    """static-axes x-y-z angles (roll, pitch, yaw) with R = Rz(yaw) Ry(pitch) Rx(roll); at the pitch = +-90 deg
    singularity yaw is set to zero, as tf.transformations does"""
    R = np.asarray(R, dtype=np.float64)
    cy = math.hypot(R[0, 0], R[1, 0])
    if cy > 4.0 * np.finfo(float).eps:
        return math.atan2(R[2, 1], R[2, 2]), math.atan2(-R[2, 0], cy), math.atan2(R[1, 0], R[0, 0])
    return math.atan2(-R[1, 2], R[1, 1]), math.atan2(-R[2, 0], cy), 0.0


def joints_from_kalibr(kalibr: dict, cameras, frame_ids, parent_frame) -> list:
    """[(parent, child, xyz, rpy)] for the chain parent_frame -> frame_ids[0] -> frame_ids[1] ..."""
    if len(cameras) != len(frame_ids):
        raise ValueError("need one frame id per camera")
    names = [parent_frame] + list(frame_ids)
    joints = []
    for i, cam in enumerate(cameras):
        T = np.eye(4)
        if "T_cn_cnm1" in kalibr[cam]:
            T = np.linalg.inv(np.asarray(kalibr[cam]["T_cn_cnm1"], dtype=np.float64))
        joints.append((names[i], names[i + 1], tuple(float(v) for v in T[:3, 3]), rpy_from_matrix(T[:3, :3])))
    return joints


def urdf_text(joints) -> str:
    out = []
    for parent, child, xyz, rpy in joints:
        out.append('<link name="%s"/>\n' % child)
        out.append('<joint name="%s_to_%s" type="fixed">\n' % (parent, child))
        out.append('  <parent link="%s"/>\n  <child link="%s"/>\n' % (parent, child))
        out.append('  <origin xyz="%.4f %.4f %.4f" rpy="%.4f %.4f %.4f"/>\n' % (xyz + tuple(rpy)))
        out.append('  <axis xyz="0 0 1"/>\n</joint>\n')
    return "".join(out)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    sub = ap.add_subparsers(dest="cmd", required=True)
    a = sub.add_parser("camerainfo", help="one camera of a Kalibr file -> camera_info YAML")
    a.add_argument("-i", "--input", required=True)
    a.add_argument("-o", "--output", required=True)
    a.add_argument("-c", "--camera", default="cam0")
    a.add_argument("-n", "--camera_name", default=None)
    b = sub.add_parser("urdf", help="T_cn_cnm1 chain -> urdf joints")
    b.add_argument("-c", "--cameras", nargs="+", required=True)
    b.add_argument("--frame_ids", nargs="+", required=True)
    b.add_argument("-p", "--parent_frame", required=True)
    b.add_argument("-f", "--input_file", required=True)
    b.add_argument("-o", "--output_file", required=True)
    args = ap.parse_args(argv)
    if args.cmd == "camerainfo":
        with open(args.input) as f:
            kal = yaml.safe_load(f)
        with open(args.output, "w") as f:
            yaml.safe_dump(camera_info_from_kalibr(kal[args.camera], args.camera_name), f)
    else:
        with open(args.input_file) as f:
            kal = yaml.safe_load(f)
        with open(args.output_file, "w") as f:
            f.write(urdf_text(joints_from_kalibr(kal, args.cameras, args.frame_ids, args.parent_frame)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
