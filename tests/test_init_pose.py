"""CPU: the homography initial pose (mcc_init_pose) against the committed cv2 fixtures.

The golden poses were produced with Python cv2 (tests/golden/make_golden.py restating get_init_pose.cpp:81-140).
OpenCV cannot be linked into the library, so the numerical kernels differ (cv2.undistortPoints stops after 5
fixed-point iterations, ours converges; cv2.findHomography's normalisation and refinement differ in detail):
tolerance 2e-4 rad on the rotation and 1e-4 relative on the translation for radtan, 1e-6 for equidistant."""
import numpy as np
import pytest

from conftest import GOLDEN
import os


@pytest.mark.parametrize("model,rot_tol,t_tol", [("radtan", 2e-4, 1e-4), ("equidistant", 1e-6, 1e-6)])
def test_init_pose_matches_cv2_fixture(built, model, rot_tol, t_tol):
    from multicam_calibration_b200 import capi, synth
    from multicam_calibration_b200.rig import Camera
    poses = np.load(os.path.join(GOLDEN, f"{model}_init_poses.npy"))
    npts, w, im = synth.make_test_corners(model)
    cam = synth.KAT_CAMERAS[model]
    c = Camera(model, np.array(cam["intrinsics"]), np.array(cam["distortion_coeffs"]))
    off = np.concatenate([[0], np.cumsum(npts)])
    for f in range(0, len(npts), 7):
        ww = np.concatenate([w[off[f]:off[f + 1]], np.zeros((npts[f], 1))], axis=1)
        T = capi.init_pose(c, ww, im[off[f]:off[f + 1]])
        R = T[:3, :3]
        assert np.allclose(R.T @ R, np.eye(3), atol=1e-12) and abs(np.linalg.det(R) - 1) < 1e-12   # a proper rotation
        ang = np.arccos(min(1.0, (np.trace(R.T @ poses[f][:3, :3]) - 1) / 2))
        assert ang <= rot_tol, (f, ang)
        assert np.linalg.norm(T[:3, 3] - poses[f][:3, 3]) <= t_tol * np.linalg.norm(poses[f][:3, 3]), f


def test_init_pose_recovers_synthetic_truth(built):
    """noise-free synthetic views of both models: the homography pose equals the generating pose"""
    from multicam_calibration_b200 import capi, synth
    rig = synth.make_synthetic_rig(2, 12, 5, detect_prob=1.0, noise_px=0.0, perturb=False)
    off = rig.frame_offsets()
    T_c_c0 = rig.truth["T_c_c0"]
    for f in range(rig.num_frames):
        o = off[f]
        for k, cam in enumerate(rig.cameras):
            n = rig.npts[f, k]
            if n >= 8:
                T = capi.init_pose(cam, rig.world[o:o + n], rig.image[o:o + n])
                want = T_c_c0[k] @ rig.truth["poses"][f]
                np.testing.assert_allclose(T[:3, :3], want[:3, :3], atol=2e-4)   # float32 observations
                np.testing.assert_allclose(T[:3, 3], want[:3, 3], atol=2e-4)
            o += n


def test_init_pose_errors(built):
    from multicam_calibration_b200 import capi
    from multicam_calibration_b200.rig import Camera
    c = Camera("equidistant", np.array([600.0, 600.0, 640.0, 512.0]), np.zeros(4))
    with pytest.raises(capi.MccError) as e:
        capi.init_pose(c, np.zeros((3, 3)), np.zeros((3, 2)))     # fewer than 4 points
    assert e.value.status == -1
    with pytest.raises(capi.MccError):
        capi.init_pose(c, np.zeros((10, 3)), np.zeros((10, 2)))   # all points identical: degenerate


def test_init_pose_batch_equals_single_views(built):
    """mcc_init_pose_batch: every view bit-identical to mcc_init_pose, whatever the thread count; a failing view gets the
    zero matrix and its own status, and does not fail the call (get_init_pose.cpp:88 returns zeros in that case)"""
    import time
    from multicam_calibration_b200 import capi, synth
    rig = synth.make_synthetic_rig(3, 60, 17, detect_prob=0.8)
    off = rig.frame_offsets()
    views = []
    for f in range(rig.num_frames):
        o = off[f]
        for k, cam in enumerate(rig.cameras):
            n = int(rig.npts[f, k])
            if n:
                views.append((cam, rig.world[o:o + n], rig.image[o:o + n]))
            o += n
    views.append((rig.cameras[0], np.zeros((3, 3)), np.zeros((3, 2))))        # too few points
    views.append((rig.cameras[1], np.zeros((10, 3)), np.zeros((10, 2))))      # degenerate
    single = []
    for cam, w, im in views[:-2]:
        single.append(capi.init_pose(cam, w, im))
    t0 = time.perf_counter()
    T1, s1 = capi.init_pose_batch(views, n_threads=1)
    t1 = time.perf_counter()
    T4, s4 = capi.init_pose_batch(views, n_threads=4)
    t4 = time.perf_counter()
    T0, s0 = capi.init_pose_batch(views, n_threads=0)
    assert np.array_equal(T1, T4) and np.array_equal(T1, T0) and np.array_equal(s1, s4) and np.array_equal(s1, s0)
    assert np.array_equal(T1[:-2], np.array(single))
    assert np.all(s1[:-2] == 0) and s1[-2] == -1 and s1[-1] != 0
    assert not T1[-2:].any()
    print("init_pose_batch: %d views, 1 thread %.1f ms, 4 threads %.1f ms" % (len(views), 1e3 * (t1 - t0), 1e3 * (t4 - t1)))
    T, s = capi.init_pose_batch([])
    assert T.shape == (0, 4, 4) and len(s) == 0


def test_cli_dry_run_writes_the_facade_poses(built, tmp_path):
    """mcc_calibrate --dry-run --init-poses-out (no GPU needed): the CLI's batched guessCameraPoses over a 2-camera corners
    file gives, frame by frame, cam0's pose from the first camera that saw the board (calibration_nodelet.cpp:510-543)"""
    import subprocess
    from conftest import ROOT
    from multicam_calibration_b200 import capi, synth
    rig = synth.make_synthetic_rig(2, 10, 23, detect_prob=0.7)
    off = rig.frame_offsets()
    with open(tmp_path / "corners.csv", "w") as fcsv:
        for f in range(rig.num_frames):
            o = off[f]
            for k in range(2):
                n = int(rig.npts[f, k])
                for i in range(n):
                    fcsv.write("%d, intrinsics, %d, %.9g, %.9g, %.9g, %.9g\n" % (f, k, rig.world[o + i, 0], rig.world[o + i, 1], rig.image[o + i, 0], rig.image[o + i, 1]))
                o += n
    with open(tmp_path / "initial.yaml", "w") as fy:
        for k, c in enumerate(rig.cameras):
            fy.write("cam%d:\n  camera_model: pinhole\n  intrinsics: [%s]\n  distortion_model: %s\n  distortion_coeffs: [%s]\n  resolution: [1280, 1024]\n  rostopic: /cam%d/image_raw\n"
                     % (k, ", ".join(repr(float(x)) for x in c.intrinsics), c.distortion_model, ", ".join(repr(float(x)) for x in c.distortion_coeffs), k))
            if k > 0:
                fy.write("  T_cn_cnm1:\n" + "".join("  - [%s]\n" % ", ".join(repr(float(x)) for x in row) for row in np.asarray(c.T_cn_cnm1)))
    exe = os.path.join(ROOT, "multicam_calibration_b200", "mcc_calibrate")
    r = subprocess.run([exe, "--calib", str(tmp_path / "initial.yaml"), "--corners", str(tmp_path / "corners.csv"), "--dry-run",
                        "--init-poses-out", str(tmp_path / "poses.txt"), "--pose-threads", "3"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "dry run" in r.stdout, r.stdout + r.stderr
    lines = open(tmp_path / "poses.txt").read().splitlines()
    assert len(lines) == rig.num_frames
    checked = 0
    for f in range(rig.num_frames):
        o = off[f]
        want = None
        T_n_0 = np.eye(4)
        for k, cam in enumerate(rig.cameras):
            T_n_0 = np.asarray(cam.T_cn_cnm1) @ T_n_0 if k > 0 else np.eye(4)
            n = int(rig.npts[f, k])
            if n >= 4 and want is None:
                # the CLI parses the csv's 9 significant digits: recompute from the same rounded values
                w = np.array([[float("%.9g" % v) for v in row] for row in rig.world[o:o + n]])
                im = np.array([[float("%.9g" % v) for v in row] for row in rig.image[o:o + n]])
                w[:, 2] = 0.0
                want = np.linalg.inv(T_n_0) @ capi.init_pose(cam, w, im)
            o += n
        if want is None:
            assert lines[f].strip() == "nan"
            continue
        got = np.array([float(x) for x in lines[f].split()]).reshape(4, 4)
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-9)
        checked += 1
    assert checked >= 5
