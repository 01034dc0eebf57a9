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
