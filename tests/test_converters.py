"""the Kalibr -> camera_info / urdf converters (CPU only)."""
import numpy as np
import yaml

from multicam_calibration_b200.tools import kalibr_convert as kc

README_T = [[0.99999965648, -0.00013331925, 0.00081808159, -0.19946344647],
            [0.00013388601, 0.99999975107, -0.00069277676, -0.00005674605],
            [-0.00081798903, 0.00069288605, 0.99999942540, 0.00010022941], [0, 0, 0, 1.0]]
KALIBR = {
    "cam0": {"camera_model": "pinhole", "intrinsics": [605.054, 604.66, 641.791, 508.728], "distortion_model": "equidistant",
             "distortion_coeffs": [-0.0146915, 0.000370396, -0.00425216, 0.0015107], "resolution": [1280, 1024], "rostopic": "/rig/left/image_mono"},
    "cam1": {"camera_model": "pinhole", "intrinsics": [605.097, 604.321, 698.772, 573.558], "distortion_model": "radtan",
             "distortion_coeffs": [-0.05, 0.02, 0.001, -0.0005], "T_cn_cnm1": README_T, "resolution": [1280, 1024], "rostopic": "/rig/right/image_mono"},
}


def test_camerainfo_fields(tmp_path):
    ci = kc.camera_info_from_kalibr(KALIBR["cam1"])
    assert ci["distortion_model"] == "plumb_bob" and ci["camera_name"] == "/rig/right/image_mono"
    assert ci["image_width"] == 1280 and ci["image_height"] == 1024
    K = np.array(ci["camera_matrix"]["data"]).reshape(3, 3)
    np.testing.assert_array_equal(K, [[605.097, 0, 698.772], [0, 604.321, 573.558], [0, 0, 1]])
    P = np.array(ci["projection_matrix"]["data"]).reshape(3, 4)
    np.testing.assert_array_equal(P[:, :3], K)
    np.testing.assert_array_equal(P[:, 3], 0)
    np.testing.assert_array_equal(np.array(ci["rectification_matrix"]["data"]).reshape(3, 3), np.eye(3))
    assert ci["distortion_coefficients"] == {"rows": 1, "cols": 4, "data": [-0.05, 0.02, 0.001, -0.0005]}
    assert kc.camera_info_from_kalibr(KALIBR["cam0"], "left")["camera_name"] == "left"
    assert kc.camera_info_from_kalibr(KALIBR["cam0"])["distortion_model"] == "equidistant"
    # command line round trip
    src, dst = tmp_path / "k.yaml", tmp_path / "ci.yaml"
    src.write_text(yaml.safe_dump(KALIBR))
    assert kc.main(["camerainfo", "-i", str(src), "-o", str(dst), "-c", "cam1"]) == 0
    assert yaml.safe_load(dst.read_text()) == ci


def _rot(roll, pitch, yaw):
    cx, sx, cy, sy, cz, sz = np.cos(roll), np.sin(roll), np.cos(pitch), np.sin(pitch), np.cos(yaw), np.sin(yaw)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def test_rpy_round_trip_and_urdf(tmp_path):
    rng = np.random.default_rng(5)
    for _ in range(50):
        a = rng.uniform([-3.1, -1.5, -3.1], [3.1, 1.5, 3.1])
        np.testing.assert_allclose(kc.rpy_from_matrix(_rot(*a)), a, atol=1e-12)
    r, p, y = kc.rpy_from_matrix(_rot(0.3, np.pi / 2, 0.0))           # singular pitch: yaw folded into roll
    assert y == 0.0 and abs(p - np.pi / 2) < 1e-7
    np.testing.assert_allclose(_rot(r, p, y), _rot(0.3, np.pi / 2, 0.0), atol=1e-7)
    joints = kc.joints_from_kalibr(KALIBR, ["cam0", "cam1"], ["left", "right"], "base")
    assert joints[0][:2] == ("base", "left") and joints[0][2] == (0.0, 0.0, 0.0) and joints[0][3] == (0.0, 0.0, 0.0)
    Tinv = np.linalg.inv(np.array(README_T))
    assert joints[1][:2] == ("left", "right")
    np.testing.assert_allclose(joints[1][2], Tinv[:3, 3], atol=1e-15)
    np.testing.assert_allclose(_rot(*joints[1][3]), Tinv[:3, :3], atol=1e-12)
    src, dst = tmp_path / "k.yaml", tmp_path / "rig.urdf"
    src.write_text(yaml.safe_dump(KALIBR))
    assert kc.main(["urdf", "-f", str(src), "-o", str(dst), "-p", "base", "-c", "cam0", "cam1", "--frame_ids", "left", "right"]) == 0
    text = dst.read_text()
    assert text.count("<joint ") == 2 and '<joint name="left_to_right" type="fixed">' in text
    assert '<origin xyz="0.1995 0.0000 0.0001" rpy="-0.0007 -0.0008 -0.0001"/>' in text
