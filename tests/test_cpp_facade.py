"""The C++ host layer above the C ABI (include/multicam_calibration/*.h + csrc/cli.cpp)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

PKG = os.path.join(ROOT, "multicam_calibration_b200")


def test_facade_cpp_host_logic(built, tmp_path):
    """compiles tests/cpp/test_facade.cpp against the facade headers and libmcc.so and runs it (no GPU needed)"""
    exe = tmp_path / "test_facade"
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([gxx, "-O1", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", str(exe),
                           os.path.join(ROOT, "tests", "cpp", "test_facade.cpp"), "-L", PKG, "-lmcc", f"-Wl,-rpath,{PKG}"])
    r = subprocess.run([str(exe), str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0 and "FACADE_OK" in r.stdout, r.stdout + r.stderr


def _write_kat_inputs(model, tmp_path):
    """corners.csv exactly as make_test_corners.py writes it, the reference's initial YAML, and the golden poses"""
    from multicam_calibration_b200 import synth
    npts, w, im = synth.make_test_corners(model)
    cam = synth.KAT_CAMERAS[model]
    frames = np.repeat(np.arange(len(npts)), npts)
    with open(tmp_path / "corners.csv", "w") as f:
        for fr, (wx, wy), (u, v) in zip(frames, w, im):
            f.write("%d, intrinsics, %d, %f, %f, %f, %f\n" % (fr, 0, wx, wy, u, v))
    with open(tmp_path / "initial.yaml", "w") as f:
        f.write("cam0:\n  camera_model: pinhole\n  intrinsics: [%s]\n  distortion_model: %s\n  distortion_coeffs: [%s]\n  resolution: [352, 287]\n  rostopic: /cam/image_raw\n"
                % (", ".join("%g" % x for x in cam["intrinsics"]), model, ", ".join(repr(x) for x in cam["distortion_coeffs"])))
    poses = np.load(os.path.join(GOLDEN, f"{model}_init_poses.npy"))
    np.savetxt(tmp_path / "poses.txt", poses.reshape(len(poses), 16), fmt="%.17g")


@pytest.mark.gpu
def test_cli_radtan_fixture_with_own_initial_poses(built, tmp_path):
    """same flow, but the initial poses come from the library's own homography step (guessCameraPose,
    calibration_nodelet.cpp:510-543) instead of a file: the solve lands on the reference's logged result"""
    _write_kat_inputs("radtan", tmp_path)
    out_yaml = tmp_path / "radtan-out.yaml"
    r = subprocess.run([os.path.join(PKG, "mcc_calibrate"), "--calib", str(tmp_path / "initial.yaml"), "--corners", str(tmp_path / "corners.csv"),
                        "--output", str(out_yaml)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    so = r.stdout
    assert "read 900 frames from" in so and "Num params: 5412" in so and "Num residuals: 715662" in so
    assert "Parameter tolerance reached" in so
    y = open(out_yaml).read()
    assert "  intrinsics: [151, 152, 173, 144]\n" in y
    assert "  distortion_coeffs: [0.155003, -0.180002, -0.005, -0.000300001, 0.0440008, 0.00500274, 0.00399778, -0.000399239]\n" in y
    assert "max error: 2.37361e-05 px at frame: 84 for cam: 0" in so


@pytest.mark.gpu
def test_cli_radtan_fixture_matches_reference_log(built, tmp_path):
    """the reference's `roslaunch test_calib.launch type:=radtan` flow, headless: same console facts and YAML as
    /root/reference/test/radtan/correct_output.txt"""
    import json
    ref = json.load(open(os.path.join(GOLDEN, "reference_logs.json")))["radtan"]
    _write_kat_inputs("radtan", tmp_path)
    out_yaml = tmp_path / "radtan-out.yaml"
    r = subprocess.run([os.path.join(PKG, "mcc_calibrate"), "--calib", str(tmp_path / "initial.yaml"), "--corners", str(tmp_path / "corners.csv"),
                        "--init-poses", str(tmp_path / "poses.txt"), "--output", str(out_yaml), "--residuals-dir", str(tmp_path)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    so = r.stdout
    assert "Found 1 cameras!" in so and "read 900 frames from corners file for cam 0" in so
    assert "Num params: 5412" in so and "Num residuals: 715662" in so
    assert "Minimizer iterations                       12" in so
    assert "Parameter tolerance reached" in so
    rows = [l.split() for l in so.splitlines() if l.strip() and l.split()[0].isdigit() and len(l.split()) == 10]
    assert len(rows) == 12 and rows[0][1] == "3.753431e-01" or rows[0][1] == "3.753430e-01"
    assert rows[11][1] == "6.693185e-06" and rows[5][1] == "6.694080e-06"
    # the YAML the reference prints (correct_output.txt:71-77)
    y = open(out_yaml).read()
    assert "cam0:\n  camera_model: pinhole\n  intrinsics: [151, 152, 173, 144]\n  distortion_model: radtan\n" in y
    assert "  distortion_coeffs: [0.155003, -0.180002, -0.005, -0.000300001, 0.0440008, 0.00500274, 0.00399778, -0.000399239]\n" in y
    assert y.endswith("  resolution: [352, 287]\n  rostopic: /cam/image_raw\n")
    assert "max error: 2.37361e-05 px at frame: 84 for cam: 0" in so
    res = np.loadtxt(tmp_path / "residuals_cam0.txt")
    assert res.shape == (ref["num_residuals"] // 2, 10)
