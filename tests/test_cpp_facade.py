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


@pytest.mark.gpu
def test_cli_staged_workflow_matches_api(built, tmp_path):
    """the staged calibration of example_calib_manager.py:22-52 (cam0 intrinsics, then cam1 intrinsics, then the
    extrinsics between them; every stage = set_parameter calls + one calibrate(), calibration_nodelet.cpp:196-285)
    through `mcc_calibrate --stage ...`, against the same sequence driven through the C ABI from Python"""
    from multicam_calibration_b200 import capi, synth
    from multicam_calibration_b200.rig import Rig
    rig = synth.make_synthetic_rig(2, 40, 20260107, readme_stereo=True)
    for k, c in enumerate(rig.cameras):
        c.name, c.rostopic, c.resolution = "cam%d" % k, "/cam%d/image_raw" % k, (1280, 1024)
    # the csv carries %f text parsed as float (calibration_nodelet.cpp:566): narrow the rig the same way
    nar = lambda a: np.array([float(np.float32(float("%f" % x))) for x in a.ravel()]).reshape(a.shape)
    world = np.concatenate([nar(rig.world[:, :2]), np.zeros((rig.num_obs, 1))], axis=1)
    image = nar(rig.image)
    rig = Rig(rig.cameras, rig.npts, world, image, rig.poses)
    with open(tmp_path / "corners.csv", "w") as f:
        o = 0
        for fr in range(rig.num_frames):
            for k in range(2):
                for i in range(rig.npts[fr, k]):
                    f.write("%d, intrinsics, %d, %f, %f, %f, %f\n" % (fr, k, world[o, 0], world[o, 1], image[o, 0], image[o, 1]))
                    o += 1
    with open(tmp_path / "initial.yaml", "w") as f:
        for k, c in enumerate(rig.cameras):
            f.write("cam%d:\n  camera_model: pinhole\n  intrinsics: [%s]\n  distortion_model: %s\n  distortion_coeffs: [%s]\n"
                    % (k, ", ".join(repr(float(x)) for x in c.intrinsics), c.distortion_model, ", ".join(repr(float(x)) for x in c.distortion_coeffs)))
            if k > 0:
                f.write("  T_cn_cnm1:\n" + "".join("  - [%s]\n" % ", ".join(repr(float(x)) for x in row) for row in c.T_cn_cnm1))
            f.write("  resolution: [1280, 1024]\n  rostopic: %s\n" % c.rostopic)
    np.savetxt(tmp_path / "poses.txt", rig.poses.reshape(rig.num_frames, 16), fmt="%.17g")
    stages = [((False, True, True), (True, True, False)), ((True, True, False), (False, True, True)), ((True, True, True), (True, False, True))]
    args = []
    for st in stages:
        args += ["--stage", ";".join("cam%d:fix_intrinsics=%d,fix_extrinsics=%d,active=%d" % ((k,) + tuple(int(b) for b in fl)) for k, fl in enumerate(st))]
    out_yaml = tmp_path / "staged.yaml"
    r = subprocess.run([os.path.join(PKG, "mcc_calibrate"), "--calib", str(tmp_path / "initial.yaml"), "--corners", str(tmp_path / "corners.csv"),
                        "--init-poses", str(tmp_path / "poses.txt"), "--output", str(out_yaml), "--quiet"] + args, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("writing calibration to") == 3
    # the same three stages through the ABI
    cal = capi.calibrator_from_rig(rig, device=0)
    cams = [c.copy() for c in rig.cameras]
    for st in stages:
        for k, fl in enumerate(st):
            cams[k].fix_intrinsics, cams[k].fix_extrinsics, cams[k].active = fl
        cal.setCameras(cams)
        cal.runCalibration()
        for k, res in enumerate(cal.getCalibrationResults()):
            cams[k].intrinsics, cams[k].distortion_coeffs, cams[k].T_cn_cnm1 = res.intrinsics, res.distortion_coeffs, res.T_cn_cnm1
    cal.close()
    y = open(out_yaml).read()
    import re
    got = [np.array([float(x) for x in m.split(",")]) for m in re.findall(r"  intrinsics: \[(.*)\]", y)]
    gotd = [np.array([float(x) for x in m.split(",")]) for m in re.findall(r"  distortion_coeffs: \[(.*)\]", y)]
    tail = y[y.index("T_cn_cnm1:"):]
    rows = [np.array([float(x) for x in m.split(",")]) for m in re.findall(r"  - \[(.*)\]", tail)]
    assert len(got) == 2 and len(rows) >= 4
    for k in range(2):
        np.testing.assert_allclose(got[k], cams[k].intrinsics, rtol=2e-5)            # the YAML prints 6 significant digits
        np.testing.assert_allclose(gotd[k], cams[k].distortion_coeffs, rtol=2e-5, atol=1e-9)
    np.testing.assert_allclose(np.array(rows[:4]), cams[1].T_cn_cnm1, rtol=2e-5, atol=1e-9)
    # and the stages did something: the generating cameras are recovered from the perturbed start
    for k, c in enumerate(rig_truth := synth.make_synthetic_rig(2, 40, 20260107, readme_stereo=True).truth["cameras"]):
        np.testing.assert_allclose(cams[k].intrinsics, c.intrinsics, rtol=5e-3)


def _write_rig_file(rig, path):
    """the text format tests/cpp/facade_gpu.cpp reads"""
    with open(path, "w") as f:
        f.write("%d %d\n" % (rig.num_cameras, rig.num_frames))
        for c in rig.cameras:
            f.write("%d %d %s %s %s %d %d %d\n" % (c.model_id, len(c.distortion_coeffs), " ".join(repr(float(x)) for x in c.intrinsics),
                                                     " ".join(repr(float(x)) for x in c.distortion_coeffs),
                                                     " ".join(repr(float(x)) for x in np.asarray(c.T_cn_cnm1).reshape(16)),
                                                     int(c.fix_intrinsics), int(c.fix_extrinsics), int(c.active)))
        o = 0
        for fr in range(rig.num_frames):
            f.write(" ".join(repr(float(x)) for x in rig.poses[fr].reshape(16)) + "\n")
            for k in range(rig.num_cameras):
                n = int(rig.npts[fr, k])
                f.write("%d\n" % n)
                for i in range(n):
                    f.write(" ".join(repr(float(x)) for x in (*rig.world[o + i], *rig.image[o + i])) + "\n")
                o += n


def _build_facade_gpu(tmp_path):
    exe = tmp_path / "facade_gpu"
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([gxx, "-O1", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", str(exe),
                           os.path.join(ROOT, "tests", "cpp", "facade_gpu.cpp"), "-L", PKG, "-lmcc", f"-Wl,-rpath,{PKG}"])
    return exe


def _parse_facade_output(out):
    res, summary, params = {}, None, None
    for line in out.splitlines():
        t = line.split()
        if not t:
            continue
        if t[0] == "RES":
            res[(int(t[1]), int(t[2]))] = np.array([float(x) for x in t[4:]]).reshape(int(t[3]), 2)
        elif t[0] == "SUMMARY":
            summary = dict(initial_cost=float(t[1]), final_cost=float(t[2]), num_iterations=int(t[3]), num_successful=int(t[4]),
                           num_unsuccessful=int(t[5]), why=int(t[6]))
        elif t[0] == "PARAMS":
            params = np.array([float(x) for x in t[2:]])
    return res, summary, params


def test_facade_gpu_driver_compiles(built, tmp_path):
    """the GPU driver of the facade builds on a CPU-only machine too (it is only run by the -m gpu tests)"""
    assert os.path.exists(_build_facade_gpu(tmp_path))


@pytest.mark.gpu
def test_facade_test_calibration_is_per_camera(built, tmp_path):
    """Calibrator::testCalibration through the C++ facade on a 2-camera rig with unequal point counts.

    The reference restarts its running index into the frame's residual vector for every camera
    (`size_t res_idx = 0;` inside the camera loop, /root/reference/src/calibrator.cpp:337-346), so for cameras >= 1
    it hands back camera 0's residuals (and reads past them when camera 1 saw more points).  This library
    deliberately does NOT reproduce that: residuals[frame][cam] holds that camera's own residuals, in the order
    FrameResidual wrote them (cameras ascending, calibrator.cpp:135-142).  This test pins the chosen behaviour against the
    oracle: camera 1's entries equal the oracle's residuals for camera 1 and differ from camera 0's."""
    from multicam_calibration_b200 import synth
    import oracle
    rig = synth.make_synthetic_rig(2, 6, 20260110, detect_prob=0.7)
    assert np.any(rig.npts[:, 0] != rig.npts[:, 1])
    _write_rig_file(rig, tmp_path / "rig.txt")
    exe = _build_facade_gpu(tmp_path)
    r = subprocess.run([str(exe), str(tmp_path / "rig.txt")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "FACADE_GPU_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    res, _, _ = _parse_facade_output(r.stdout)
    ro = oracle.Oracle(rig).residuals().reshape(-1, 2)      # frame -> camera -> point order
    o = 0
    for fr in range(rig.num_frames):
        for k in range(2):
            n = int(rig.npts[fr, k])
            got = res[(fr, k)]
            assert got.shape == (n, 2)
            assert np.max(np.abs(got - ro[o:o + n])) <= 1e-10
            o += n
        n0, n1 = int(rig.npts[fr, 0]), int(rig.npts[fr, 1])
        m = min(n0, n1)
        assert m > 0 and not np.allclose(res[(fr, 1)][:m], res[(fr, 0)][:m])   # what the reference's res_idx reset would return


@pytest.mark.gpu
def test_facade_run_calibration_matches_oracle(built, tmp_path):
    """the reference's call sequence through the C++ facade (setCameras, addPoints per frame, runCalibration,
    testCalibration; calibration_nodelet.cpp:196-208, 642) against the oracle at the north-star tolerances"""
    from multicam_calibration_b200 import synth
    import oracle
    rig = synth.make_synthetic_rig(3, 30, 20260111, detect_prob=0.85)
    _write_rig_file(rig, tmp_path / "rig.txt")
    exe = _build_facade_gpu(tmp_path)
    r = subprocess.run([str(exe), str(tmp_path / "rig.txt"), "-", "run"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "FACADE_GPU_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    _, s, p = _parse_facade_output(r.stdout)
    orc = oracle.Oracle(rig)
    so = orc.solve()
    po = orc.get_params()
    assert abs(s["final_cost"] - so.final_cost) <= 1e-9 * so.final_cost
    assert np.max(np.abs(p - po) / np.maximum(np.abs(po), 1e-3)) <= 1e-7
    assert s["num_iterations"] == so.num_iterations and s["why"] == so.why
