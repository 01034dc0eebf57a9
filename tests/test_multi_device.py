"""Several GPUs behind ONE handle (mcc_set_devices / Calibrator::setDevices / mcc_calibrate --devices): one process, the
reference's own call sequence, frames dealt out over the devices inside the library.  Needs >= 2 GPUs; on a one-GPU box
these tests skip (the one-process-per-GPU mode is covered by tests/test_multi_rank.py)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

PKG = os.path.join(ROOT, "multicam_calibration_b200")


def _devices():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run under gpurun --gpus 2)")
    return list(range(8 if n >= 8 else (4 if n >= 4 else 2)))


def test_set_devices_needs_a_gpu_or_fails_loudly(built):
    """without a CUDA device the call reports MCC_ERR_NO_DEVICE (no CPU fallback); with one, bad lists are refused"""
    import torch
    from multicam_calibration_b200 import capi
    if not torch.cuda.is_available():
        with pytest.raises(capi.MccError) as e:
            capi.Calibrator([0, 1])
        assert e.value.status == -7
    else:
        with pytest.raises(capi.MccError):
            capi.Calibrator([0, 0])
        with pytest.raises(capi.MccError):
            capi.Calibrator([0, 99])


@pytest.mark.gpu
@pytest.mark.parametrize("how", ["bulk", "per_frame"])
def test_multi_device_handle_matches_oracle(built, how):
    """solve, testCalibration, removeOutliers + re-solve through ONE handle over all GPUs of the box against the oracle"""
    from multicam_calibration_b200 import capi, synth
    from multicam_calibration_b200.rig import Rig
    import oracle
    devs = _devices()
    base = synth.make_synthetic_rig(3, 40, 20260112, detect_prob=0.8)
    rig = Rig([c.copy() for c in base.cameras], base.npts, base.world, base.image.copy(), base.poses)
    rng = np.random.default_rng(9)
    bad = rng.choice(rig.num_obs, 25, replace=False)
    rig.image[bad] += rng.uniform(8, 20, (25, 2))
    cal = capi.Calibrator(devs)
    cal.setCameras(rig.cameras)
    if how == "bulk":
        cal.addFrames(rig)
    else:                                   # the nodelet's way: addPoints once per frame (calibration_nodelet.cpp:642)
        off = rig.frame_offsets()
        for f in range(rig.num_frames):
            lo = off[f]
            wp, ip = [], []
            for k in range(rig.num_cameras):
                n = int(rig.npts[f, k])
                wp.append(rig.world[lo:lo + n]); ip.append(rig.image[lo:lo + n]); lo += n
            cal.addPoints(f, wp, ip, rig.poses[f])
    orc = oracle.Oracle(rig)
    sg, so = cal.runCalibration(), orc.solve()
    assert cal.commTransport() == 1                     # peer windows between the devices, no NCCL
    assert abs(sg.final_cost - so.final_cost) <= 1e-9 * so.final_cost
    assert (sg.num_iterations, sg.num_successful, sg.num_unsuccessful, sg.why) == (so.num_iterations, so.num_successful, so.num_unsuccessful, so.why)
    assert sg.num_params == so.num_params and sg.num_residuals == so.num_residuals and sg.num_reduced_params == so.num_reduced_params
    pg, po = cal.getParams(), orc.get_params()
    assert np.max(np.abs(pg - po) / np.maximum(np.abs(po), 1e-3)) <= 1e-7
    # testCalibration: frame -> camera -> point order of the whole problem, whichever device holds the frame
    rg, ro = cal.testCalibration().reshape(-1), orc.residuals()
    assert rg.shape == ro.shape and np.max(np.abs(rg - ro)) <= 1e-8
    assert abs(cal.evalCost(True) - sg.final_cost) <= 1e-12 * sg.final_cost
    # removeOutliers + second solve (calibration_nodelet.cpp:202-207)
    ng, no = cal.removeOutliers(3.0), orc.remove_outliers(3.0)
    assert ng == no and ng >= 20
    assert np.array_equal(cal.getMask(), orc.get_mask())
    orc.init_params()
    sg, so = cal.runCalibration(), orc.solve()
    assert abs(sg.final_cost - so.final_cost) <= 1e-9 * so.final_cost and sg.num_iterations == so.num_iterations
    res = cal.getCalibrationResults()
    for k in range(rig.num_cameras):
        intr, dist, T = orc.get_camera(k)
        np.testing.assert_allclose(res[k].intrinsics, intr, rtol=1e-7)
        if k > 0:
            np.testing.assert_allclose(res[k].T_cn_cnm1, T, rtol=1e-7, atol=1e-10)
    cal.close()


@pytest.mark.gpu
def test_multi_device_rejected_steps_and_restart(built):
    """the rejected-steps fixture through one multi-device handle, then a restart from the solution (keep_params)"""
    from multicam_calibration_b200 import capi
    import oracle
    from fixtures_util import rejected_steps_rig
    devs = _devices()
    rig, mrd = rejected_steps_rig()
    cal = capi.calibrator_from_rig(rig, device=devs)
    og, oo = capi.default_options(), oracle.default_options()
    og.min_relative_decrease = oo.min_relative_decrease = mrd
    orc = oracle.Oracle(rig)
    sg, so = cal.runCalibration(og), orc.solve(oo)
    assert (sg.num_iterations, sg.num_unsuccessful) == (so.num_iterations, so.num_unsuccessful) and so.num_unsuccessful == 5
    assert abs(sg.final_cost - so.final_cost) <= 1e-9 * so.final_cost
    p = cal.getParams()
    cal.setParams(p)                      # goes back out to the devices
    og.keep_params = 1
    s2 = cal.runCalibration(og)
    assert s2.num_iterations <= 4 and abs(s2.final_cost - sg.final_cost) <= 1e-10 * sg.final_cost
    cal.close()


@pytest.mark.gpu
def test_facade_and_cli_on_several_devices(built, tmp_path):
    """Calibrator::setDevices through the C++ facade driver, and mcc_calibrate --devices on the reference's radtan fixture:
    the same logged result as on one device (/root/reference/test/radtan/correct_output.txt)"""
    from multicam_calibration_b200 import synth
    import oracle
    import test_cpp_facade as tf
    devs = _devices()
    dl = ",".join(str(d) for d in devs)
    rig = synth.make_synthetic_rig(3, 30, 20260111, detect_prob=0.85)
    tf._write_rig_file(rig, tmp_path / "rig.txt")
    exe = tf._build_facade_gpu(tmp_path)
    r = subprocess.run([str(exe), str(tmp_path / "rig.txt"), dl, "run"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "FACADE_GPU_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    res, s, p = tf._parse_facade_output(r.stdout)
    orc = oracle.Oracle(rig)
    so = orc.solve()
    po = orc.get_params()
    assert abs(s["final_cost"] - so.final_cost) <= 1e-9 * so.final_cost and s["num_iterations"] == so.num_iterations
    assert np.max(np.abs(p - po) / np.maximum(np.abs(po), 1e-3)) <= 1e-7
    ro = orc.residuals().reshape(-1, 2)
    o = 0
    for fr in range(rig.num_frames):
        for k in range(rig.num_cameras):
            n = int(rig.npts[fr, k])
            assert np.max(np.abs(res[(fr, k)] - ro[o:o + n])) <= 1e-8 if n else True
            o += n
    # the CLI
    tf._write_kat_inputs("radtan", tmp_path)
    out_yaml = tmp_path / "radtan-out.yaml"
    r = subprocess.run([os.path.join(PKG, "mcc_calibrate"), "--calib", str(tmp_path / "initial.yaml"), "--corners", str(tmp_path / "corners.csv"),
                        "--init-poses", str(tmp_path / "poses.txt"), "--output", str(out_yaml), "--devices", dl], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "Parameter tolerance reached" in r.stdout and "Minimizer iterations                       12" in r.stdout
    y = open(out_yaml).read()
    assert "  distortion_coeffs: [0.155003, -0.180002, -0.005, -0.000300001, 0.0440008, 0.00500274, 0.00399778, -0.000399239]\n" in y
    assert "max error: 2.37361e-05 px at frame: 84 for cam: 0" in r.stdout
