"""CPU: the C-ABI library loads, exports every symbol include/mcc.h declares, and its host-side logic (parameter
packing, result unpacking, sharding) agrees with the oracle.  No compute call needs a GPU here; the compute
entry points must fail loudly without one."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def test_header_symbols_exported(built):
    from multicam_calibration_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "mcc.h")).read()
    declared = set(re.findall(r"\b(mcc_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"mcc_handle", "mcc_status"}     # "(mcc_status (*fn)..." style false positives
    L = ctypes.CDLL(capi.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    L.mcc_version.restype = ctypes.c_char_p
    assert b"sm_100a" in L.mcc_version()


def test_library_is_sm100a_only(built):
    """the product is built for sm_100a and nothing else"""
    import subprocess
    from multicam_calibration_b200 import capi
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("no cuobjdump")
    out = subprocess.run([cuobjdump, "-lelf", capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_param_packing_matches_oracle(small_rig):
    """initializeVariables (calibrator.cpp:251-283) and transform_to_rvec_tvec (calibrator.cpp:157-170)"""
    from multicam_calibration_b200 import capi
    import oracle
    cal = capi.calibrator_from_rig(small_rig)
    cal.initParams()
    o = oracle.Oracle(small_rig)
    assert cal.num_params == o.num_params == small_rig.num_cam_params + 6 * small_rig.num_frames
    np.testing.assert_array_equal(cal.getParams(), o.get_params())
    cal.close()


def test_rvec_round_trip_and_branches():
    """the quaternion branches of the matrix -> angle-axis route, incl. rotations by ~pi (the test fixtures' poses)"""
    from multicam_calibration_b200 import capi
    import oracle
    rng = np.random.default_rng(0)
    mats = []
    for _ in range(50):
        v = rng.normal(size=3)
        v *= rng.uniform(0.01, np.pi) / np.linalg.norm(v)
        mats.append(v)
    mats += [np.array([np.pi - 1e-9, 0, 0]), np.array([0, np.pi - 1e-7, 0]), np.array([0, 0, 3.1]), np.array([2.2, 2.2, 0.0]), np.zeros(3)]
    for v in mats:
        T = capi.rvec_tvec_to_transform(np.concatenate([v, [0.1, -0.2, 0.3]]))
        np.testing.assert_allclose(T, oracle.rvec_tvec_to_transform(np.concatenate([v, [0.1, -0.2, 0.3]])), atol=1e-15)
        a, b = capi.transform_to_rvec_tvec(T), oracle.transform_to_rvec_tvec(T)
        np.testing.assert_allclose(a, b, atol=1e-15)
        np.testing.assert_allclose(capi.rvec_tvec_to_transform(a), T, atol=1e-12)


def test_get_camera_unpacks_chain(small_rig):
    """getCalibrationResults (calibrator.cpp:415-463): T_cn_cnm1 = T_cn_c0 * inv(T_cnm1_c0)"""
    from multicam_calibration_b200 import capi
    import oracle
    cal = capi.calibrator_from_rig(small_rig)
    cal.initParams()
    o = oracle.Oracle(small_rig)
    res = cal.getCalibrationResults()
    for k, c in enumerate(small_rig.cameras):
        intr, dist, T = o.get_camera(k)
        np.testing.assert_allclose(res[k].intrinsics, intr, rtol=0, atol=0)
        np.testing.assert_allclose(res[k].distortion_coeffs, dist, rtol=0, atol=0)
        if k > 0:
            np.testing.assert_allclose(res[k].T_cn_cnm1, T, atol=1e-14)
            np.testing.assert_allclose(res[k].T_cn_cnm1, c.T_cn_cnm1, atol=1e-12)   # unpack(pack(x)) == x
    cal.close()


def test_compute_fails_loudly_without_gpu(small_rig):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from multicam_calibration_b200 import capi
    cal = capi.calibrator_from_rig(small_rig)
    with pytest.raises(capi.MccError) as e:
        cal.runCalibration()
    assert e.value.status == -7 and "no CPU fallback" in str(e.value)
    with pytest.raises(capi.MccError):
        cal.testCalibration()
    cal.close()


def test_error_conventions_host(small_rig):
    from multicam_calibration_b200 import capi
    cal = capi.Calibrator()
    with pytest.raises(capi.MccError) as e:
        cal.runCalibration()
    assert e.value.status == -1            # "no data to run on!"
    cal.setCameras(small_rig.cameras)
    with pytest.raises(capi.MccError) as e:
        cal.addPoints(0, [np.zeros((0, 3))], [np.zeros((0, 2))], np.eye(4))
    assert e.value.status == -2            # camera-count mismatch
    bad = small_rig.cameras[0].copy()
    bad.distortion_coeffs = np.zeros(9)
    with pytest.raises(capi.MccError):
        cal.setCameras([bad])
    cal.close()


def test_shard_bounds_balanced():
    from multicam_calibration_b200.rig import shard_bounds
    rng = np.random.default_rng(1)
    npts = rng.integers(0, 140, size=(1000, 8))
    for ws in (1, 2, 4, 8):
        b = shard_bounds(npts, ws)
        assert b[0] == 0 and b[-1] == 1000 and np.all(np.diff(b) >= 0)
        per = [npts[b[r]:b[r + 1]].sum() for r in range(ws)]
        assert max(per) - min(per) <= 2 * 8 * 140
