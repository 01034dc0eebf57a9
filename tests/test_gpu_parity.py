"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same inputs.

Tolerances (BASELINE.json north_star): final cost <= 1e-9 relative, parameters <= 1e-7 relative, same LM
iteration count.  Unit-level tolerances are stated in each test.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b, floor=1e-300):
    return np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(np.asarray(b))), floor)


def _param_rel(pg, po):
    # relative per parameter, with an absolute floor for parameters that are ~0 (tangential coefficients, angles)
    return np.max(np.abs(pg - po) / np.maximum(np.abs(po), 1e-3))


def _pair(rig):
    from multicam_calibration_b200 import capi
    import oracle
    cal = capi.calibrator_from_rig(rig, device=0)
    cal.initParams()
    return cal, oracle.Oracle(rig)


def test_residuals_match_oracle(small_rig):
    """testCalibration (calibrator.cpp:316-351): raw residuals, same order, |diff| <= 1e-10 px"""
    cal, orc = _pair(small_rig)
    rg = cal.testCalibration().reshape(-1)
    ro = orc.residuals()
    assert rg.shape == ro.shape
    assert np.max(np.abs(rg - ro)) <= 1e-10
    cal.close()


@pytest.mark.parametrize("use_loss", [True, False])
def test_normal_blocks_match_oracle(small_rig, use_loss):
    """K1: analytic Jacobian blocks vs the oracle's Jet autodiff blocks; relative to each block's max, <= 1e-10"""
    cal, orc = _pair(small_rig)
    g = cal.normalBlocks(use_loss)
    o = orc.normal_blocks(use_loss)
    assert abs(g["cost"] - o["cost"]) <= 1e-12 * o["cost"]
    assert list(g["frames"]) == list(o["frames"])
    for key in ("Hcc", "gc", "Hff", "gf", "Hfc"):
        assert g[key].shape == o[key].shape, key
        assert _rel(g[key], o[key]) <= 1e-10, (key, _rel(g[key], o[key]))
    # per frame as well (small frames must not hide behind large ones)
    for s in range(len(o["frames"])):
        assert _rel(g["Hff"][s], o["Hff"][s]) <= 1e-10
        assert _rel(g["Hfc"][s], o["Hfc"][s]) <= 1e-10
    cal.close()


def test_jacobian_guards(built):
    """rotation vector below 1e-8 (utils.h:31-38: identity, zero derivative) and a point on the optical axis
    (utils.h:69-72) -- blocks must match the oracle's Jets there too"""
    from multicam_calibration_b200 import synth, capi
    import oracle
    rig = synth.make_synthetic_rig(2, 6, 3, detect_prob=1.0, noise_px=0.1, mixed_models=False)
    # frame 0: zero rotation pose; camera 1 extrinsic rotation zero as well
    rig.poses[0, :3, :3] = np.eye(3)
    rig.cameras[1].T_cn_cnm1[:3, :3] = np.eye(3)
    # put the first world point of frame 0 / cam 0 exactly on the optical axis of cam0
    off = rig.frame_offsets()
    t = rig.poses[0, :3, 3]
    rig.world[off[0], :] = [-t[0], -t[1], 0.0]
    cal, orc = _pair(rig)
    g, o = cal.normalBlocks(True), orc.normal_blocks(True)
    for key in ("Hcc", "gc", "Hff", "gf", "Hfc"):
        assert _rel(g[key], o[key]) <= 1e-10, key
    # the rotation block of frame 0 and of the cam1 extrinsics must be exactly zero
    assert np.all(g["Hff"][0][:3, :] == 0.0)
    assert np.all(o["Hff"][0][:3, :] == 0.0)
    cal.close()


def _check_solve(cal, orc, opt_g=None, opt_o=None, same_iters=True):
    from multicam_calibration_b200 import capi
    sg = cal.runCalibration(opt_g)
    so = orc.solve(opt_o)
    assert abs(sg.initial_cost - so.initial_cost) <= 1e-9 * so.initial_cost
    assert abs(sg.final_cost - so.final_cost) <= 1e-9 * so.final_cost, (sg.final_cost, so.final_cost)
    pg, po = cal.getParams(), orc.get_params()
    assert _param_rel(pg, po) <= 1e-7, _param_rel(pg, po)
    if same_iters:
        assert sg.num_iterations == so.num_iterations, (sg.num_iterations, so.num_iterations)
        assert sg.why == so.why
        assert sg.num_successful == so.num_successful and sg.num_unsuccessful == so.num_unsuccessful
        tg, to = sg.trace(), so.trace()
        # the radius depends continuously on tr_ratio = cost_change / model_cost_change, and cost_change of the last
        # rows is a difference of nearly equal costs: compare loosely
        np.testing.assert_allclose(tg["radius"], to["radius"], rtol=2e-3)
        np.testing.assert_allclose(tg["cost"], to["cost"], rtol=1e-9)
    return sg, so


def test_solve_small_rig(small_rig):
    cal, orc = _pair(small_rig)
    _check_solve(cal, orc)
    # getCalibrationResults (calibrator.cpp:415-463)
    res = cal.getCalibrationResults()
    for k in range(small_rig.num_cameras):
        intr, dist, T = orc.get_camera(k)
        np.testing.assert_allclose(res[k].intrinsics, intr, rtol=1e-7)
        np.testing.assert_allclose(res[k].distortion_coeffs, dist, rtol=1e-7, atol=1e-10)
        if k > 0:
            np.testing.assert_allclose(res[k].T_cn_cnm1, T, rtol=1e-7, atol=1e-10)
    cal.close()


def test_solve_radtan_kat(kat_rigs):
    """config c2: the reference's test/radtan fixture; must land on the reference's logged result
    (/root/reference/test/radtan/correct_output.txt) exactly like the oracle does"""
    import json, os
    from conftest import GOLDEN
    ref = json.load(open(os.path.join(GOLDEN, "reference_logs.json")))["radtan"]
    cal, orc = _pair(kat_rigs["radtan"])
    sg, so = _check_solve(cal, orc)
    assert sg.num_residuals == ref["num_residuals"] and sg.num_params == ref["num_params"]
    assert sg.num_iterations == ref["iterations"] == 12
    assert sg.why == 1  # parameter tolerance
    assert abs(sg.final_cost - ref["final_cost"]) <= 5e-7 * ref["final_cost"]   # printed to 7 digits
    res = cal.getCalibrationResults()[0]
    for got, want in zip(res.distortion_coeffs, ref["distortion_coeffs"]):
        assert abs(got - want) <= 5.1e-6 * abs(want) + 1e-12   # printed to 6 significant digits
    cal.close()


def test_solve_equidistant_kat(kat_rigs):
    """config c1: test/equidistant.  The tail of this solve sits at the rounding floor (|cost change|/cost ~1e-11),
    so the iteration count is compared within +-3 (as the oracle is against the log) and the results at full tolerance."""
    import json, os
    from conftest import GOLDEN
    ref = json.load(open(os.path.join(GOLDEN, "reference_logs.json")))["equidistant"]
    cal, orc = _pair(kat_rigs["equidistant"])
    sg, so = _check_solve(cal, orc, same_iters=False)
    assert abs(sg.num_iterations - so.num_iterations) <= 3
    assert abs(sg.final_cost - ref["final_cost"]) <= 5e-7 * ref["final_cost"]
    res = cal.getCalibrationResults()[0]
    for got, want in zip(res.distortion_coeffs, ref["distortion_coeffs"]):
        assert abs(got - want) <= 5.1e-6 * abs(want) + 1e-12
    r = cal.testCalibration()
    assert abs(np.hypot(r[:, 0], r[:, 1]).max() - ref["max_reproj_err"]) <= 1e-9
    cal.close()


def test_solve_c3_stereo(c3_rig):
    """config c3: README-scale stereo rig, 409 frames, ~202k residuals, 2476 params"""
    cal, orc = _pair(c3_rig)
    sg, so = _check_solve(cal, orc)
    assert sg.num_params == 2476
    cal.close()


@pytest.mark.parametrize("flags", [
    dict(fix_all=True),                                     # node-level fix_intrinsics
    dict(fix_intr=[1, 0, 0]),                               # one camera's intrinsics fixed
    dict(fix_extr=[0, 1, 0]),                               # one extrinsic block fixed
    dict(active=[1, 0, 1]),                                 # example_calib_manager.py: a camera switched off
    dict(fix_intr=[1, 1, 1], fix_extr=[0, 1, 1]),           # extrinsics-only stage
])
def test_fix_and_active_flags(small_rig, flags):
    """the fix / active matrix of example_calib_manager.py:22-52 and calibrator.cpp:510-521"""
    from multicam_calibration_b200 import capi
    from multicam_calibration_b200.rig import Rig
    import oracle
    cams = [c.copy() for c in small_rig.cameras]
    for k, c in enumerate(cams):
        c.fix_intrinsics = bool(flags.get("fix_intr", [0, 0, 0])[k])
        c.fix_extrinsics = bool(flags.get("fix_extr", [0, 0, 0])[k])
        c.active = bool(flags.get("active", [1, 1, 1])[k])
    rig = Rig(cams, small_rig.npts, small_rig.world, small_rig.image, small_rig.poses, fix_all_intrinsics=flags.get("fix_all", False))
    cal, orc = _pair(rig)
    sg, so = _check_solve(cal, orc)
    assert sg.num_reduced_params == so.num_reduced_params
    assert sg.num_residuals == so.num_residuals
    cal.close()


def test_remove_outliers_then_resolve(small_rig):
    """calibrate() with outlier_pixel_threshold (calibration_nodelet.cpp:202-207): solve, drop, solve again"""
    from multicam_calibration_b200.rig import Rig
    rig = Rig([c.copy() for c in small_rig.cameras], small_rig.npts, small_rig.world, small_rig.image.copy(), small_rig.poses)
    rng = np.random.default_rng(5)
    bad = rng.choice(rig.num_obs, 25, replace=False)
    rig.image[bad] += rng.uniform(8, 20, (25, 2))
    cal, orc = _pair(rig)
    _check_solve(cal, orc)
    ng, no = cal.removeOutliers(3.0), orc.remove_outliers(3.0)
    assert ng == no and ng >= 20
    assert np.array_equal(cal.getMask(), orc.get_mask())
    orc.init_params()          # runCalibration restarts from the initial values (calibrator.cpp:293-294)
    _check_solve(cal, orc)
    cal.close()


def test_empty_and_ragged_frames(built):
    """frames where some / all cameras saw nothing (calibrator.cpp:482-485: no residual block, pose untouched)"""
    from multicam_calibration_b200 import synth
    from multicam_calibration_b200.rig import Rig
    base = synth.make_synthetic_rig(3, 16, 21, detect_prob=0.7)
    off = base.frame_offsets()
    keep = np.ones(base.num_obs, dtype=bool)
    npts = base.npts.copy()
    # frame 2: nothing at all; frame 5: only camera 1; frame 7: camera 0 missing
    def drop(f, k):
        o = off[f] + npts[f, :k].sum()
        keep[o:o + npts[f, k]] = False
    for f, ks in ((2, (0, 1, 2)), (5, (0, 2)), (7, (0,))):
        for k in ks:
            drop(f, k)
    for f, ks in ((2, (0, 1, 2)), (5, (0, 2)), (7, (0,))):
        for k in ks:
            npts[f, k] = 0
    rig = Rig(base.cameras, npts, base.world[keep], base.image[keep], base.poses)
    cal, orc = _pair(rig)
    p0 = cal.getParams().copy()
    _check_solve(cal, orc)
    pg = cal.getParams()
    C = rig.num_cam_params
    assert np.array_equal(pg[C + 12:C + 18], p0[C + 12:C + 18])   # pose of the empty frame is untouched
    cal.close()


def test_add_points_one_by_one_equals_bulk(small_rig):
    """Calibrator::addPoints frame by frame == the bulk entry point"""
    from multicam_calibration_b200 import capi
    rig = small_rig
    cal = capi.Calibrator(0)
    cal.setCameras(rig.cameras)
    off = rig.frame_offsets()
    for f in range(rig.num_frames):
        wp, ip, o = [], [], off[f]
        for k in range(rig.num_cameras):
            n = rig.npts[f, k]
            wp.append(rig.world[o:o + n]); ip.append(rig.image[o:o + n]); o += n
        cal.addPoints(f, wp, ip, rig.poses[f])
    bulk = capi.calibrator_from_rig(rig, 0)
    s1, s2 = cal.runCalibration(), bulk.runCalibration()
    assert s1.final_cost == s2.final_cost and s1.num_iterations == s2.num_iterations   # bit-reproducible
    assert np.array_equal(cal.getParams(), bulk.getParams())
    cal.close(); bulk.close()


def test_error_conventions(small_rig):
    from multicam_calibration_b200 import capi
    cal = capi.Calibrator(0)
    with pytest.raises(capi.MccError) as e:
        cal.runCalibration()                       # "no data to run on!" (calibrator.cpp:286-289)
    assert e.value.status == -1
    cal.setCameras(small_rig.cameras)
    with pytest.raises(capi.MccError) as e:
        cal.addPoints(0, [np.zeros((0, 3))], [np.zeros((0, 2))], np.eye(4))   # camera-count mismatch (calibrator.cpp:236-240)
    assert e.value.status == -2
    cal.close()


def test_pinned_and_pageable_inputs_agree(small_rig, c3_rig):
    """mcc_add_frames DMAs page-locked caller arrays straight to the device and stages pageable ones on the host;
    mixing both in one handle must give the bit-identical solve"""
    import torch
    from multicam_calibration_b200 import capi
    from multicam_calibration_b200.rig import Rig
    rig = c3_rig
    ref = capi.calibrator_from_rig(rig, 0)
    s0 = ref.runCalibration()
    half = rig.num_frames // 2
    a, b = rig.slice_frames(0, half), rig.slice_frames(half, rig.num_frames)
    wp, ip = torch.from_numpy(a.world).pin_memory(), torch.from_numpy(a.image).pin_memory()
    a_pinned = Rig(a.cameras, a.npts, wp.numpy(), ip.numpy(), a.poses)
    cal = capi.Calibrator(0)
    cal.setCameras(rig.cameras)
    cal.addFrames(a_pinned)      # direct DMA
    cal.addFrames(b)             # pageable -> staged (pulls the first half back into the staging buffers)
    s1 = cal.runCalibration()
    assert s1.final_cost == s0.final_cost and s1.num_iterations == s0.num_iterations
    assert np.array_equal(cal.getParams(), ref.getParams())
    # and again after clear, pageable first then pinned
    cal.clearFrames()
    wp2, ip2 = torch.from_numpy(b.world).pin_memory(), torch.from_numpy(b.image).pin_memory()
    cal.addFrames(a)
    cal.addFrames(Rig(b.cameras, b.npts, wp2.numpy(), ip2.numpy(), b.poses))
    s2 = cal.runCalibration()
    assert s2.final_cost == s0.final_cost and np.array_equal(cal.getParams(), ref.getParams())
    cal.close(); ref.close()


def test_frames_with_thousands_of_points(built):
    """a frame whose tiles do not fit the Jacobian pass's shared memory (30 000 points of one camera = 188 tiles) takes the
    global-memory parking of the tile Grams: same blocks, same solve as the oracle"""
    from multicam_calibration_b200 import synth
    from multicam_calibration_b200.rig import Camera, Rig
    rng = np.random.default_rng(12)
    cam = Camera("equidistant", np.array([600.0, 601.0, 640.0, 512.0]), np.array([-0.0147, 0.0004, -0.004, 0.0015]))
    worlds, images, poses = [], [], []
    for f in range(3):
        n = 30000 if f < 2 else 500
        w = np.concatenate([rng.uniform(-0.3, 0.3, (n, 2)), np.zeros((n, 1))], axis=1)
        T = np.eye(4)
        T[:3, :3] = synth.rodrigues(np.array([0.11 - 0.1 * f, -0.07, 0.2 * f]))
        T[:3, 3] = [0.02, -0.03, 0.7 + 0.1 * f]
        uv = synth.project("equidistant", w @ T[:3, :3].T + T[:3, 3], cam.intrinsics, cam.distortion_coeffs) + rng.normal(0, 0.3, (n, 2))
        Tg = T.copy()
        Tg[:3, 3] += [0.004, -0.003, 0.006]               # start a little off
        worlds.append(w); images.append(uv); poses.append(Tg)
    start = Camera("equidistant", cam.intrinsics * 1.002, cam.distortion_coeffs + 1e-4)
    rig = Rig([start], np.array([[30000], [30000], [500]], dtype=np.int32), np.concatenate(worlds), np.concatenate(images), np.array(poses))
    cal, orc = _pair(rig)
    g, o = cal.normalBlocks(True), orc.normal_blocks(True)
    for key in ("Hcc", "gc", "Hff", "gf", "Hfc"):
        assert _rel(g[key], o[key]) <= 1e-10, (key, _rel(g[key], o[key]))
    _check_solve(cal, orc)
    cal.close()


def test_full_size_c4_properties(built):
    """config c4 at BASELINE.json's full size (8 cameras, 10 000 frames, ~20 M residuals, 60 106 parameters): too big
    for the oracle in a test, so size-independent properties are checked instead"""
    from multicam_calibration_b200 import capi, synth
    rig = synth.config_c4()
    assert rig.num_cam_params == 106 and rig.num_cam_params + 6 * rig.num_frames == 60106
    cal = capi.calibrator_from_rig(rig, device=0)
    s = cal.runCalibration()
    assert s.termination == 0 and s.num_residuals == 2 * rig.num_obs
    tr = s.trace()
    # accepted rows: strictly decreasing cost; gradient drops by orders of magnitude
    acc = tr["cost"][np.concatenate([[True], tr["ratio"][1:] > 1e-3])]
    assert np.all(np.diff(acc) < 0)
    assert tr["gradient_max"][-1] < 1e-4 * tr["gradient_max"][0]
    p = cal.getParams()
    # (1) the cost the solver reports is the cost of the parameters it returns, by two independent device paths:
    #     the Jacobian pass (summary) and the residual-only pass (mcc_eval_cost)
    c_eval = cal.evalCost(use_loss=True)
    assert abs(c_eval - s.final_cost) <= 1e-12 * s.final_cost
    # (2) ... and by the residuals themselves, Huber applied per frame on the host (checksum of checksums)
    r = cal.testCalibration()
    e2 = r[:, 0] ** 2 + r[:, 1] ** 2
    per_frame = np.add.reduceat(e2, np.concatenate([[0], np.cumsum(rig.npts.sum(axis=1))[:-1]]))
    rho = np.where(per_frame > 1.0, 2.0 * np.sqrt(per_frame) - 1.0, per_frame)
    assert abs(0.5 * rho.sum() - s.final_cost) <= 1e-11 * s.final_cost
    # (3) idempotence: solving again from the initial values gives the same bits
    s2 = cal.runCalibration()
    assert s2.final_cost == s.final_cost and s2.num_iterations == s.num_iterations
    assert np.array_equal(cal.getParams(), p)
    # (4) restarting from the solution: nothing left to gain (a handful of noise-floor iterations at most)
    opt = capi.default_options()
    opt.keep_params = 1
    s3 = cal.runCalibration(opt)
    assert s3.num_iterations <= 4 and abs(s3.final_cost - s.final_cost) <= 1e-10 * s.final_cost
    # (5) the generating cameras are recovered (noise 0.2 px, 10 M corners)
    res = cal.getCalibrationResults()
    for k, c in enumerate(rig.truth["cameras"]):
        np.testing.assert_allclose(res[k].intrinsics, c.intrinsics, rtol=2e-4)
        np.testing.assert_allclose(res[k].distortion_coeffs, c.distortion_coeffs, atol=2e-3)
        if k > 0:
            np.testing.assert_allclose(res[k].T_cn_cnm1[:3, 3], c.T_cn_cnm1[:3, 3], atol=2e-4)
    cal.close()


@pytest.mark.parametrize("models", [
    [("equidistant", 2), ("radtan", 5), ("radtan", 8)],     # odd coefficient counts (utils.h:73-78 any number; :115-118 zero-padded to 8)
    [("radtan", 0), ("equidistant", 0)],                    # no distortion parameters at all
    [("equidistant", 4), ("equidistant", 8), ("radtan", 2), ("radtan", 8)],   # 25 local columns on the last camera
])
def test_distortion_coefficient_counts_and_nonplanar_points(built, models):
    """every n_dist the reference's templates accept, mixed in one rig, with world points off the z = 0 plane
    (Point3 has a z; calibrator.cpp:229-249 copies it): blocks and the whole solve against the oracle"""
    from multicam_calibration_b200 import synth
    from multicam_calibration_b200.rig import Rig
    base = synth.make_synthetic_rig(len(models), 14, 17, detect_prob=0.9, mixed_models=False)
    rng = np.random.default_rng(2)
    cams = []
    for (model, nd), c in zip(models, base.cameras):
        c = c.copy()
        c.distortion_model = model
        c.distortion_coeffs = rng.uniform(-2e-3, 2e-3, nd)
        cams.append(c)
    world = base.world.copy()
    world[:, 2] = rng.uniform(-0.01, 0.01, len(world))
    rig = Rig(cams, base.npts, world, base.image, base.poses)
    cal, orc = _pair(rig)
    g, o = cal.normalBlocks(True), orc.normal_blocks(True)
    for key in ("Hcc", "gc", "Hff", "gf", "Hfc"):
        assert _rel(g[key], o[key]) <= 1e-10, key
    _check_solve(cal, orc)
    cal.close()


def test_camera_that_never_saw_the_board(small_rig):
    """an active camera without a single observation: its blocks stay in the problem with zero columns
    (Ceres keeps the parameter blocks; Jacobi scale 1, LM diagonal clamped to 1e-6 / radius)"""
    from multicam_calibration_b200.rig import Rig
    off = small_rig.frame_offsets()
    keep = np.ones(small_rig.num_obs, dtype=bool)
    npts = small_rig.npts.copy()
    for f in range(small_rig.num_frames):
        o = off[f] + npts[f, 0]
        keep[o:o + npts[f, 1]] = False
    npts[:, 1] = 0
    rig = Rig(small_rig.cameras, npts, small_rig.world[keep], small_rig.image[keep], small_rig.poses)
    cal, orc = _pair(rig)
    sg, so = _check_solve(cal, orc)
    assert sg.num_reduced_params == so.num_reduced_params
    p0 = orc.get_params()
    cal.close()


@pytest.mark.parametrize("ncam", [16, 24])
def test_wide_rigs_match_oracle(built, ncam):
    """16 cameras (config c5's rig, 250 camera columns) and 24 cameras (> 256 columns): the reduced system takes the
    kernel variants that the 8-camera tests never reach (Schur rows read from global memory, Cholesky panels not
    transposed)."""
    from multicam_calibration_b200 import synth
    rig = synth.make_synthetic_rig(ncam, 24, 20260105 + ncam)
    cal, orc = _pair(rig)
    _check_solve(cal, orc)
    cal.close()


@pytest.mark.parametrize("config,nframes", [("c4", 200), ("c5", 64)])
def test_bench_configs_match_oracle(built, config, nframes):
    """the two configurations every performance number is quoted on (BASELINE.json configs[3], configs[4]), at a frame
    count the oracle solves in seconds: same generator, same seed, same cameras as bench.py's workload.  Blocks
    <= 1e-10 relative, solve at the north-star tolerances with the same iteration count."""
    from multicam_calibration_b200 import synth
    rig = (synth.config_c4 if config == "c4" else synth.config_c5)(nframes=nframes)
    assert rig.num_cameras == (8 if config == "c4" else 16)
    cal, orc = _pair(rig)
    g, o = cal.normalBlocks(True), orc.normal_blocks(True)
    assert abs(g["cost"] - o["cost"]) <= 1e-12 * o["cost"]
    for key in ("Hcc", "gc", "Hff", "gf", "Hfc"):
        assert _rel(g[key], o[key]) <= 1e-10, (key, _rel(g[key], o[key]))
    _check_solve(cal, orc)
    cal.close()


def test_rejected_steps_match_oracle(built):
    """a solve with five rejected steps in a row (tests/fixtures_util.py): radius halving, the cost-only candidate pass and the
    return to speculative iterations must follow the oracle row by row"""
    from multicam_calibration_b200 import capi
    import oracle
    from fixtures_util import rejected_steps_rig
    rig, mrd = rejected_steps_rig()
    cal, orc = _pair(rig)
    og, oo = capi.default_options(), oracle.default_options()
    og.min_relative_decrease = oo.min_relative_decrease = mrd
    sg, so = _check_solve(cal, orc, og, oo)
    assert so.num_unsuccessful == 5 and sg.num_unsuccessful == 5
    cal.close()


def test_queued_iterations_match_synchronous_path(small_rig, kat_rigs, monkeypatch):
    """solve() queues iterations ahead of their accept / reject decision (decided on the device, replayed as CUDA
    graphs); MCC_PIPELINE=0 / MCC_GRAPHS=0 take the ordinary synchronous path.  Same arithmetic, same decisions:
    costs, parameters and iteration tables are bit-identical, also on a fixture with rejected steps."""
    from multicam_calibration_b200 import capi
    for rig in (small_rig, kat_rigs["equidistant"]):
        results = []
        for env in ({}, {"MCC_GRAPHS": "0"}, {"MCC_PIPELINE": "0"}):
            for k in ("MCC_GRAPHS", "MCC_PIPELINE"):
                monkeypatch.delenv(k, raising=False)
            for k, v in env.items():
                monkeypatch.setenv(k, v)
            cal = capi.calibrator_from_rig(rig, device=0)
            s = cal.runCalibration()
            results.append((s.final_cost, s.num_iterations, s.num_successful, s.num_unsuccessful, s.why, cal.getParams().copy(), s.trace()))
            cal.close()
        ref = results[-1]
        for r in results[:-1]:
            assert r[:5] == ref[:5]
            assert np.array_equal(r[5], ref[5])
            for key in ("cost", "radius", "gradient_max", "ratio"):
                assert np.array_equal(r[6][key], ref[6][key]), key
