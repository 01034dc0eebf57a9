"""GPU: mcc_reprojection_cost_noalloc against the handle-based mcc_eval_cost and against the CPU oracle.

The no-alloc entry point exists for callers that cannot allocate device memory when they need a cost (a live
per-frame monitor that must not hit cudaMalloc's device-wide sync; a process whose GPU memory belongs to another
allocator).  It keeps its partial sums in a fixed device array inside the library, shared by all calls, so this
test makes two calls in a row -- a large one, then a much smaller one -- and checks each against the existing
entry point: leftovers of the first call must not leak into the second.  Device buffers are staged here with
torch (the test may allocate; the entry point may not).  Named test_zz_* so pytest collects it last.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12     # same bound the block-level cost check in test_gpu_parity.py uses


def _one_frame_rig(model, n, seed):
    """one camera, one frame, n observations scattered over the image, noisy"""
    from multicam_calibration_b200 import synth
    from multicam_calibration_b200.rig import Camera, Rig
    rng = np.random.default_rng(seed)
    if model == "equidistant":
        cam = Camera("equidistant", np.array([600.0, 601.0, 640.0, 512.0]), np.array([-0.0147, 0.0004, -0.004, 0.0015]))
    else:
        cam = Camera("radtan", np.array([600.0, 601.0, 640.0, 512.0]), np.array([-0.05, 0.02, 0.001, -0.0005, 0.003]))
    world = np.concatenate([rng.uniform(-0.3, 0.3, (n, 2)), rng.uniform(-0.01, 0.01, (n, 1))], axis=1)
    T = np.eye(4)
    T[:3, :3] = synth.rodrigues(np.array([0.11, -0.07, 0.2]))
    T[:3, 3] = [0.02, -0.03, 0.7]
    Xc = world @ T[:3, :3].T + T[:3, 3]
    uv = synth.project(model, Xc, cam.intrinsics, cam.distortion_coeffs) + rng.normal(0, 0.3, (n, 2))
    return Rig([cam], np.array([[n]], dtype=np.int32), world, uv, T[None])


def _existing(rig, mask):
    """the handle-based path: 1/2 sum r^2 through mcc_eval_cost(use_loss = 0)"""
    from multicam_calibration_b200 import capi
    cal = capi.calibrator_from_rig(rig, device=0)
    cal.initParams()
    if mask is not None:
        cal.setMask(mask)
    c = cal.evalCost(use_loss=False)
    cal.close()
    return c


def _oracle(rig, mask):
    """the CPU oracle's cost without loss, 1/2 sum r^2 (oracle.cpp: FrameResidual with T = double)"""
    import oracle
    o = oracle.Oracle(rig)
    if mask is not None:
        o.set_mask(mask)
    return o.cost(use_loss=False)


def _noalloc(rig, mask, check_no_alloc=False):
    import torch
    from multicam_calibration_b200 import capi
    cam = rig.cameras[0]
    dev = torch.device("cuda", 0)
    # the handle turns the 4x4 pose into (rvec, t) and back (calibrator.cpp:157-170, utils.h:29-39): do the same
    T = capi.rvec_tvec_to_transform(capi.transform_to_rvec_tvec(rig.poses[0]))
    Rt = torch.tensor(np.concatenate([T[:3, :3].reshape(9), T[:3, 3]]), dtype=torch.float64, device=dev)
    intr = torch.tensor(np.concatenate([cam.intrinsics, cam.distortion_coeffs]), dtype=torch.float64, device=dev)
    cols = [torch.tensor(np.ascontiguousarray(a), dtype=torch.float64, device=dev)
            for a in (rig.world[:, 0], rig.world[:, 1], rig.world[:, 2], rig.image[:, 0], rig.image[:, 1])]
    m = torch.tensor(mask, dtype=torch.uint8, device=dev) if mask is not None else None
    out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    capi.reprojection_cost_noalloc(cam.model_id, len(cam.distortion_coeffs), intr.data_ptr(), Rt.data_ptr(), rig.num_obs,
                                   *[c.data_ptr() for c in cols], m.data_ptr() if m is not None else None, out.data_ptr(), 0)
    torch.cuda.synchronize()
    if check_no_alloc:
        # the device's free memory is what it was: the call took nothing (only meaningful once the module is loaded,
        # i.e. not for the very first call of the process, see include/mcc.h)
        assert torch.cuda.mem_get_info(0)[0] == free_before
    return 0.5 * out.item()


@pytest.mark.parametrize("model", ["equidistant", "radtan"])
def test_noalloc_matches_eval_cost_twice_in_a_row(built, model):
    big = _one_frame_rig(model, 200_000, 1)
    small = _one_frame_rig(model, 3_000, 2)
    rng = np.random.default_rng(3)
    mask_small = (rng.random(small.num_obs) > 0.2).astype(np.uint8)
    want_big, want_small = _existing(big, None), _existing(small, mask_small)
    got_big = _noalloc(big, None)             # first call: 1563 partial sums land in the shared scratch
    got_small = _noalloc(small, mask_small, check_no_alloc=True)   # second call: 24 partial sums; the other 1539 slots are stale
    assert abs(got_big - want_big) <= REL_TOL * want_big, (got_big, want_big)
    assert abs(got_small - want_small) <= REL_TOL * want_small, (got_small, want_small)
    # ... and directly against the CPU oracle, not only against another CUDA path of this library
    for got, rig, mask in ((got_big, big, None), (got_small, small, mask_small)):
        want = _oracle(rig, mask)
        assert abs(got - want) <= REL_TOL * want, (got, want)
    assert want_small < 0.1 * want_big        # the two calls really differ, so a stale sum would show


def test_noalloc_rejects_oversize_and_handles_empty(built):
    import torch
    from multicam_calibration_b200 import capi
    dev = torch.device("cuda", 0)
    buf = torch.zeros(16, dtype=torch.float64, device=dev)
    out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    p = buf.data_ptr()
    with pytest.raises(capi.MccError) as e:
        capi.reprojection_cost_noalloc(0, 4, p, p, capi.NOALLOC_MAX_OBS + 1, p, p, p, p, p, None, out.data_ptr(), 0)
    assert e.value.status == -8 and "MCC_NOALLOC_MAX_OBS" in str(e.value)
    capi.reprojection_cost_noalloc(0, 4, p, p, 0, None, None, None, None, None, None, out.data_ptr(), 0)
    torch.cuda.synchronize()
    assert out.item() == 0.0
