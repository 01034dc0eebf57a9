"""ctypes binding of the C ABI in ``include/mcc.h`` (``libmcc.so``).

``Calibrator`` mirrors the method list of the reference's ``multicam_calibration::Calibrator``
(/root/reference/include/multicam_calibration/calibrator.h:17-49) so tests read like the reference's usage:
``setCameras``, ``addPoints``, ``runCalibration``, ``getCalibrationResults``, ``testCalibration``,
``removeOutliers``, ``setFixAllIntrinsics``.  Everything computes through the CUDA library; if it is missing
or no GPU is present the calls raise -- there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence

import numpy as np

from .rig import Camera, Rig

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MCC_LIB") or os.path.join(_HERE, "libmcc.so")   # MCC_LIB: another build of the same library (kernel experiments)

MAX_TRACE = 256
UNIQUE_ID_BYTES = 128
WHY = {0: "gradient", 1: "parameter", 2: "function", 3: "max_iterations", 4: "invalid_steps", 5: "min_radius", 6: "linear_solver", 7: "no_data"}


class MccError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"mcc status {status}: {msg}")
        self.status = status


class McCamera(C.Structure):
    _fields_ = [("model", C.c_int), ("n_intr", C.c_int), ("n_dist", C.c_int),
                ("intr", C.POINTER(C.c_double)), ("dist", C.POINTER(C.c_double)),
                ("T_cn_cnm1", C.c_double * 16), ("fix_intr", C.c_int), ("fix_extr", C.c_int), ("active", C.c_int)]


class View(C.Structure):
    """mcc_view (include/mcc.h)"""
    _fields_ = [("model", C.c_int), ("n_dist", C.c_int), ("npts", C.c_int),
                ("intr", C.POINTER(C.c_double)), ("dist", C.POINTER(C.c_double)),
                ("world_xyz", C.POINTER(C.c_double)), ("image_uv", C.POINTER(C.c_double))]


class Options(C.Structure):
    _fields_ = [("max_num_iterations", C.c_int),
                ("function_tolerance", C.c_double), ("parameter_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
                ("initial_trust_region_radius", C.c_double), ("max_trust_region_radius", C.c_double), ("min_trust_region_radius", C.c_double),
                ("min_relative_decrease", C.c_double), ("min_lm_diagonal", C.c_double), ("max_lm_diagonal", C.c_double),
                ("max_num_consecutive_invalid_steps", C.c_int), ("jacobi_scaling", C.c_int), ("use_loss", C.c_int),
                ("verbose", C.c_int), ("keep_params", C.c_int)]


class Summary(C.Structure):
    _fields_ = [("initial_cost", C.c_double), ("final_cost", C.c_double),
                ("num_iterations", C.c_int), ("num_successful", C.c_int), ("num_unsuccessful", C.c_int),
                ("termination", C.c_int), ("why", C.c_int), ("why_value", C.c_double),
                ("num_params", C.c_int), ("num_reduced_params", C.c_int), ("num_residuals", C.c_longlong),
                ("t_upload", C.c_double), ("t_jacobian", C.c_double), ("t_linear", C.c_double), ("t_residual", C.c_double), ("t_total", C.c_double),
                ("n_jacobian_evals", C.c_int), ("n_residual_evals", C.c_int), ("n_linear_solves", C.c_int),
                ("n_kernel_launches", C.c_longlong), ("h2d_bytes", C.c_longlong), ("d2h_bytes", C.c_longlong),
                ("trace_len", C.c_int),
                ("tr_cost", C.c_double * MAX_TRACE), ("tr_cost_change", C.c_double * MAX_TRACE), ("tr_gradient_max", C.c_double * MAX_TRACE),
                ("tr_step_norm", C.c_double * MAX_TRACE), ("tr_ratio", C.c_double * MAX_TRACE), ("tr_radius", C.c_double * MAX_TRACE),
                ("tr_iter_time", C.c_double * MAX_TRACE)]

    def trace(self):
        n = self.trace_len
        return {k: np.array(getattr(self, "tr_" + k)[:n]) for k in ("cost", "cost_change", "gradient_max", "step_norm", "ratio", "radius", "iter_time")}


# every symbol include/mcc.h declares (tests check the built library exports all of them)
SYMBOLS = ["mcc_version", "mcc_default_options", "mcc_summary_size", "mcc_create", "mcc_destroy", "mcc_last_error", "mcc_set_device", "mcc_set_devices",
           "mcc_set_cameras", "mcc_set_fix_all_intrinsics", "mcc_set_camera_flags", "mcc_add_frame", "mcc_add_frames", "mcc_clear_frames",
           "mcc_num_frames", "mcc_num_observations", "mcc_init_params", "mcc_num_params", "mcc_get_params", "mcc_set_params", "mcc_upload",
           "mcc_run", "mcc_get_camera", "mcc_num_residuals", "mcc_eval_residuals", "mcc_eval_cost", "mcc_remove_outliers", "mcc_get_mask",
           "mcc_set_mask", "mcc_num_reduced_cam_cols", "mcc_num_active_frames", "mcc_eval_normal_blocks", "mcc_comm_unique_id", "mcc_comm_init", "mcc_comm_transport",
           "mcc_transform_to_rvec_tvec", "mcc_rvec_tvec_to_transform", "mcc_timer_start", "mcc_timer_stop", "mcc_set_kernel_timing",
           "mcc_num_kernel_classes", "mcc_kernel_class_name", "mcc_get_kernel_times", "mcc_measure_fp64_peak", "mcc_debug_lane_map",
           "mcc_reprojection_cost_noalloc", "mcc_init_pose", "mcc_init_pose_batch"]


def build(force: bool = False) -> str:
    """compile libmcc.so in-tree for sm_100a (nvcc cross-compiles without a GPU)"""
    src_dir = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(src_dir, f) for f in os.listdir(src_dir)] + [os.path.join(_HERE, "..", "include", "mcc.h")]
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", src_dir, "-s"] + (["-B"] if force else []))
    return LIB_PATH


ASAN_LIB_PATH = os.path.join(_HERE, "libmcc_asan.so")


def build_sanitized() -> str:
    """the same sources with the host code under AddressSanitizer + UBSan (test infrastructure: tests/test_sanitizers.py)"""
    src_dir = os.path.join(_HERE, "csrc")
    subprocess.check_call(["make", "-C", src_dir, "-s", "asan"])
    return ASAN_LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MccError(-4, f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'` (no CPU fallback exists)")
    L = C.CDLL(LIB_PATH)
    dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p
    u8p = C.POINTER(C.c_uint8)
    L.mcc_version.restype = C.c_char_p
    L.mcc_last_error.restype = C.c_char_p
    L.mcc_last_error.argtypes = [vp]
    L.mcc_default_options.argtypes = [C.POINTER(Options)]
    L.mcc_create.argtypes = [C.POINTER(vp)]
    L.mcc_destroy.argtypes = [vp]
    L.mcc_destroy.restype = None
    L.mcc_set_device.argtypes = [vp, C.c_int]
    L.mcc_set_devices.argtypes = [vp, C.c_int, ip]
    L.mcc_set_cameras.argtypes = [vp, C.c_int, C.POINTER(McCamera)]
    L.mcc_set_fix_all_intrinsics.argtypes = [vp, C.c_int]
    L.mcc_set_camera_flags.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int]
    L.mcc_add_frame.argtypes = [vp, C.c_int, ip, C.POINTER(dp), C.POINTER(dp), dp]
    L.mcc_add_frames.argtypes = [vp, C.c_int, C.c_int, ip, dp, dp, dp]
    L.mcc_clear_frames.argtypes = [vp]
    for f in ("mcc_num_frames", "mcc_num_observations", "mcc_num_params", "mcc_num_residuals", "mcc_num_active_frames"):
        getattr(L, f).argtypes = [vp]
        getattr(L, f).restype = C.c_longlong
    L.mcc_init_params.argtypes = [vp]
    L.mcc_get_params.argtypes = [vp, dp, C.c_size_t]
    L.mcc_set_params.argtypes = [vp, dp, C.c_size_t]
    L.mcc_upload.argtypes = [vp]
    L.mcc_run.argtypes = [vp, C.POINTER(Options), C.POINTER(Summary)]
    L.mcc_get_camera.argtypes = [vp, C.c_int, dp, dp, dp]
    L.mcc_eval_residuals.argtypes = [vp, dp, C.c_size_t]
    L.mcc_eval_cost.argtypes = [vp, C.c_int, dp]
    L.mcc_remove_outliers.argtypes = [vp, C.c_double, C.POINTER(C.c_longlong)]
    L.mcc_get_mask.argtypes = [vp, u8p, C.c_size_t]
    L.mcc_set_mask.argtypes = [vp, u8p, C.c_size_t]
    L.mcc_num_reduced_cam_cols.argtypes = [vp]
    L.mcc_eval_normal_blocks.argtypes = [vp, C.c_int, dp, dp, dp, dp, dp, dp, ip]
    L.mcc_comm_unique_id.argtypes = [C.POINTER(C.c_ubyte)]
    L.mcc_comm_init.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_ubyte)]
    L.mcc_comm_transport.argtypes = [vp]
    L.mcc_transform_to_rvec_tvec.argtypes = [dp, dp]
    L.mcc_transform_to_rvec_tvec.restype = None
    L.mcc_rvec_tvec_to_transform.argtypes = [dp, dp]
    L.mcc_rvec_tvec_to_transform.restype = None
    L.mcc_timer_start.argtypes = [vp]
    L.mcc_timer_stop.argtypes = [vp, dp]
    L.mcc_set_kernel_timing.argtypes = [vp, C.c_int]
    L.mcc_kernel_class_name.argtypes = [C.c_int]
    L.mcc_kernel_class_name.restype = C.c_char_p
    L.mcc_get_kernel_times.argtypes = [vp, dp, C.POINTER(C.c_longlong), C.c_int]
    L.mcc_measure_fp64_peak.argtypes = [C.c_int, dp]
    L.mcc_init_pose.argtypes = [C.c_int, dp, C.c_int, dp, C.c_int, dp, dp, dp]
    L.mcc_init_pose_batch.argtypes = [C.c_int, C.POINTER(View), dp, ip, C.c_int]
    L.mcc_reprojection_cost_noalloc.argtypes = [C.c_int, C.c_int] + [vp, vp, C.c_size_t] + [vp] * 8 + [C.c_char_p, C.c_size_t]
    assert L.mcc_summary_size() == C.sizeof(Summary), (L.mcc_summary_size(), C.sizeof(Summary))
    _lib = L
    return L


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def default_options() -> Options:
    o = Options()
    lib().mcc_default_options(C.byref(o))
    return o


def comm_unique_id() -> bytes:
    buf = (C.c_ubyte * UNIQUE_ID_BYTES)()
    st = lib().mcc_comm_unique_id(buf)
    if st != 0:
        raise MccError(st, "ncclGetUniqueId failed (libnccl.so.2 not loadable?)")
    return bytes(buf)


def measure_fp64_peak(device: int = 0) -> float:
    """measured FP64 FMA throughput of the device in TFLOP/s"""
    t = C.c_double()
    st = lib().mcc_measure_fp64_peak(int(device), C.byref(t))
    if st != 0:
        raise MccError(st, "mcc_measure_fp64_peak failed")
    return t.value


NOALLOC_MAX_OBS = 4194304


def reprojection_cost_noalloc(model: int, n_dist: int, d_intr: int, d_Rt: int, n: int, d_X: int, d_Y: int, d_Z: int, d_U: int, d_V: int,
                              d_mask: Optional[int], d_sum_sq: int, stream: int = 0) -> None:
    """mcc_reprojection_cost_noalloc over raw DEVICE addresses (ints); raises MccError with the library's message.
    The library allocates nothing in this call; the caller owns every buffer."""
    err = C.create_string_buffer(512)
    st = lib().mcc_reprojection_cost_noalloc(int(model), int(n_dist), d_intr, d_Rt, int(n), d_X, d_Y, d_Z, d_U, d_V, d_mask, d_sum_sq,
                                             stream, err, len(err))
    if st != 0:
        raise MccError(st, err.value.decode())


def init_pose(camera: Camera, world_xyz: np.ndarray, image_uv: np.ndarray) -> np.ndarray:
    """get_init_pose (/root/reference/src/get_init_pose.cpp:81-140): 4x4 T_cam_world of one camera view"""
    w = np.ascontiguousarray(world_xyz, dtype=np.float64).reshape(-1, 3)
    im = np.ascontiguousarray(image_uv, dtype=np.float64).reshape(-1, 2)
    intr = np.ascontiguousarray(camera.intrinsics, dtype=np.float64)
    dist = np.ascontiguousarray(camera.distortion_coeffs, dtype=np.float64)
    T = np.zeros(16)
    st = lib().mcc_init_pose(camera.model_id, _dp(intr), len(dist), _dp(dist), len(w), _dp(w), _dp(im), _dp(T))
    if st != 0:
        raise MccError(st, "mcc_init_pose failed")
    return T.reshape(4, 4)


def init_pose_batch(views, n_threads: int = 0):
    """mcc_init_pose_batch: views = [(camera, world_points Nx3, image_points Nx2), ...] -> (T [n,4,4], status [n])"""
    n = len(views)
    arr = (View * max(n, 1))()
    keep = []
    for i, (camera, w, im) in enumerate(views):
        w = np.ascontiguousarray(w, dtype=np.float64)
        im = np.ascontiguousarray(im, dtype=np.float64)
        intr = np.ascontiguousarray(camera.intrinsics, dtype=np.float64)
        dist = np.ascontiguousarray(camera.distortion_coeffs, dtype=np.float64)
        keep.append((w, im, intr, dist))
        arr[i].model, arr[i].n_dist, arr[i].npts = camera.model_id, len(dist), len(w)
        arr[i].intr, arr[i].dist, arr[i].world_xyz, arr[i].image_uv = _dp(intr), _dp(dist), _dp(w), _dp(im)
    T = np.zeros((n, 16))
    status = np.zeros(n, dtype=np.int32)
    st = lib().mcc_init_pose_batch(n, arr, _dp(T), _ip(status), n_threads)
    if st != 0:
        raise MccError(st, "mcc_init_pose_batch failed")
    return T.reshape(n, 4, 4), status


def transform_to_rvec_tvec(T) -> np.ndarray:
    T = np.ascontiguousarray(T, dtype=np.float64).reshape(16)
    o = np.zeros(6)
    lib().mcc_transform_to_rvec_tvec(_dp(T), _dp(o))
    return o


def rvec_tvec_to_transform(v) -> np.ndarray:
    v = np.ascontiguousarray(v, dtype=np.float64)
    T = np.zeros(16)
    lib().mcc_rvec_tvec_to_transform(_dp(v), _dp(T))
    return T.reshape(4, 4)


class Calibrator:
    """Host-side mirror of the reference's Calibrator over the C ABI."""

    def __init__(self, device=None):
        """device: None (current device), an ordinal, or a list of ordinals (several GPUs behind this one handle)"""
        self._L = lib()
        self.h = C.c_void_p()
        self._check(self._L.mcc_create(C.byref(self.h)))
        if isinstance(device, (list, tuple)):
            ids = np.asarray(device, dtype=np.int32)
            self._check(self._L.mcc_set_devices(self.h, len(ids), _ip(ids)))
        elif device is not None:
            self._check(self._L.mcc_set_device(self.h, int(device)))
        self.cameras: List[Camera] = []
        self.summary: Optional[Summary] = None

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self._L.mcc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st):
        if st != 0:
            raise MccError(st, self._L.mcc_last_error(self.h).decode())

    # ---- reference API -------------------------------------------------------------------------
    def setFixAllIntrinsics(self, f: bool):
        self._check(self._L.mcc_set_fix_all_intrinsics(self.h, int(bool(f))))

    def setCameras(self, cameras: Sequence[Camera]):
        self.cameras = [c.copy() for c in cameras]
        arr = (McCamera * len(cameras))()
        keep = []
        for k, c in enumerate(cameras):
            intr = np.ascontiguousarray(c.intrinsics, dtype=np.float64)
            dist = np.ascontiguousarray(c.distortion_coeffs, dtype=np.float64)
            keep += [intr, dist]
            arr[k].model = c.model_id
            arr[k].n_intr = len(intr)
            arr[k].n_dist = len(dist)
            arr[k].intr = _dp(intr)
            arr[k].dist = _dp(dist)
            arr[k].T_cn_cnm1 = (C.c_double * 16)(*np.asarray(c.T_cn_cnm1, dtype=np.float64).reshape(16))
            arr[k].fix_intr, arr[k].fix_extr, arr[k].active = int(c.fix_intrinsics), int(c.fix_extrinsics), int(c.active)
        self._check(self._L.mcc_set_cameras(self.h, len(cameras), arr))

    def setCameraFlags(self, cam: int, fix_intrinsics: bool, fix_extrinsics: bool, active: bool):
        self._check(self._L.mcc_set_camera_flags(self.h, cam, int(fix_intrinsics), int(fix_extrinsics), int(active)))
        c = self.cameras[cam]
        c.fix_intrinsics, c.fix_extrinsics, c.active = bool(fix_intrinsics), bool(fix_extrinsics), bool(active)

    def addPoints(self, frameNum: int, wp: Sequence[np.ndarray], ip: Sequence[np.ndarray], cam0PoseGuess: np.ndarray):
        """wp[k]: [n,3] world points, ip[k]: [n,2] image points per camera (frameNum is unused, as in the reference)"""
        ncam = len(wp)
        npts = np.array([len(w) for w in wp], dtype=np.int32)
        ws = [np.ascontiguousarray(w, dtype=np.float64).reshape(-1) for w in wp]
        is_ = [np.ascontiguousarray(i, dtype=np.float64).reshape(-1) for i in ip]
        wpp = (C.POINTER(C.c_double) * ncam)(*[_dp(w) for w in ws])
        ipp = (C.POINTER(C.c_double) * ncam)(*[_dp(i) for i in is_])
        pose = np.ascontiguousarray(cam0PoseGuess, dtype=np.float64).reshape(16)
        self._check(self._L.mcc_add_frame(self.h, ncam, _ip(npts), wpp, ipp, _dp(pose)))

    def addFrames(self, rig: Rig):
        """bulk addPoints for every frame of a flat Rig"""
        self._check(self._L.mcc_add_frames(self.h, rig.num_frames, rig.num_cameras, _ip(rig.npts), _dp(rig.world), _dp(rig.image), _dp(rig.poses)))

    def clearFrames(self):
        self._check(self._L.mcc_clear_frames(self.h))

    def runCalibration(self, options: Optional[Options] = None) -> Summary:
        s = Summary()
        st = self._L.mcc_run(self.h, C.byref(options) if options is not None else None, C.byref(s))
        self.summary = s
        self._check(st)
        return s

    def getCalibrationResults(self) -> List[Camera]:
        out = []
        for k, c in enumerate(self.cameras):
            intr = np.zeros(4)
            dist = np.zeros(max(1, len(c.distortion_coeffs)))
            T = np.zeros(16)
            self._check(self._L.mcc_get_camera(self.h, k, _dp(intr), _dp(dist), _dp(T)))
            r = c.copy()
            r.intrinsics = intr
            r.distortion_coeffs = dist[:len(c.distortion_coeffs)]
            if k > 0:
                r.T_cn_cnm1 = T.reshape(4, 4)
            out.append(r)
        # T_cam_imu chain (calibrator.cpp:444-446)
        if out and np.asarray(out[0].T_cam_imu).sum() != 0:
            T_n_0 = np.eye(4)
            for k in range(1, len(out)):
                T_n_0 = out[k].T_cn_cnm1 @ T_n_0
                out[k].T_cam_imu = T_n_0 @ out[0].T_cam_imu
        return out

    def testCalibration(self) -> np.ndarray:
        """raw residuals [(n_valid), 2] in frame -> camera -> point order"""
        n = self._L.mcc_num_residuals(self.h)
        r = np.zeros(max(n, 1))
        self._check(self._L.mcc_eval_residuals(self.h, _dp(r), n))
        return r[:n].reshape(-1, 2)

    def removeOutliers(self, maxErr: float) -> int:
        nd = C.c_longlong()
        self._check(self._L.mcc_remove_outliers(self.h, float(maxErr), C.byref(nd)))
        return nd.value

    # ---- extras over the ABI -------------------------------------------------------------------
    def upload(self):
        self._check(self._L.mcc_upload(self.h))

    def initParams(self):
        self._check(self._L.mcc_init_params(self.h))

    @property
    def num_params(self) -> int:
        return self._L.mcc_num_params(self.h)

    def getParams(self) -> np.ndarray:
        p = np.zeros(self.num_params)
        self._check(self._L.mcc_get_params(self.h, _dp(p), p.size))
        return p

    def setParams(self, p):
        p = np.ascontiguousarray(p, dtype=np.float64)
        self._check(self._L.mcc_set_params(self.h, _dp(p), p.size))

    def evalCost(self, use_loss=True) -> float:
        c = C.c_double()
        self._check(self._L.mcc_eval_cost(self.h, int(use_loss), C.byref(c)))
        return c.value

    def getMask(self) -> np.ndarray:
        n = self._L.mcc_num_observations(self.h)
        m = np.zeros(max(n, 1), dtype=np.uint8)
        self._check(self._L.mcc_get_mask(self.h, m.ctypes.data_as(C.POINTER(C.c_uint8)), n))
        return m[:n]

    def setMask(self, m):
        m = np.ascontiguousarray(m, dtype=np.uint8)
        self._check(self._L.mcc_set_mask(self.h, m.ctypes.data_as(C.POINTER(C.c_uint8)), m.size))

    def normalBlocks(self, use_loss=True):
        self.upload()
        Cc = self._L.mcc_num_reduced_cam_cols(self.h)
        ns = self._L.mcc_num_active_frames(self.h)
        cost = C.c_double()
        Hcc = np.zeros((Cc, Cc)); gc = np.zeros(max(Cc, 1))
        Hff = np.zeros((max(ns, 1), 6, 6)); gf = np.zeros((max(ns, 1), 6)); Hfc = np.zeros((max(ns, 1), 6, max(Cc, 1))); ids = np.zeros(max(ns, 1), dtype=np.int32)
        self._check(self._L.mcc_eval_normal_blocks(self.h, int(use_loss), C.byref(cost), _dp(Hcc) if Cc else None, _dp(gc), _dp(Hff), _dp(gf), _dp(Hfc), _ip(ids)))
        return dict(cost=cost.value, Hcc=Hcc, gc=gc[:Cc], Hff=Hff[:ns], gf=gf[:ns], Hfc=Hfc.reshape(-1)[:ns * 6 * Cc].reshape(ns, 6, Cc), frames=ids[:ns])

    def timerStart(self):
        self._check(self._L.mcc_timer_start(self.h))

    def timerStop(self) -> float:
        """milliseconds between timerStart and now, by CUDA events on the library's stream"""
        ms = C.c_double()
        self._check(self._L.mcc_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def setKernelTiming(self, enabled: bool):
        self._check(self._L.mcc_set_kernel_timing(self.h, int(enabled)))

    def kernelTimes(self) -> dict:
        n = self._L.mcc_num_kernel_classes()
        ms = np.zeros(n)
        cnt = np.zeros(n, dtype=np.int64)
        self._check(self._L.mcc_get_kernel_times(self.h, _dp(ms), cnt.ctypes.data_as(C.POINTER(C.c_longlong)), n))
        return {self._L.mcc_kernel_class_name(i).decode(): dict(ms=float(ms[i]), launches=int(cnt[i])) for i in range(n)}

    def commInit(self, nranks: int, rank: int, unique_id: bytes):
        buf = (C.c_ubyte * UNIQUE_ID_BYTES).from_buffer_copy(unique_id)
        self._check(self._L.mcc_comm_init(self.h, nranks, rank, buf))

    def commTransport(self) -> int:
        """0: single rank, 1: peer windows, 2: NCCL (valid after the first upload / run)"""
        return int(self._L.mcc_comm_transport(self.h))


def calibrator_from_rig(rig: Rig, device=None) -> Calibrator:
    cal = Calibrator(device)
    cal.setCameras(rig.cameras)
    cal.setFixAllIntrinsics(rig.fix_all_intrinsics)
    cal.addFrames(rig)
    return cal
