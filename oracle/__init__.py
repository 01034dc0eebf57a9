"""ctypes loader for the CPU oracle (oracle/oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs -- never by the product package ``multicam_calibration_b200``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "liboracle.so")

WHY = {0: "gradient", 1: "parameter", 2: "function", 3: "max_iterations", 4: "invalid_steps", 5: "min_radius", 6: "linear_solver", 7: "no_data"}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


class Options(C.Structure):
    _fields_ = [("max_num_iterations", C.c_int),
                ("function_tolerance", C.c_double), ("parameter_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
                ("initial_trust_region_radius", C.c_double), ("max_trust_region_radius", C.c_double), ("min_trust_region_radius", C.c_double),
                ("min_relative_decrease", C.c_double), ("min_lm_diagonal", C.c_double), ("max_lm_diagonal", C.c_double),
                ("max_num_consecutive_invalid_steps", C.c_int), ("jacobi_scaling", C.c_int), ("num_threads", C.c_int),
                ("use_loss", C.c_int), ("autodiff_all_params", C.c_int), ("verbose", C.c_int)]


MAX_TRACE = 256


class Summary(C.Structure):
    _fields_ = [("initial_cost", C.c_double), ("final_cost", C.c_double),
                ("num_iterations", C.c_int), ("num_successful", C.c_int), ("num_unsuccessful", C.c_int),
                ("termination", C.c_int), ("why", C.c_int), ("why_value", C.c_double),
                ("num_params", C.c_int), ("num_reduced_params", C.c_int), ("num_residuals", C.c_longlong),
                ("t_jacobian", C.c_double), ("t_linear", C.c_double), ("t_residual", C.c_double), ("t_total", C.c_double),
                ("n_jacobian_evals", C.c_int), ("n_residual_evals", C.c_int), ("n_linear_solves", C.c_int),
                ("trace_len", C.c_int),
                ("tr_cost", C.c_double * MAX_TRACE), ("tr_cost_change", C.c_double * MAX_TRACE), ("tr_gradient_max", C.c_double * MAX_TRACE),
                ("tr_step_norm", C.c_double * MAX_TRACE), ("tr_ratio", C.c_double * MAX_TRACE), ("tr_radius", C.c_double * MAX_TRACE)]

    def trace(self):
        n = self.trace_len
        return {k: np.array(getattr(self, "tr_" + k)[:n]) for k in ("cost", "cost_change", "gradient_max", "step_norm", "ratio", "radius")}


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p
        L.orc_create.restype = vp
        L.orc_create.argtypes = [C.c_int, ip, ip, dp, dp, dp, ip, ip, ip, C.c_int]
        L.orc_destroy.argtypes = [vp]
        L.orc_add_frame.argtypes = [vp, ip, dp, dp, dp]
        L.orc_set_flags.argtypes = [vp, ip, ip, ip, C.c_int]
        L.orc_init_params.argtypes = [vp]
        L.orc_num_params.argtypes = [vp]
        L.orc_num_cam_params.argtypes = [vp]
        L.orc_set_params.argtypes = [vp, dp]
        L.orc_get_params.argtypes = [vp, dp]
        L.orc_num_residuals.argtypes = [vp]; L.orc_num_residuals.restype = C.c_longlong
        L.orc_num_reduced_cam_cols.argtypes = [vp]
        L.orc_reduced_cam_cols.argtypes = [vp, ip]
        L.orc_get_mask.argtypes = [vp, C.POINTER(C.c_uint8)]
        L.orc_set_mask.argtypes = [vp, C.POINTER(C.c_uint8)]
        L.orc_eval_residuals.argtypes = [vp, dp]
        L.orc_cost.argtypes = [vp, C.c_int, dp]; L.orc_cost.restype = C.c_double
        L.orc_frame_jacobian.argtypes = [vp, C.c_int, dp, dp]
        L.orc_frame_num_valid.argtypes = [vp, C.c_int]; L.orc_frame_num_valid.restype = C.c_longlong
        L.orc_normal_blocks.argtypes = [vp, C.c_int, C.c_int, dp, dp, dp, dp, dp, dp, ip]
        L.orc_solve.argtypes = [vp, C.POINTER(Options), C.POINTER(Summary)]
        L.orc_remove_outliers.argtypes = [vp, C.c_double]; L.orc_remove_outliers.restype = C.c_longlong
        L.orc_get_camera.argtypes = [vp, C.c_int, dp, dp, dp]
        L.orc_transform_to_rvec_tvec.argtypes = [dp, dp]
        L.orc_rvec_tvec_to_transform.argtypes = [dp, dp]
        L.orc_rotation_matrix.argtypes = [dp, dp]
        L.orc_default_options.argtypes = [C.POINTER(Options)]
        assert L.orc_summary_size() == C.sizeof(Summary), (L.orc_summary_size(), C.sizeof(Summary))
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def default_options() -> Options:
    o = Options()
    lib().orc_default_options(C.byref(o))
    return o


class Oracle:
    """CPU oracle over a ``multicam_calibration_b200.rig.Rig``-shaped problem."""

    def __init__(self, rig):
        L = lib()
        self.rig = rig
        ncam = rig.num_cameras
        model = np.array([c.model_id for c in rig.cameras], dtype=np.int32)
        ndist = np.array([len(c.distortion_coeffs) for c in rig.cameras], dtype=np.int32)
        intr = np.zeros((ncam, 4)); dist = np.zeros((ncam, 8)); T = np.zeros((ncam, 16))
        for k, c in enumerate(rig.cameras):
            intr[k] = c.intrinsics
            dist[k, :ndist[k]] = c.distortion_coeffs
            T[k] = np.asarray(c.T_cn_cnm1, dtype=np.float64).reshape(16)
        fi = np.array([c.fix_intrinsics for c in rig.cameras], dtype=np.int32)
        fe = np.array([c.fix_extrinsics for c in rig.cameras], dtype=np.int32)
        ac = np.array([c.active for c in rig.cameras], dtype=np.int32)
        self.h = L.orc_create(ncam, _ip(model), _ip(ndist), _dp(intr), _dp(dist), _dp(T), _ip(fi), _ip(fe), _ip(ac), int(rig.fix_all_intrinsics))
        off = rig.frame_offsets()
        for f in range(rig.num_frames):
            w = np.ascontiguousarray(rig.world[off[f]:off[f + 1]])
            im = np.ascontiguousarray(rig.image[off[f]:off[f + 1]])
            n = np.ascontiguousarray(rig.npts[f])
            pose = np.ascontiguousarray(rig.poses[f].reshape(16))
            L.orc_add_frame(self.h, _ip(n), _dp(w), _dp(im), _dp(pose))
        L.orc_init_params(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(self.h)
            self.h = None

    # -- parameters
    @property
    def num_params(self):
        return lib().orc_num_params(self.h)

    @property
    def num_cam_params(self):
        return lib().orc_num_cam_params(self.h)

    def init_params(self):
        lib().orc_init_params(self.h)

    def get_params(self) -> np.ndarray:
        p = np.zeros(self.num_params)
        lib().orc_get_params(self.h, _dp(p))
        return p

    def set_params(self, p):
        p = np.ascontiguousarray(p, dtype=np.float64)
        assert p.size == self.num_params
        lib().orc_set_params(self.h, _dp(p))

    def set_flags(self, fix_intr, fix_extr, active, fix_all=False):
        a = [np.ascontiguousarray(v, dtype=np.int32) for v in (fix_intr, fix_extr, active)]
        lib().orc_set_flags(self.h, _ip(a[0]), _ip(a[1]), _ip(a[2]), int(fix_all))

    def reduced_cam_cols(self) -> np.ndarray:
        c = np.zeros(self.num_cam_params, dtype=np.int32)
        lib().orc_reduced_cam_cols(self.h, _ip(c))
        return c

    # -- evaluation
    @property
    def num_residuals(self):
        return lib().orc_num_residuals(self.h)

    def residuals(self) -> np.ndarray:
        r = np.zeros(self.num_residuals)
        lib().orc_eval_residuals(self.h, _dp(r))
        return r

    def cost(self, use_loss=True, per_frame=False):
        pf = np.zeros(self.rig.num_frames)
        c = lib().orc_cost(self.h, int(use_loss), _dp(pf))
        return (c, pf) if per_frame else c

    def frame_jacobian(self, f):
        nv = lib().orc_frame_num_valid(self.h, f)
        W = self.num_cam_params + 6
        J = np.zeros((2 * nv, W)); r = np.zeros(2 * nv)
        if nv:
            lib().orc_frame_jacobian(self.h, f, _dp(J), _dp(r))
        return J, r

    def normal_blocks(self, use_loss=True, num_threads=8):
        Cc = lib().orc_num_reduced_cam_cols(self.h)
        F = self.rig.num_frames
        cost = C.c_double()
        Hcc = np.zeros((Cc, Cc)); gc = np.zeros(Cc)
        Hff = np.zeros((F, 6, 6)); gf = np.zeros((F, 6)); Hfc = np.zeros((F, 6, Cc)); ids = np.zeros(F, dtype=np.int32)
        nf = lib().orc_normal_blocks(self.h, int(use_loss), num_threads, C.byref(cost), _dp(Hcc), _dp(gc), _dp(Hff), _dp(gf), _dp(Hfc), _ip(ids))
        return dict(cost=cost.value, Hcc=Hcc, gc=gc, Hff=Hff[:nf], gf=gf[:nf], Hfc=Hfc.reshape(-1)[:nf * 6 * Cc].reshape(nf, 6, Cc), frames=ids[:nf])

    def solve(self, options: Optional[Options] = None) -> Summary:
        s = Summary()
        lib().orc_solve(self.h, C.byref(options) if options is not None else None, C.byref(s))
        return s

    def remove_outliers(self, max_err) -> int:
        return lib().orc_remove_outliers(self.h, float(max_err))

    def get_mask(self):
        m = np.zeros(self.rig.num_obs, dtype=np.uint8)
        lib().orc_get_mask(self.h, m.ctypes.data_as(C.POINTER(C.c_uint8)))
        return m

    def set_mask(self, m):
        m = np.ascontiguousarray(m, dtype=np.uint8)
        lib().orc_set_mask(self.h, m.ctypes.data_as(C.POINTER(C.c_uint8)))

    def get_camera(self, k):
        nd = len(self.rig.cameras[k].distortion_coeffs)
        intr = np.zeros(4); dist = np.zeros(max(nd, 1)); T = np.zeros(16)
        lib().orc_get_camera(self.h, k, _dp(intr), _dp(dist), _dp(T))
        return intr, dist[:nd], T.reshape(4, 4)


def transform_to_rvec_tvec(T) -> np.ndarray:
    T = np.ascontiguousarray(T, dtype=np.float64).reshape(16)
    o = np.zeros(6)
    lib().orc_transform_to_rvec_tvec(_dp(T), _dp(o))
    return o


def rvec_tvec_to_transform(v) -> np.ndarray:
    v = np.ascontiguousarray(v, dtype=np.float64)
    T = np.zeros(16)
    lib().orc_rvec_tvec_to_transform(_dp(v), _dp(T))
    return T.reshape(4, 4)
