"""Input generators: the reference's two test fixtures and synthetic multi-camera rigs.

* ``make_test_corners`` restates /root/reference/src/make_test_corners.py (deterministic 20x20 grid seen from
  9 x 10 x 10 vantage points) so the ``corners.csv`` blobs missing from /root/reference/test/*/ can be rebuilt.
* ``init_pose_homography`` restates /root/reference/src/get_init_pose.cpp:46-140 with Python cv2 (used only by
  tests/golden/make_golden.py to produce the committed initial-pose fixtures).
* ``make_synthetic_rig`` builds aprilgrid rigs "in the spirit of make_test_corners.py" for the
  README-scale (c3), 8-camera (c4) and 16-camera (c5) configurations of BASELINE.json.

Nothing here is on the timed path; it only produces inputs.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence

import numpy as np

from .rig import Camera, Rig

# ----------------------------------------------------------------------------------------------------
# forward camera models (numpy, float64) -- generator-side truth
# ----------------------------------------------------------------------------------------------------


def rodrigues(rvec: np.ndarray) -> np.ndarray:
    """rotation vector(s) [...,3] -> matrices [...,3,3]"""
    rvec = np.asarray(rvec, dtype=np.float64)
    th = np.linalg.norm(rvec, axis=-1)
    small = th < 1e-12
    ths = np.where(small, 1.0, th)
    a = rvec / ths[..., None]
    K = np.zeros(rvec.shape[:-1] + (3, 3))
    K[..., 0, 1], K[..., 0, 2] = -a[..., 2], a[..., 1]
    K[..., 1, 0], K[..., 1, 2] = a[..., 2], -a[..., 0]
    K[..., 2, 0], K[..., 2, 1] = -a[..., 1], a[..., 0]
    s, c = np.sin(th)[..., None, None], np.cos(th)[..., None, None]
    I = np.broadcast_to(np.eye(3), K.shape)
    R = c * I + (1 - c) * a[..., :, None] * a[..., None, :] + s * K
    R = np.where(small[..., None, None], I, R)
    return R


def rot_to_rvec(R: np.ndarray) -> np.ndarray:
    """single 3x3 rotation -> rotation vector with angle in [0, pi]"""
    R = np.asarray(R, dtype=np.float64)
    c = (np.trace(R) - 1.0) / 2.0
    c = min(1.0, max(-1.0, c))
    th = math.acos(c)
    if th < 1e-12:
        return np.zeros(3)
    if math.pi - th < 1e-6:
        # near pi: axis from the symmetric part
        A = (R + np.eye(3)) / 2.0
        i = int(np.argmax(np.diag(A)))
        ax = A[:, i] / math.sqrt(A[i, i])
        w = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
        if np.dot(w, ax) < 0:
            ax = -ax
        return ax * th
    w = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    return w / (2.0 * math.sin(th)) * th


def project_equidistant(Xc: np.ndarray, K4: np.ndarray, D: np.ndarray) -> np.ndarray:
    x, y, z = Xc[..., 0], Xc[..., 1], Xc[..., 2]
    h = np.hypot(x, y)
    th = np.arctan2(h, z)
    th2 = th * th
    fac = np.ones_like(th)
    p = th2.copy()
    for k in np.atleast_1d(D):
        fac = fac + k * p
        p = p * th2
    r = th * fac
    hs = np.where(h > 0, h, 1.0)
    c = np.where(h > 0, x / hs, 1.0)
    s = np.where(h > 0, y / hs, 0.0)
    return np.stack([K4[0] * r * c + K4[2], K4[1] * r * s + K4[3]], axis=-1)


def project_radtan(Xc: np.ndarray, K4: np.ndarray, D: np.ndarray) -> np.ndarray:
    d = np.zeros(8)
    d[:len(D)] = D
    x = Xc[..., 0] / Xc[..., 2]
    y = Xc[..., 1] / Xc[..., 2]
    x2, y2, xy = x * x, y * y, x * y
    r2 = x2 + y2
    r4 = r2 * r2
    r6 = r2 * r4
    ratio = (1 + d[0] * r2 + d[1] * r4 + d[4] * r6) / (1 + d[5] * r2 + d[6] * r4 + d[7] * r6)
    xpp = x * ratio + 2 * d[2] * xy + d[3] * (r2 + 2 * x2)
    ypp = y * ratio + d[2] * (r2 + 2 * y2) + 2 * d[3] * xy
    return np.stack([K4[0] * xpp + K4[2], K4[1] * ypp + K4[3]], axis=-1)


def project(model: str, Xc, K4, D):
    return project_equidistant(Xc, K4, D) if model == "equidistant" else project_radtan(Xc, K4, D)


# ----------------------------------------------------------------------------------------------------
# the reference's test fixtures (test/equidistant, test/radtan)
# ----------------------------------------------------------------------------------------------------

KAT_CAMERAS = {
    # /root/reference/test/equidistant/equidistant-initial.yaml:1-7
    "equidistant": dict(intrinsics=[151.0, 152.0, 173.0, 144.0], distortion_coeffs=[0.155, -0.18, -0.005, -0.0003],
                        resolution=(352, 287)),
    # /root/reference/test/radtan/radtan-initial.yaml:1-7
    "radtan": dict(intrinsics=[151.0, 152.0, 173.0, 144.0],
                   distortion_coeffs=[0.155, -0.18, -0.005, -0.0003, 0.044, 0.005, 0.004, -0.0004], resolution=(352, 287)),
}


def make_test_corners(model: str):
    """Restates make_test_corners.py.  Returns (npts[900], world_xy[N,2], image_uv[N,2]) with the values as
    they read back from the ``%f`` csv through the reference's ``float`` sscanf
    (/root/reference/src/calibration_nodelet.cpp:564-567): rounded to 6 decimals, then to float32."""
    cam = KAT_CAMERAS[model]
    K4 = np.array(cam["intrinsics"])
    D = np.array(cam["distortion_coeffs"])
    res = cam["resolution"]
    n, L = 20, 2.0
    d = L / n
    iy, ix = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    wp = np.stack([(ix.ravel() - n // 2) * d, (iy.ravel() - n // 2) * d, np.zeros(n * n)], axis=1)  # iy outer, ix inner
    Rbase = np.diag([1.0, -1.0, -1.0])
    npts, wl, il = [], [], []
    for dist in np.arange(1.0, 10.0, 1):
        for theta in np.arange(-0.1, 0.1, 0.02):
            for phi in np.arange(-0.1, 0.1, 0.02):
                ct, st, cp, sp = math.cos(theta), math.sin(theta), math.cos(phi), math.sin(phi)
                Rx = np.array([[1, 0, 0], [0, ct, -st], [0, st, ct]])
                Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
                R = Rx @ Ry @ Rbase            # euler_matrix(theta, phi, 0, 'rxyz') * Rbase
                # the script converts to angle/axis and back (both projectors rebuild R from rvec)
                R = rodrigues(rot_to_rvec(R))
                Xc = wp @ R.T + np.array([0.0, 0.0, dist])
                uv = project(model, Xc, K4, D)
                keep = (uv[:, 0] > 0) & (uv[:, 0] < res[0]) & (uv[:, 1] > 0) & (uv[:, 1] < res[1])
                npts.append(int(keep.sum()))
                wl.append(wp[keep, :2])
                il.append(uv[keep])
    w = np.concatenate(wl)
    im = np.concatenate(il)
    # '%f' -> 6 decimals; sscanf("%f") into float -> float32
    w = np.round(w, 6).astype(np.float32).astype(np.float64)
    im = np.round(im, 6).astype(np.float32).astype(np.float64)
    return np.asarray(npts, dtype=np.int32), w, im


def init_pose_homography(world_xy: np.ndarray, image_uv: np.ndarray, K4, D, model: str) -> np.ndarray:
    """get_init_pose.cpp:81-140 with Python cv2: undistort -> findHomography -> decompose. Returns 4x4."""
    import cv2
    K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1.0]])
    wp = np.asarray(world_xy, dtype=np.float32).reshape(-1, 1, 2)
    ip = np.asarray(image_uv, dtype=np.float32).reshape(-1, 1, 2)
    dist = np.asarray(D, dtype=np.float64).reshape(-1, 1)
    if model == "equidistant":
        ipu = cv2.fisheye.undistortPoints(ip, K, dist[:4])
    else:
        ipu = cv2.undistortPoints(ip, K, dist)
    H = cv2.findHomography(wp, ipu)[0]
    if H[2, 2] < 0:
        H = -H
    sv = np.linalg.svd(H, compute_uv=False)
    H = H / sv[1]                      # SL.at<double>(0,1) on a 3x1 Mat == second singular value
    r1 = H[:, 0] / np.linalg.norm(H[:, 0])
    r2 = H[:, 1] - np.dot(H[:, 1], r1) * r1
    r2 = r2 / np.linalg.norm(r2)
    r3 = np.cross(r1, r2)
    T = np.eye(4)
    T[:3, 0], T[:3, 1], T[:3, 2], T[:3, 3] = r1, r2, r3, H[:, 2]
    return T


def kat_rig(model: str, poses: np.ndarray) -> Rig:
    """The 1-camera, 900-frame rig of /root/reference/test/<model>/ with the given initial poses [900,4,4]."""
    npts, w, im = make_test_corners(model)
    cam = KAT_CAMERAS[model]
    c = Camera(distortion_model=model, intrinsics=np.array(cam["intrinsics"]), distortion_coeffs=np.array(cam["distortion_coeffs"]),
               resolution=cam["resolution"], name="cam0", rostopic="/cam/image_raw")
    world = np.concatenate([w, np.zeros((len(w), 1))], axis=1)
    return Rig([c], npts.reshape(-1, 1), world, im, poses)


# ----------------------------------------------------------------------------------------------------
# synthetic aprilgrid rigs
# ----------------------------------------------------------------------------------------------------


def aprilgrid_corners(rows: int = 5, cols: int = 7, size: float = 0.04, spacing: float = 0.25) -> np.ndarray:
    """[rows*cols*4, 2] board coordinates, tag-major, corner order of
    /root/reference/src/multicam_apriltag_detector.cpp:191-208 (defaults from config/aprilgrid.yaml)."""
    out = []
    for tag in range(rows * cols):
        col, row = tag % cols, tag // cols
        x0, y0 = col * size * (1 + spacing), row * size * (1 + spacing)
        out += [(x0, y0), (x0 + size, y0), (x0 + size, y0 + size), (x0, y0 + size)]
    return np.asarray(out, dtype=np.float64)


def _euler_xyz(ax, ay, az):
    """Rz(az) @ Ry(ay) @ Rx(ax), vectorised over leading dims"""
    ax, ay, az = np.broadcast_arrays(ax, ay, az)
    cx, sx, cy, sy, cz, sz = np.cos(ax), np.sin(ax), np.cos(ay), np.sin(ay), np.cos(az), np.sin(az)
    R = np.empty(ax.shape + (3, 3))
    R[..., 0, 0] = cz * cy; R[..., 0, 1] = cz * sy * sx - sz * cx; R[..., 0, 2] = cz * sy * cx + sz * sx
    R[..., 1, 0] = sz * cy; R[..., 1, 1] = sz * sy * sx + cz * cx; R[..., 1, 2] = sz * sy * cx - cz * sx
    R[..., 2, 0] = -sy;     R[..., 2, 1] = cy * sx;                R[..., 2, 2] = cy * cx
    return R


README_CAMS = [
    # /root/reference/README.md:35-51
    dict(intrinsics=[605.054, 604.66, 641.791, 508.728], dist=[-0.0146915, 0.000370396, -0.00425216, 0.0015107]),
    dict(intrinsics=[605.097, 604.321, 698.772, 573.558], dist=[-0.0146155, -0.00291494, -0.000681981, 0.000221855]),
]
README_T_C1_C0 = np.array([[0.99999965648, -0.00013331925, 0.00081808159, -0.19946344647],
                           [0.00013388601, 0.99999975107, -0.00069277676, -0.00005674605],
                           [-0.00081798903, 0.00069288605, 0.99999942540, 0.00010022941],
                           [0.0, 0.0, 0.0, 1.0]])


def make_synthetic_rig(ncam: int, nframes: int, seed: int, *, readme_stereo: bool = False, detect_prob: float = 0.893,
                       noise_px: float = 0.2, mixed_models: bool = True, resolution=(1280, 1024),
                       perturb: bool = True, chunk: int = 2000, pose_seed: Optional[int] = None) -> Rig:
    """Synthetic aprilgrid rig.  Cameras (truth + perturbed initial values) depend only on ``seed``; the frames
    (board poses, detections, noise) depend on ``pose_seed`` (default: ``seed``) so that several ranks can own
    different frames of the SAME rig.  Observations are rounded to float32 like the csv path.
    ``rig.truth`` holds the generating parameters."""
    rng_c = np.random.default_rng(seed)
    rng = np.random.default_rng(seed if pose_seed is None else [seed, pose_seed])
    board = aprilgrid_corners()
    ntag = len(board) // 4
    board3 = np.concatenate([board, np.zeros((len(board), 1))], axis=1)
    bc = board.mean(axis=0)

    # ---- cameras: truth
    cams_true: List[Camera] = []
    T_c_c0 = [np.eye(4)]
    centres = [np.zeros(3)]
    if readme_stereo:
        assert ncam == 2
        for k in range(2):
            cams_true.append(Camera("equidistant", np.array(README_CAMS[k]["intrinsics"]), np.array(README_CAMS[k]["dist"]),
                                    resolution=resolution, name=f"cam{k}", rostopic=f"/rig/cam{k}/image_mono"))
        T_c_c0.append(README_T_C1_C0.copy())
        centres.append(np.array([0.2, 0.0, 0.0]))
    else:
        ncol = 4
        for k in range(ncam):
            model = "radtan" if (mixed_models and k % 2 == 1) else "equidistant"
            f = 600.0 + rng_c.uniform(-10, 10)
            intr = np.array([f, f + rng_c.uniform(-2, 2), resolution[0] / 2 + rng_c.uniform(-20, 20), resolution[1] / 2 + rng_c.uniform(-20, 20)])
            if model == "equidistant":
                dist = np.array([-0.0147, 0.0004, -0.004, 0.0015]) * (1 + rng_c.uniform(-0.2, 0.2, 4))
            else:
                dist = np.array([-0.05, 0.02, 0.001, -0.0005]) * (1 + rng_c.uniform(-0.2, 0.2, 4))
            cams_true.append(Camera(model, intr, dist, resolution=resolution, name=f"cam{k}", rostopic=f"/rig/cam{k}/image_raw"))
            if k > 0:
                # camera centre on a lattice with 5 cm pitch, small relative rotation (never exact identity)
                cpos = np.array([(k % ncol) * 0.05, (k // ncol) * 0.05, 0.0])
                ang = np.deg2rad(rng_c.uniform(-3, 3, 3))
                Rk = _euler_xyz(ang[0], ang[1], ang[2])
                T = np.eye(4)
                T[:3, :3] = Rk
                T[:3, 3] = -Rk @ cpos
                T_c_c0.append(T)
                centres.append(cpos)
    for k in range(ncam):
        cams_true[k].T_cn_cnm1 = np.eye(4) if k == 0 else T_c_c0[k] @ np.linalg.inv(T_c_c0[k - 1])

    # ---- cameras: initial values
    cams_init = [c.copy() for c in cams_true]
    T_init_c_c0 = [np.eye(4)]
    if perturb:
        for k, c in enumerate(cams_init):
            c.intrinsics = c.intrinsics * (1 + rng_c.uniform(-5e-3, 5e-3, 4))
            c.distortion_coeffs = c.distortion_coeffs + rng_c.uniform(-1e-3, 1e-3, len(c.distortion_coeffs))
            if k > 0:
                ang = np.deg2rad(rng_c.uniform(-0.3, 0.3, 3))
                dT = np.eye(4)
                dT[:3, :3] = _euler_xyz(ang[0], ang[1], ang[2])
                dT[:3, 3] = rng_c.uniform(-5e-3, 5e-3, 3)
                T_init_c_c0.append(dT @ T_c_c0[k])
        for k in range(1, ncam):
            cams_init[k].T_cn_cnm1 = T_init_c_c0[k] @ np.linalg.inv(T_init_c_c0[k - 1])

    # ---- frames
    npts_l, world_l, image_l, poses_true_l, poses_init_l = [], [], [], [], []
    for f0 in range(0, nframes, chunk):
        nf = min(chunk, nframes - f0)
        dist = rng.uniform(0.25, 0.6, nf)
        tx, ty = rng.uniform(-0.5, 0.5, nf), rng.uniform(-0.5, 0.5, nf)
        roll = rng.uniform(-math.pi, math.pi, nf) * 0.1
        lat = rng.uniform(-0.1, 0.1, (nf, 2))
        Rf = _euler_xyz(tx, ty, roll)                                   # [nf,3,3]
        # board centre placed at (lat, dist) in front of the rig centre
        rig_c = np.mean(np.asarray(centres), axis=0)
        centre = np.concatenate([lat, dist[:, None]], axis=1) + rig_c
        tf = centre - np.einsum("fij,j->fi", Rf, np.array([bc[0], bc[1], 0.0]))
        Y = np.einsum("fij,pj->fpi", Rf, board3) + tf[:, None, :]       # [nf, P, 3] in cam0
        det = rng.random((nf, ncam, ntag)) < detect_prob
        noise = rng.normal(0.0, noise_px, (nf, ncam, len(board), 2)) if noise_px > 0 else np.zeros((nf, ncam, len(board), 2))
        keep_all = np.zeros((nf, ncam, len(board)), dtype=bool)
        uv_all = np.zeros((nf, ncam, len(board), 2))
        for k in range(ncam):
            T = T_c_c0[k]
            Xc = Y @ T[:3, :3].T + T[:3, 3]
            uv = project(cams_true[k].distortion_model, Xc, cams_true[k].intrinsics, cams_true[k].distortion_coeffs)
            inside = (uv[..., 0] > 0) & (uv[..., 0] < resolution[0]) & (uv[..., 1] > 0) & (uv[..., 1] < resolution[1]) & (Xc[..., 2] > 0.05)
            tag_ok = inside.reshape(nf, ntag, 4).all(axis=2) & det[:, k, :]
            keep_all[:, k, :] = np.repeat(tag_ok, 4, axis=1)
            uv_all[:, k] = uv + noise[:, k]
        npts_l.append(keep_all.sum(axis=2).astype(np.int32))
        flat = keep_all.reshape(-1)
        image_l.append(uv_all.reshape(-1, 2)[flat].astype(np.float32).astype(np.float64))
        world_l.append(np.broadcast_to(board3, (nf, ncam, len(board), 3)).reshape(-1, 3)[flat].astype(np.float32).astype(np.float64))
        Tt = np.zeros((nf, 4, 4)); Tt[:, :3, :3] = Rf; Tt[:, :3, 3] = tf; Tt[:, 3, 3] = 1.0
        poses_true_l.append(Tt)
        if perturb:
            ang = np.deg2rad(rng.uniform(-1, 1, (nf, 3)))
            dR = _euler_xyz(ang[:, 0], ang[:, 1], ang[:, 2])
            Ti = np.zeros((nf, 4, 4)); Ti[:, :3, :3] = dR @ Rf; Ti[:, :3, 3] = tf + rng.uniform(-0.01, 0.01, (nf, 3)); Ti[:, 3, 3] = 1.0
            poses_init_l.append(Ti)
        else:
            poses_init_l.append(Tt.copy())
    rig = Rig(cams_init, np.concatenate(npts_l), np.concatenate(world_l), np.concatenate(image_l), np.concatenate(poses_init_l))
    rig.truth = dict(cameras=cams_true, poses=np.concatenate(poses_true_l), T_c_c0=T_c_c0)
    return rig


# named configurations of BASELINE.json (c3..c5)
def config_c3() -> Rig:
    return make_synthetic_rig(2, 409, 20260103, readme_stereo=True, detect_prob=0.88)


def config_c4(nframes: int = 10000, pose_seed: Optional[int] = None) -> Rig:
    return make_synthetic_rig(8, nframes, 20260104, pose_seed=pose_seed)


def config_c5(nframes: int = 100000, pose_seed: Optional[int] = None) -> Rig:
    return make_synthetic_rig(16, nframes, 20260105, pose_seed=pose_seed)
