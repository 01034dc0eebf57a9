"""Flat, frame-major description of a calibration problem.

This is the host-side container the Python binding feeds to the C ABI (``include/mcc.h``).  It holds
what the reference's ``Calibrator`` accumulates through ``setCameras`` / ``addPoints``
(/root/reference/src/calibrator.cpp:225-249): per-camera ``CalibrationData`` and, per frame and camera,
world points (x, y, z), image points (u, v) and the cam0 pose guess ``T_c0_w`` -- but stored as flat
arrays (observations sorted frame-major, camera-minor) instead of vectors of vectors.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

MODEL_IDS = {"equidistant": 0, "radtan": 1, "plumb_bob": 1}
MODEL_NAMES = {0: "equidistant", 1: "radtan"}


@dataclass
class Camera:
    """Mirror of ``CalibrationData`` (/root/reference/include/multicam_calibration/calibration_data.h:20-36)."""
    distortion_model: str = "equidistant"
    intrinsics: np.ndarray = field(default_factory=lambda: np.zeros(4))
    distortion_coeffs: np.ndarray = field(default_factory=lambda: np.zeros(4))
    T_cn_cnm1: np.ndarray = field(default_factory=lambda: np.eye(4))
    T_cam_imu: np.ndarray = field(default_factory=lambda: np.zeros((4, 4)))
    resolution: tuple = (0, 0)
    camera_model: str = "pinhole"
    name: str = ""
    rostopic: str = ""
    fix_intrinsics: bool = False
    fix_extrinsics: bool = False
    active: bool = True

    @property
    def model_id(self) -> int:
        return MODEL_IDS[self.distortion_model]

    def copy(self) -> "Camera":
        return Camera(self.distortion_model, np.array(self.intrinsics, dtype=np.float64),
                      np.array(self.distortion_coeffs, dtype=np.float64), np.array(self.T_cn_cnm1, dtype=np.float64),
                      np.array(self.T_cam_imu, dtype=np.float64), tuple(self.resolution), self.camera_model,
                      self.name, self.rostopic, self.fix_intrinsics, self.fix_extrinsics, self.active)


@dataclass
class Rig:
    """cameras + observations.

    npts   int32  [F, ncam]   number of corners camera k saw in frame f
    world  float64 [N, 3]     board coordinates, frame-major / camera-minor / insertion order
    image  float64 [N, 2]     detected pixel coordinates, same order
    poses  float64 [F, 4, 4]  initial cam0 pose guess T_c0_w per frame
    """
    cameras: List[Camera]
    npts: np.ndarray
    world: np.ndarray
    image: np.ndarray
    poses: np.ndarray
    fix_all_intrinsics: bool = False
    truth: Optional[dict] = None  # generator ground truth, if synthetic

    def __post_init__(self):
        self.npts = np.ascontiguousarray(self.npts, dtype=np.int32)
        self.world = np.ascontiguousarray(self.world, dtype=np.float64)
        self.image = np.ascontiguousarray(self.image, dtype=np.float64)
        self.poses = np.ascontiguousarray(self.poses, dtype=np.float64)
        assert self.npts.ndim == 2 and self.npts.shape[1] == len(self.cameras)
        assert self.world.shape == (self.num_obs, 3) and self.image.shape == (self.num_obs, 2)
        assert self.poses.shape == (self.num_frames, 4, 4)

    @property
    def num_frames(self) -> int:
        return int(self.npts.shape[0])

    @property
    def num_cameras(self) -> int:
        return len(self.cameras)

    @property
    def num_obs(self) -> int:
        return int(self.npts.sum())

    @property
    def num_cam_params(self) -> int:
        return sum(4 + len(c.distortion_coeffs) for c in self.cameras) + 6 * (self.num_cameras - 1)

    def frame_offsets(self) -> np.ndarray:
        """observation offset of each frame, length F+1"""
        per_frame = self.npts.sum(axis=1, dtype=np.int64)
        return np.concatenate([[0], np.cumsum(per_frame)])

    def slice_frames(self, lo: int, hi: int) -> "Rig":
        """contiguous frame range [lo, hi) as its own rig (how frames are sharded over ranks)"""
        off = self.frame_offsets()
        return Rig([c.copy() for c in self.cameras], self.npts[lo:hi], self.world[off[lo]:off[hi]],
                   self.image[off[lo]:off[hi]], self.poses[lo:hi], self.fix_all_intrinsics, None)


def shard_bounds(npts: np.ndarray, world_size: int) -> np.ndarray:
    """Contiguous frame ranges balanced by observation count.

    Returns ``bounds`` of length world_size+1: rank r owns frames [bounds[r], bounds[r+1]).
    Frames are the independent unit of the reference's problem (one residual block per frame,
    /root/reference/src/calibrator.cpp:466-506), so the split needs no data-path exchange.
    """
    per_frame = np.asarray(npts).sum(axis=1, dtype=np.int64)
    cum = np.concatenate([[0], np.cumsum(per_frame)])
    total = cum[-1]
    F = len(per_frame)
    bounds = [0]
    for r in range(1, world_size):
        target = total * r / world_size
        b = int(np.searchsorted(cum, target, side="left"))
        b = min(max(b, bounds[-1]), F)
        bounds.append(b)
    bounds.append(F)
    return np.asarray(bounds, dtype=np.int64)
