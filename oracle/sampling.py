"""Depth-map sampling at keypoints and the z-depth projection helper.  TEST INFRASTRUCTURE ONLY.

Restates, in numpy:

* ``PriorUtils._data_at_kps`` (/root/reference/mpsfm/sfm/scene/image/mixins/priorutils.py:49-62): fp64
  bilinear ``torch.nn.functional.grid_sample(..., mode="bilinear", padding_mode="zeros",
  align_corners=True)`` of an H x W map at the keypoints scaled by ``camera.sx, camera.sy``.  The
  normalise / un-normalise round trip of the reference is kept operation for operation so that the sample
  position carries the same rounding;
* ``project3D_colmap`` / ``project3D`` (/root/reference/mpsfm/utils/geometry.py:6-19): pinhole pixel
  coordinates through ``calibration_matrix()`` (distortion ignored) and the z-depth of the points.

PINNED against the reference itself: tests/golden/make_golden.py imports those two reference files
(they only need numpy and torch) and stores their outputs in tests/golden/sampling.npz; tests/test_golden.py
compares this restatement with them.
"""

from __future__ import annotations

import numpy as np


def data_at_kps(kps: np.ndarray, data: np.ndarray, sx: float, sy: float) -> np.ndarray:
    kps = np.asarray(kps, dtype=np.float64)
    if kps.ndim == 1:
        kps = kps[None]
    data = np.asarray(data, dtype=np.float64)
    H, W = data.shape[:2]
    kp = kps * np.array([sx, sy])
    # priorutils.py:55-56 (normalise to [-1, 1]) followed by grid_sample's align_corners=True un-normalise
    xn = (kp[:, 0] / (W - 1)) * 2 - 1
    yn = (kp[:, 1] / (H - 1)) * 2 - 1
    ix = ((xn + 1) / 2) * (W - 1)
    iy = ((yn + 1) / 2) * (H - 1)
    x0 = np.floor(ix)
    y0 = np.floor(iy)
    x1, y1 = x0 + 1, y0 + 1
    w_nw = (x1 - ix) * (y1 - iy)
    w_ne = (ix - x0) * (y1 - iy)
    w_sw = (x1 - ix) * (iy - y0)
    w_se = (ix - x0) * (iy - y0)

    def tap(xx, yy):
        inside = (xx >= 0) & (xx <= W - 1) & (yy >= 0) & (yy <= H - 1)
        xi = np.clip(xx, 0, W - 1).astype(np.int64)
        yi = np.clip(yy, 0, H - 1).astype(np.int64)
        return np.where(inside, data[yi, xi], 0.0)

    return tap(x0, y0) * w_nw + tap(x1, y0) * w_ne + tap(x0, y1) * w_sw + tap(x1, y1) * w_se


def valid_at_kps(kps, valid, sx, sy):
    """priorutils.py:42-47: all four neighbours valid <=> the interpolated mask equals 1."""
    return data_at_kps(kps, np.asarray(valid, dtype=np.float64), sx, sy) == 1


def project3D(points3D: np.ndarray, H: np.ndarray, K: np.ndarray):
    """geometry.py:13-19: H is the 4x4 cam_from_world, K the calibration matrix."""
    points3D_h = np.hstack([points3D, np.ones((points3D.shape[0], 1))])
    points_cam = (H @ points3D_h.T)[:3, :].T
    depth = points_cam[:, 2].copy()
    pts = ((K @ (points_cam / depth[:, None]).T).T)[:, :2]
    return pts, depth


def project3D_colmap(cam_from_world_3x4: np.ndarray, K: np.ndarray, points3D: np.ndarray):
    """geometry.py:6-10"""
    H = np.concatenate([cam_from_world_3x4, np.array([[0, 0, 0, 1]])], axis=0)
    return project3D(points3D, H, K)
