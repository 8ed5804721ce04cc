"""Point-to-plane rows from monocular normals -- EXTENSION, no counterpart in the reference.

The reference uses monocular surface normals only inside its depth-integration solver
(/root/reference/mpsfm/extraction/imagewise/geometry/models/utils/integrator.py, the ``mpba_integrate`` twin of this
package); its bundle adjustment (mpsfm/sfm/mapper/bundle_adjustment.py:60-100) has reprojection and log-depth rows
and nothing for normals.  SURVEY.md section 8 row a-9 lists a normal row as a possible extension; this file holds its host
side.  One row per observation that has a depth prior AND a normal:

    r = m . p_cam - c,     m = sqrt(weight) n,     c = m . q,     q = depth * (u, v, 1)

with (u, v) the keypoint in normalised camera coordinates (undistorted here, once, on the host), ``depth`` the prior
z-depth after the image's constant shift / scale (exp(s) d + b, the same alignment the log-depth row uses) and ``n``
the unit normal in the camera frame.  The row says: the triangulated point lies on the tangent plane of the monocular
surface at the keypoint.  It constrains the point along the normal only, so it complements the log-depth row (which
constrains it along the optical axis) where the surface is slanted.

Everything downstream (ba_math.cuh: linearize_obs, oracle/residuals.py: evaluate) only sees ``{m, c}`` per observation.
"""

from __future__ import annotations

import dataclasses

import numpy as np

from .problem import BundleProblem, CameraModel
from .synthetic import _distort


def cam_from_img(model: CameraModel, params: np.ndarray, xy: np.ndarray, iterations: int = 12) -> np.ndarray:
    """Normalised camera coordinates [n,2] of pixels ``xy`` [n,2]: Newton on the forward model with a central-difference
    Jacobian (COLMAP's ``IterativeUndistortion`` does the same with 100 iterations and a 1e-10 step; the forward models
    are smooth and the start is the pinhole answer, a dozen steps reach 1e-13)."""
    model = CameraModel(model)
    p = np.asarray(params, dtype=np.float64)
    xy = np.asarray(xy, dtype=np.float64).reshape(-1, 2)
    if model in (CameraModel.SIMPLE_PINHOLE, CameraModel.SIMPLE_RADIAL, CameraModel.RADIAL,
                 CameraModel.SIMPLE_RADIAL_FISHEYE, CameraModel.RADIAL_FISHEYE):
        fx, fy, cx, cy = p[0], p[0], p[1], p[2]
    else:
        fx, fy, cx, cy = p[0], p[1], p[2], p[3]
    uv = np.stack([(xy[:, 0] - cx) / fx, (xy[:, 1] - cy) / fy], axis=1)
    if model in (CameraModel.SIMPLE_PINHOLE, CameraModel.PINHOLE) or uv.shape[0] == 0:
        return uv
    eps = 1e-6
    for _ in range(iterations):
        u, v = uv[:, 0], uv[:, 1]
        # the fisheye family divides by r: keep the exact centre off zero
        tiny = (u * u + v * v) < 1e-24
        u = np.where(tiny, 1e-12, u)
        x0, y0 = _distort(model, p, u, v)
        xu1, yu1 = _distort(model, p, u + eps, v)
        xu0, yu0 = _distort(model, p, u - eps, v)
        xv1, yv1 = _distort(model, p, u, v + eps)
        xv0, yv0 = _distort(model, p, u, v - eps)
        a, b = (xu1 - xu0) / (2 * eps), (xv1 - xv0) / (2 * eps)
        c, d = (yu1 - yu0) / (2 * eps), (yv1 - yv0) / (2 * eps)
        ex, ey = x0 - xy[:, 0], y0 - xy[:, 1]
        det = a * d - b * c
        du, dv = (d * ex - b * ey) / det, (a * ey - c * ex) / det
        uv = np.stack([u - du, v - dv], axis=1)
        if max(float(np.max(np.abs(du))), float(np.max(np.abs(dv)))) < 1e-14:
            break
    return uv


def plane_rows(problem: BundleProblem, obs: np.ndarray, normals_cam: np.ndarray, weight) -> tuple:
    """``(plane_obs, plane_m, plane_c)`` for observations ``obs`` (a subset of ``problem.depth_obs``) with unit normals
    ``normals_cam`` [n,3] in the camera frame and row weights ``weight`` (scalar or [n], 1 / sigma^2 in scene units)."""
    obs = np.asarray(obs, dtype=np.int32).reshape(-1)
    n = np.asarray(normals_cam, dtype=np.float64).reshape(-1, 3)
    if n.shape[0] != obs.shape[0]:
        raise ValueError("one normal per listed observation")
    w = np.broadcast_to(np.asarray(weight, dtype=np.float64), obs.shape)
    if np.any(w < 0):
        raise ValueError("weights must be >= 0")
    # prior depth of each listed observation
    slot = np.full(problem.num_obs, -1, dtype=np.int64)
    slot[problem.depth_obs] = np.arange(problem.num_depth)
    k = slot[obs]
    if np.any(k < 0):
        raise ValueError("a normal prior needs a depth prior on the same observation")
    img = problem.obs_img[obs]
    b, s = problem.shift_scale[img, 0], problem.shift_scale[img, 1]
    depth = np.exp(s) * problem.depth_d[k] + b
    uv = np.zeros((obs.shape[0], 2))
    cam = problem.img_camera[img]
    for c in np.unique(cam):
        sel = np.flatnonzero(cam == c)
        uv[sel] = cam_from_img(CameraModel(int(problem.cam_model[c])), problem.cam_params[c], problem.obs_xy[obs[sel]])
    q = depth[:, None] * np.concatenate([uv, np.ones((obs.shape[0], 1))], axis=1)
    norm = np.linalg.norm(n, axis=1, keepdims=True)
    m = np.sqrt(w)[:, None] * n / np.where(norm > 0, norm, 1.0)
    return obs, m, np.einsum("ni,ni->n", m, q)


def with_normal_priors(problem: BundleProblem, obs, normals_cam, weight, loss_scale: float = 1.0) -> BundleProblem:
    """A copy of ``problem`` carrying the rows of :func:`plane_rows`."""
    o, m, c = plane_rows(problem, obs, normals_cam, weight)
    return dataclasses.replace(problem, plane_obs=o, plane_m=m, plane_c=c,
                               plane_loss_scale=float(loss_scale)).normalized()


def synthetic_normal_priors(problem: BundleProblem, frac: float = 0.5, seed: int = 0, rel_sigma: float = 0.05,
                            loss_scale: float = 2.0, tilt: float = 0.6) -> BundleProblem:
    """Seeded normal priors on a fraction of the depth-prior observations of a synthetic scene: normals facing the
    camera, tilted by up to ``tilt`` rad off the optical axis; plane-distance std ``rel_sigma`` x the prior depth (the
    order of the scenes' depth noise, so the rows sit at ~1 sigma at the optimum); ``loss_scale`` in units of sigma."""
    rng = np.random.Generator(np.random.PCG64(seed + 9176))
    keep = rng.random(problem.num_depth) < frac
    pick = problem.depth_obs[keep]
    n = pick.shape[0]
    axis = rng.normal(size=(n, 3))
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    normals = np.array([0.0, 0.0, -1.0]) + np.tan(rng.uniform(0, tilt, size=(n, 1))) * axis * np.array([1.0, 1.0, 0.0])
    sigma = rel_sigma * problem.depth_d[keep]
    return with_normal_priors(problem, pick, normals, 1.0 / sigma**2, loss_scale)
