"""Post-BA track filter.  TEST INFRASTRUCTURE ONLY (imported by tests/, never by mpsfm_b200/).

The reference filters after every bundle through pycolmap (mpsfm/sfm/mapper/base.py:686-711 ``filter_all`` /
``filter_bundle``, :772-797 ``filter_local_points3D``):

    rec.obs.filter_points3D(max_reproj_error, min_tri_angle, point3D_ids)

``obs`` is COLMAP's ``ObservationManager``; its C++ is a third-party dependency that is NOT under /root/reference
(pycolmap, the fork the reference's requirements pin; no wheel here) -- PARITY UNPINNED.  What follows restates the
algorithm COLMAP publishes (``ObservationManager::FilterPoints3D`` = ``FilterPoints3DWithLargeReprojectionError`` then
``FilterPoints3DWithSmallTriangulationAngle``, ``CalculateSquaredReprojectionError``, ``CalculateTriangulationAngle``,
``Image::ProjectionCenter``) and is anchored on the reference's call sites above:

* phase 1, per selected point: a track shorter than 2 deletes the point (``num_filtered += len``).  Otherwise every
  track element with squared reprojection error ``> max_reproj_error**2`` is marked (a point with camera-frame
  ``z < DBL_EPSILON`` has error ``DBL_MAX``), the others add ``sqrt(err2)`` to a sum.  If ``marked >= len - 1`` the
  point is deleted (``num_filtered += len``), else the marked observations are deleted (``num_filtered += marked``)
  and ``point.error = sum / remaining``.
* phase 2, per selected point that still exists: with projection centres ``c = -R^T t`` of its (remaining) track
  images, the point is kept as soon as one pair has ``CalculateTriangulationAngle(c1, c2, X) >= min_tri_angle``;
  otherwise it is deleted (``num_filtered += remaining``).  The angle is ``acos((r1 + r2 - b) / (2 sqrt(r1 r2)))`` on
  squared lengths, folded with ``min(angle, pi - angle)``; a zero denominator gives 0.

Plain per-track Python loops: meant for scenes of a few thousand observations.
"""

from __future__ import annotations

import numpy as np

from . import camera_models, geometry

DBL_MAX = np.finfo(np.float64).max
DBL_EPS = np.finfo(np.float64).eps


def squared_reprojection_errors(problem, q, t, X):
    """[N] squared reprojection error of every observation (natural order), DBL_MAX behind the camera."""
    oi, op = problem.obs_img, problem.obs_pt
    pc = geometry.quat_rotate(q[oi], X[op]) + t[oi]
    cam = problem.img_camera[oi]
    front = pc[:, 2] >= DBL_EPS
    safe = pc.copy()
    safe[~front] = (0.0, 0.0, 1.0)
    xy = camera_models.img_from_cam_values(problem.cam_model[cam], problem.cam_params[cam], safe)
    e2 = np.sum((xy - problem.obs_xy) ** 2, axis=1)
    e2[~front] = DBL_MAX
    return e2


def triangulation_angle(c1, c2, X):
    b = float(np.sum((c1 - c2) ** 2))
    r1 = float(np.sum((X - c1) ** 2))
    r2 = float(np.sum((X - c2) ** 2))
    den = 2.0 * np.sqrt(r1 * r2)
    if den == 0.0:
        return 0.0
    with np.errstate(invalid="ignore"):
        ang = abs(np.arccos((r1 + r2 - b) / den))
    oth = np.pi - ang
    return oth if oth < ang else ang  # std::min(angle, pi - angle): nan stays nan


def filter_points3D(problem, q, t, X, max_reproj_error, min_tri_angle_deg, select=None):
    """Returns (obs_keep [N] bool natural order, pt_keep [P] bool, pt_error [P], num_filtered).

    ``select`` [P] bool: the ``point3D_ids`` argument (None = every point); unselected points are kept,
    their error is reported as -1 (COLMAP leaves it untouched)."""
    q = q / np.linalg.norm(q, axis=1, keepdims=True)
    N, P = problem.num_obs, problem.num_points
    e2 = squared_reprojection_errors(problem, q, t, X)
    centers = -np.einsum("nji,nj->ni", geometry.rotmat(q), t)
    order = np.argsort(problem.obs_pt, kind="stable")
    counts = np.bincount(problem.obs_pt, minlength=P)
    ptr = np.concatenate([[0], np.cumsum(counts)])
    obs_keep = np.ones(N, dtype=bool)
    pt_keep = np.ones(P, dtype=bool)
    pt_error = np.full(P, -1.0)
    max2 = max_reproj_error * max_reproj_error
    min_angle = np.deg2rad(min_tri_angle_deg)
    num_filtered = 0
    for p in range(P):
        if select is not None and not select[p]:
            continue
        track = order[ptr[p]:ptr[p + 1]]
        n = len(track)
        # phase 1
        marked = [k for k in track if e2[k] > max2]
        if n < 2 or len(marked) >= n - 1:
            num_filtered += n
            obs_keep[track] = False
            pt_keep[p] = False
            continue
        num_filtered += len(marked)
        obs_keep[marked] = False
        rest = [k for k in track if not e2[k] > max2]
        pt_error[p] = float(np.sum(np.sqrt(e2[rest]))) / len(rest)
        # phase 2
        keep = False
        for i1 in range(len(rest)):
            for i2 in range(i1):
                a = triangulation_angle(centers[problem.obs_img[rest[i1]]], centers[problem.obs_img[rest[i2]]], X[p])
                if a >= min_angle:
                    keep = True
                    break
            if keep:
                break
        if not keep:
            num_filtered += len(rest)
            obs_keep[rest] = False
            pt_keep[p] = False
            pt_error[p] = -1.0
    return obs_keep, pt_keep, pt_error, num_filtered
