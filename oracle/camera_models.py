"""COLMAP camera models ``ImgFromCam`` (upstream ``src/colmap/sensor/models.h``; not in
/root/reference -- pycolmap fork, docker/install_colmap.sh:28).  TEST INFRASTRUCTURE ONLY.

Each function maps normalised camera coordinates ``(u, v) = (X/Z, Y/Z)`` to pixels and works on
plain arrays or on :class:`oracle.jet.Jet`, so the Jacobians come from forward-mode duals exactly
as Ceres autodiff produces them for the reference (bundle_adjustment.py:104).  ``params`` is [n,12].
Model ids follow COLMAP; the intrinsics are constant in every reference call site
(bundle_adjustment.py:92-94), so no derivative w.r.t. ``params`` is needed.
"""

from __future__ import annotations

import numpy as np

from . import jet as J

EPS = np.finfo(np.float64).eps

SIMPLE_PINHOLE, PINHOLE, SIMPLE_RADIAL, RADIAL, OPENCV = 0, 1, 2, 3, 4
OPENCV_FISHEYE, FULL_OPENCV, FOV, SIMPLE_RADIAL_FISHEYE, RADIAL_FISHEYE, THIN_PRISM_FISHEYE = 5, 6, 7, 8, 9, 10


def _simple_pinhole(p, u, v):
    return p[:, 0] * u + p[:, 1], p[:, 0] * v + p[:, 2]


def _pinhole(p, u, v):
    return p[:, 0] * u + p[:, 2], p[:, 1] * v + p[:, 3]


def _simple_radial(p, u, v):
    r2 = u * u + v * v
    radial = p[:, 3] * r2
    return p[:, 0] * (u + u * radial) + p[:, 1], p[:, 0] * (v + v * radial) + p[:, 2]


def _radial(p, u, v):
    r2 = u * u + v * v
    radial = p[:, 3] * r2 + p[:, 4] * r2 * r2
    return p[:, 0] * (u + u * radial) + p[:, 1], p[:, 0] * (v + v * radial) + p[:, 2]


def _opencv(p, u, v):
    k1, k2, p1, p2 = p[:, 4], p[:, 5], p[:, 6], p[:, 7]
    u2, v2, uv = u * u, v * v, u * v
    r2 = u2 + v2
    radial = k1 * r2 + k2 * r2 * r2
    du = u * radial + 2.0 * p1 * uv + p2 * (r2 + 2.0 * u2)
    dv = v * radial + 2.0 * p2 * uv + p1 * (r2 + 2.0 * v2)
    return p[:, 0] * (u + du) + p[:, 2], p[:, 1] * (v + dv) + p[:, 3]


def _fisheye_thetad_over_r(u, v, poly):
    """theta_d / r with theta = atan(r); equals 1 where r <= eps (COLMAP leaves du = dv = 0)."""
    r2v = J.value(u * u + v * v)
    small = np.sqrt(r2v) <= EPS
    # keep the jet arithmetic finite on the masked-out lanes
    safe_u = J.where(small, u * 0.0 + 1.0, u)
    r = J.sqrt(safe_u * safe_u + v * v)
    theta = J.atan(r)
    thetad = theta * poly(theta * theta)
    return J.where(small, u * 0.0 + 1.0, thetad / r)


def _opencv_fisheye(p, u, v):
    k1, k2, k3, k4 = p[:, 4], p[:, 5], p[:, 6], p[:, 7]
    s = _fisheye_thetad_over_r(u, v, lambda t2: 1.0 + t2 * (k1 + t2 * (k2 + t2 * (k3 + t2 * k4))))
    return p[:, 0] * (u * s) + p[:, 2], p[:, 1] * (v * s) + p[:, 3]


def _full_opencv(p, u, v):
    k1, k2, p1, p2, k3, k4, k5, k6 = (p[:, i] for i in range(4, 12))
    u2, v2, uv = u * u, v * v, u * v
    r2 = u2 + v2
    r4 = r2 * r2
    r6 = r4 * r2
    radial = (1.0 + k1 * r2 + k2 * r4 + k3 * r6) / (1.0 + k4 * r2 + k5 * r4 + k6 * r6)
    du = u * radial + 2.0 * p1 * uv + p2 * (r2 + 2.0 * u2) - u
    dv = v * radial + 2.0 * p2 * uv + p1 * (r2 + 2.0 * v2) - v
    return p[:, 0] * (u + du) + p[:, 2], p[:, 1] * (v + dv) + p[:, 3]


def _fov(p, u, v):
    omega = p[:, 4]
    omega2 = omega * omega
    r2 = u * u + v * v
    r2v = J.value(r2)
    kEps = 1e-4
    small_w = omega2 < kEps
    small_r = (~small_w) & (r2v < kEps)
    tan_half = np.tan(np.where(small_w, 1.0, omega) / 2.0)
    f_taylor_w = (omega2 * r2) / 3.0 - omega2 / 12.0 + 1.0
    f_taylor_r = (-2.0 * tan_half * (4.0 * r2 * tan_half * tan_half - 3.0)) / (3.0 * np.where(small_w, 1.0, omega))
    safe_r2 = J.where(r2v < kEps, r2 * 0.0 + 1.0, r2)
    radius = J.sqrt(safe_r2)
    f_full = J.atan(radius * 2.0 * tan_half) / (radius * np.where(small_w, 1.0, omega))
    factor = J.where(small_w, f_taylor_w, J.where(small_r, f_taylor_r, f_full))
    return p[:, 0] * (u * factor) + p[:, 2], p[:, 1] * (v * factor) + p[:, 3]


def _simple_radial_fisheye(p, u, v):
    k = p[:, 3]
    s = _fisheye_thetad_over_r(u, v, lambda t2: 1.0 + k * t2)
    return p[:, 0] * (u * s) + p[:, 1], p[:, 0] * (v * s) + p[:, 2]


def _radial_fisheye(p, u, v):
    k1, k2 = p[:, 3], p[:, 4]
    s = _fisheye_thetad_over_r(u, v, lambda t2: 1.0 + t2 * (k1 + t2 * k2))
    return p[:, 0] * (u * s) + p[:, 1], p[:, 0] * (v * s) + p[:, 2]


def _thin_prism_fisheye(p, u, v):
    k1, k2, p1, p2, k3, k4, sx1, sy1 = (p[:, i] for i in range(4, 12))
    s = _fisheye_thetad_over_r(u, v, lambda t2: 1.0 + 0.0 * t2)  # theta / r
    uu, vv = u * s, v * s
    u2, v2, uv = uu * uu, vv * vv, uu * vv
    t2 = u2 + v2
    t4 = t2 * t2
    radial = k1 * t2 + k2 * t4 + k3 * t4 * t2 + k4 * t4 * t4
    du = uu * radial + 2.0 * p1 * uv + p2 * (t2 + 2.0 * u2) + sx1 * t2
    dv = vv * radial + 2.0 * p2 * uv + p1 * (t2 + 2.0 * v2) + sy1 * t2
    return p[:, 0] * (uu + du) + p[:, 2], p[:, 1] * (vv + dv) + p[:, 3]


_MODELS = {
    SIMPLE_PINHOLE: _simple_pinhole,
    PINHOLE: _pinhole,
    SIMPLE_RADIAL: _simple_radial,
    RADIAL: _radial,
    OPENCV: _opencv,
    OPENCV_FISHEYE: _opencv_fisheye,
    FULL_OPENCV: _full_opencv,
    FOV: _fov,
    SIMPLE_RADIAL_FISHEYE: _simple_radial_fisheye,
    RADIAL_FISHEYE: _radial_fisheye,
    THIN_PRISM_FISHEYE: _thin_prism_fisheye,
}


def img_from_cam(model_ids, params, pc):
    """Project camera-frame points.

    model_ids [n] int, params [n,12], pc [n,3].  Returns (xy [n,2], dxy_dpc [n,2,3]) with the
    derivative taken through ``u = X/Z, v = Y/Z`` and the model, as Ceres' jets would."""
    model_ids = np.asarray(model_ids)
    n = pc.shape[0]
    xy = np.zeros((n, 2))
    jac = np.zeros((n, 2, 3))
    for m in np.unique(model_ids):
        sel = np.flatnonzero(model_ids == m)
        X, Y, Z = J.Jet.variables(pc[sel])
        u, v = X / Z, Y / Z
        x, y = _MODELS[int(m)](params[sel], u, v)
        x, y = X._lift(x), X._lift(y)
        xy[sel, 0], xy[sel, 1] = x.a, y.a
        jac[sel, 0, :], jac[sel, 1, :] = x.v, y.v
    return xy, jac


def img_from_cam_values(model_ids, params, pc):
    """Value-only projection (plain arrays)."""
    model_ids = np.asarray(model_ids)
    xy = np.zeros((pc.shape[0], 2))
    for m in np.unique(model_ids):
        sel = np.flatnonzero(model_ids == m)
        u, v = pc[sel, 0] / pc[sel, 2], pc[sel, 1] / pc[sel, 2]
        x, y = _MODELS[int(m)](params[sel], u, v)
        xy[sel, 0], xy[sel, 1] = x, y
    return xy
