"""Quaternion / manifold arithmetic of the reference's Ceres problem.  TEST INFRASTRUCTURE ONLY.

* quaternion storage x,y,z,w (Eigen / ``pycolmap.Rotation3d.quat``,
  /root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:115,122),
* ``q * v`` as Eigen evaluates it inside COLMAP's ``ReprojErrorCostFunctor``
  (``EigenQuaternionMap<T>(q) * X + t``; upstream cost_functions.h),
* ``ceres::EigenQuaternionManifold::Plus`` (upstream manifold.cc): ``q (+) d = dq(d) (x) q`` with
  ``dq = [sin|d| d/|d|, cos|d|]`` -- left multiplication, ``d`` a half-angle vector.
"""

from __future__ import annotations

import numpy as np


def quat_rotate(q, v):
    """Eigen ``Quaternion::_transformVector``: v + 2w (u x v) + 2 u x (u x v), batched."""
    u, w = q[..., :3], q[..., 3:4]
    uv = np.cross(u, v)
    uv = uv + uv
    return v + w * uv + np.cross(u, uv)


def quat_mul(a, b):
    """Hamilton product a (x) b, storage x,y,z,w."""
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack(
        [
            aw * bx + ax * bw + ay * bz - az * by,
            aw * by - ax * bz + ay * bw + az * bx,
            aw * bz + ax * by - ay * bx + az * bw,
            aw * bw - ax * bx - ay * by - az * bz,
        ],
        axis=-1,
    )


def quat_plus(q, delta):
    """EigenQuaternionManifold::Plus, batched; delta [n,3]."""
    nrm = np.linalg.norm(delta, axis=-1)
    safe = np.where(nrm == 0.0, 1.0, nrm)
    sbd = np.sin(nrm) / safe
    dq = np.concatenate([delta * sbd[..., None], np.cos(nrm)[..., None]], axis=-1)
    out = quat_mul(dq, q)
    return np.where((nrm == 0.0)[..., None], q, out)


def quat_plus_jacobian(q):
    """EigenQuaternionManifold::PlusJacobian, [n,4,3] (rows x,y,z,w)."""
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    J = np.empty(q.shape[:-1] + (4, 3))
    J[..., 0, 0], J[..., 0, 1], J[..., 0, 2] = w, z, -y
    J[..., 1, 0], J[..., 1, 1], J[..., 1, 2] = -z, w, x
    J[..., 2, 0], J[..., 2, 1], J[..., 2, 2] = y, -x, w
    J[..., 3, 0], J[..., 3, 1], J[..., 3, 2] = -x, -y, -z
    return J


def rotmat(q):
    """Rotation matrix of a unit quaternion (x,y,z,w), [n,3,3]."""
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    R = np.empty(q.shape[:-1] + (3, 3))
    R[..., 0, 0] = 1 - 2 * (y * y + z * z)
    R[..., 0, 1] = 2 * (x * y - z * w)
    R[..., 0, 2] = 2 * (x * z + y * w)
    R[..., 1, 0] = 2 * (x * y + z * w)
    R[..., 1, 1] = 1 - 2 * (x * x + z * z)
    R[..., 1, 2] = 2 * (y * z - x * w)
    R[..., 2, 0] = 2 * (x * z - y * w)
    R[..., 2, 1] = 2 * (y * z + x * w)
    R[..., 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def skew(v):
    S = np.zeros(v.shape[:-1] + (3, 3))
    S[..., 0, 1], S[..., 0, 2] = -v[..., 2], v[..., 1]
    S[..., 1, 0], S[..., 1, 2] = v[..., 2], -v[..., 0]
    S[..., 2, 0], S[..., 2, 1] = -v[..., 1], v[..., 0]
    return S
