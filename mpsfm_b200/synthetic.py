"""Seeded synthetic scenes of the shapes BASELINE.json names (SURVEY.md section 8 row d).

Images sit on a perturbed ring looking at a box of points; every point is observed by a window
of ring-neighbouring images, which gives the banded co-visibility of a real reconstruction (the
reduced camera system is block-sparse).  Observations carry pixel noise and outliers, depth priors
carry log-normal noise and gross outliers, and the start point is the truth perturbed in poses and
points.  The gauge is the reference's: first image constant, second image translation-x constant
(/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:111-122).
"""

from __future__ import annotations

import numpy as np

from .problem import FREE_ALL, FREE_FIX_TX, FREE_NONE, BundleProblem, CameraModel, LossType

DEFAULT_PARAMS = {
    CameraModel.SIMPLE_PINHOLE: [1200.0, 800.0, 600.0],
    CameraModel.PINHOLE: [1200.0, 1200.0, 800.0, 600.0],
    CameraModel.SIMPLE_RADIAL: [1200.0, 800.0, 600.0, 0.02],
    CameraModel.RADIAL: [1200.0, 800.0, 600.0, 0.02, -0.005],
    CameraModel.OPENCV: [1200.0, 1190.0, 800.0, 600.0, 0.02, -0.005, 1e-3, -5e-4],
    CameraModel.OPENCV_FISHEYE: [1200.0, 1190.0, 800.0, 600.0, 0.02, -0.01, 3e-3, -1e-3],
    CameraModel.FULL_OPENCV: [1200.0, 1190.0, 800.0, 600.0, 0.05, -0.01, 1e-3, -2e-3, 3e-3, 0.02, -4e-3, 1e-3],
    CameraModel.FOV: [1200.0, 1190.0, 800.0, 600.0, 0.9],
    CameraModel.SIMPLE_RADIAL_FISHEYE: [1200.0, 800.0, 600.0, 0.03],
    CameraModel.RADIAL_FISHEYE: [1200.0, 800.0, 600.0, 0.03, -0.004],
    CameraModel.THIN_PRISM_FISHEYE: [1200.0, 1190.0, 800.0, 600.0, 0.02, -0.01, 1e-3, -2e-3, 3e-3, -1e-3, 2e-3, -3e-3],
}

# named shapes: (images, points, observations per point)
CONFIGS = {
    "tiny": (6, 80, 4),
    "small": (20, 2000, 5),
    "cfg2": (100, 50_000, 6),  # 300k obs, PINHOLE
    "cfg3": (1000, 1_000_000, 5),  # 5M obs, SIMPLE_RADIAL
    "target10m": (2000, 2_000_000, 5),  # 10M obs
    "cfg4": (5000, 5_000_000, 5),  # 25M obs
    "cfg5": (20000, 20_000_000, 5),  # 100M obs
}


def quat_from_rotmat(R: np.ndarray) -> np.ndarray:
    """Batch rotation matrices [n,3,3] -> unit quaternions [n,4] in x,y,z,w order."""
    R = np.asarray(R, dtype=np.float64)
    n = R.shape[0]
    q = np.empty((n, 4))
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    # robust branchless-ish conversion: pick the largest of (w,x,y,z) per row
    cand = np.stack([tr, R[:, 0, 0], R[:, 1, 1], R[:, 2, 2]], axis=1)
    k = np.argmax(cand, axis=1)
    for case in range(4):
        m = k == case
        if not m.any():
            continue
        Rm = R[m]
        if case == 0:
            s = np.sqrt(Rm[:, 0, 0] + Rm[:, 1, 1] + Rm[:, 2, 2] + 1.0) * 2
            q[m, 3] = 0.25 * s
            q[m, 0] = (Rm[:, 2, 1] - Rm[:, 1, 2]) / s
            q[m, 1] = (Rm[:, 0, 2] - Rm[:, 2, 0]) / s
            q[m, 2] = (Rm[:, 1, 0] - Rm[:, 0, 1]) / s
        else:
            i = case - 1
            j, l = (i + 1) % 3, (i + 2) % 3
            s = np.sqrt(1.0 + Rm[:, i, i] - Rm[:, j, j] - Rm[:, l, l]) * 2
            q[m, i] = 0.25 * s
            q[m, 3] = (Rm[:, l, j] - Rm[:, j, l]) / s
            q[m, j] = (Rm[:, j, i] + Rm[:, i, j]) / s
            q[m, l] = (Rm[:, l, i] + Rm[:, i, l]) / s
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q


def rotmat_from_quat(q: np.ndarray) -> np.ndarray:
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.empty((q.shape[0], 3, 3))
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - z * w)
    R[:, 0, 2] = 2 * (x * z + y * w)
    R[:, 1, 0] = 2 * (x * y + z * w)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - x * w)
    R[:, 2, 0] = 2 * (x * z - y * w)
    R[:, 2, 1] = 2 * (y * z + x * w)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def _rodrigues(rv: np.ndarray) -> np.ndarray:
    th = np.linalg.norm(rv, axis=1)
    k = rv / np.maximum(th, 1e-300)[:, None]
    K = np.zeros((rv.shape[0], 3, 3))
    K[:, 0, 1], K[:, 0, 2] = -k[:, 2], k[:, 1]
    K[:, 1, 0], K[:, 1, 2] = k[:, 2], -k[:, 0]
    K[:, 2, 0], K[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = np.sin(th)[:, None, None], np.cos(th)[:, None, None]
    return np.eye(3)[None] + s * K + (1 - c) * (K @ K)


def _distort(model: CameraModel, params, u, v):
    """Pixel coordinates for the models the generator supports (COLMAP ``ImgFromCam``)."""
    p = params
    r2 = u * u + v * v
    if model == CameraModel.SIMPLE_PINHOLE:
        return p[0] * u + p[1], p[0] * v + p[2]
    if model == CameraModel.PINHOLE:
        return p[0] * u + p[2], p[1] * v + p[3]
    if model == CameraModel.SIMPLE_RADIAL:
        g = 1 + p[3] * r2
        return p[0] * u * g + p[1], p[0] * v * g + p[2]
    if model == CameraModel.RADIAL:
        g = 1 + p[3] * r2 + p[4] * r2 * r2
        return p[0] * u * g + p[1], p[0] * v * g + p[2]
    if model == CameraModel.OPENCV:
        rad = p[4] * r2 + p[5] * r2 * r2
        du = u * rad + 2 * p[6] * u * v + p[7] * (r2 + 2 * u * u)
        dv = v * rad + 2 * p[7] * u * v + p[6] * (r2 + 2 * v * v)
        return p[0] * (u + du) + p[2], p[1] * (v + dv) + p[3]
    if model == CameraModel.FULL_OPENCV:
        r4 = r2 * r2
        rad = (1 + p[4] * r2 + p[5] * r4 + p[8] * r4 * r2) / (1 + p[9] * r2 + p[10] * r4 + p[11] * r4 * r2)
        x = u * rad + 2 * p[6] * u * v + p[7] * (r2 + 2 * u * u)
        y = v * rad + 2 * p[7] * u * v + p[6] * (r2 + 2 * v * v)
        return p[0] * x + p[2], p[1] * y + p[3]
    if model == CameraModel.FOV:
        # the generator only draws omega, r well inside the closed-form branch (COLMAP switches to series below 1e-4)
        r = np.sqrt(r2)
        f = np.arctan(2 * r * np.tan(p[4] / 2)) / (r * p[4])
        return p[0] * u * f + p[2], p[1] * v * f + p[3]
    # fisheye family: (u, v) scaled by theta_d / r, theta = atan(r)
    r = np.sqrt(r2)
    th = np.arctan(r)
    t2 = th * th
    if model == CameraModel.OPENCV_FISHEYE:
        s = th * (1 + t2 * (p[4] + t2 * (p[5] + t2 * (p[6] + t2 * p[7])))) / r
        return p[0] * u * s + p[2], p[1] * v * s + p[3]
    if model == CameraModel.SIMPLE_RADIAL_FISHEYE:
        s = th * (1 + p[3] * t2) / r
        return p[0] * u * s + p[1], p[0] * v * s + p[2]
    if model == CameraModel.RADIAL_FISHEYE:
        s = th * (1 + t2 * (p[3] + t2 * p[4])) / r
        return p[0] * u * s + p[1], p[0] * v * s + p[2]
    if model == CameraModel.THIN_PRISM_FISHEYE:
        uu, vv = u * th / r, v * th / r
        q2 = uu * uu + vv * vv
        q4 = q2 * q2
        rad = p[4] * q2 + p[5] * q4 + p[8] * q4 * q2 + p[9] * q4 * q4
        du = uu * rad + 2 * p[6] * uu * vv + p[7] * (q2 + 2 * uu * uu) + p[10] * q2
        dv = vv * rad + 2 * p[7] * uu * vv + p[6] * (q2 + 2 * vv * vv) + p[11] * q2
        return p[0] * (uu + du) + p[2], p[1] * (vv + dv) + p[3]
    raise NotImplementedError(f"synthetic generator does not draw {model!r}")


def make_scene(
    num_images: int,
    num_points: int,
    obs_per_point: int,
    model: CameraModel = CameraModel.PINHOLE,
    seed: int = 0,
    *,
    noise_px: float = 1.0,
    outlier_frac: float = 0.05,
    depth_prior: bool = True,
    depth_outlier_frac: float = 0.05,
    pose_rot_noise: float = 0.01,
    pose_pos_noise: float = 0.05,
    point_noise: float = 0.1,
    reproj_loss: LossType = LossType.SOFT_L1,
    depth_loss: LossType = LossType.CAUCHY,
    reproj_loss_scale: float = 1.5,
    kp_std: float = 1.0,
    const_point_frac: float = 0.0,
    depth_frac: float = 1.0,
    return_truth: bool = False,
):
    """Draw one scene; returns a :class:`BundleProblem` (and the ground truth if asked)."""
    C, P, L = int(num_images), int(num_points), int(obs_per_point)
    if L > C:
        raise ValueError("obs_per_point cannot exceed the number of images")
    rng = np.random.Generator(np.random.PCG64(seed))
    model = CameraModel(model)
    params = np.asarray(DEFAULT_PARAMS[model], dtype=np.float64)

    # ---- true images on a perturbed ring of radius 40 looking at the origin
    ring = 40.0
    theta = 2 * np.pi * np.arange(C) / C
    centre = np.stack([ring * np.cos(theta), 1.5 * np.sin(3 * theta), ring * np.sin(theta)], axis=1)
    centre += rng.normal(0, 0.3, size=(C, 3))
    look = -centre + rng.normal(0, 1.0, size=(C, 3))
    zc = look / np.linalg.norm(look, axis=1, keepdims=True)
    up = np.array([0.0, 1.0, 0.0])
    xc = np.cross(np.broadcast_to(up, zc.shape), zc)
    xc /= np.linalg.norm(xc, axis=1, keepdims=True)
    yc = np.cross(zc, xc)
    R_true = np.stack([xc, yc, zc], axis=1)  # rows = camera axes: p_cam = R X + t
    t_true = -np.einsum("nij,nj->ni", R_true, centre)
    q_true = quat_from_rotmat(R_true)

    # ---- true points in a box around the origin
    X_true = rng.uniform(-10.0, 10.0, size=(P, 3))

    # ---- tracks: a window of ring-neighbouring images per point
    phi = np.arctan2(X_true[:, 2], X_true[:, 0])
    near = np.rint(phi / (2 * np.pi) * C).astype(np.int64)
    jitter_w = max(1, min(C // 8, 12))
    near += rng.integers(-jitter_w, jitter_w + 1, size=P)
    max_stride = max(1, min(3, (C - 1) // max(L - 1, 1)))
    stride = rng.integers(1, max_stride + 1, size=P)
    k = np.arange(L)[None, :] - L // 2
    obs_img_pm = np.mod(near[:, None] + stride[:, None] * k, C).astype(np.int32)  # [P,L] point-major
    obs_pt_pm = np.repeat(np.arange(P, dtype=np.int32), L).reshape(P, L)

    # natural order = image-major, the order an extraction loop over images produces
    order = np.argsort(obs_img_pm.reshape(-1), kind="stable")
    obs_img = obs_img_pm.reshape(-1)[order]
    obs_pt = obs_pt_pm.reshape(-1)[order]
    del obs_img_pm, obs_pt_pm, order
    N = obs_img.shape[0]

    # ---- observations from the truth
    Rm = rotmat_from_quat(q_true)
    pc = np.einsum("nij,nj->ni", Rm[obs_img], X_true[obs_pt]) + t_true[obs_img]
    z_true = pc[:, 2].copy()
    x, y = _distort(model, params, pc[:, 0] / pc[:, 2], pc[:, 1] / pc[:, 2])
    del pc
    xy = np.stack([x, y], axis=1) + rng.normal(0, noise_px, size=(N, 2))
    out = rng.random(N) < outlier_frac
    xy[out] += rng.uniform(-50.0, 50.0, size=(int(out.sum()), 2))

    # ---- depth priors  d = z exp(eps), v = (sigma d)^2, w = d^2/v, a = rob_std sqrt(v)/d
    if depth_prior:
        if depth_frac >= 1.0:
            depth_obs = np.arange(N, dtype=np.int32)
        else:
            depth_obs = np.flatnonzero(rng.random(N) < depth_frac).astype(np.int32)
        nd = depth_obs.shape[0]
        sig = rng.uniform(0.03, 0.08, size=nd)
        d = z_true[depth_obs] * np.exp(rng.normal(0, 1.0, size=nd) * sig)
        gross = rng.random(nd) < depth_outlier_frac
        d[gross] *= rng.uniform(0.5, 2.0, size=int(gross.sum()))
        v = (sig * d) ** 2
        depth_w = d * d / np.clip(v, 1e-6, None)
        depth_a = 2.0 * np.sqrt(v) / d
    else:
        depth_obs = np.zeros(0, np.int32)
        d = depth_w = depth_a = np.zeros(0)

    # ---- start point: perturbed truth
    dR = _rodrigues(rng.normal(0, pose_rot_noise, size=(C, 3)))
    R0 = dR @ R_true
    c0 = centre + rng.normal(0, pose_pos_noise, size=(C, 3))
    t0 = -np.einsum("nij,nj->ni", R0, c0)
    q0 = quat_from_rotmat(R0)
    X0 = X_true + rng.normal(0, point_noise, size=(P, 3))

    free = np.full(C, FREE_ALL, dtype=np.uint8)
    free[0] = FREE_NONE
    if C > 1:
        free[1] = FREE_FIX_TX
    pt_const = (rng.random(P) < const_point_frac).astype(np.uint8) if const_point_frac > 0 else np.zeros(P, np.uint8)

    prob = BundleProblem(
        img_q=q0,
        img_t=t0,
        img_camera=np.zeros(C, np.int32),
        img_free_mask=free,
        cam_model=np.array([int(model)], np.int32),
        cam_params=params[None, :],
        pt_xyz=X0,
        pt_const=pt_const,
        obs_img=obs_img,
        obs_pt=obs_pt,
        obs_xy=xy,
        depth_obs=depth_obs,
        depth_d=d,
        depth_w=depth_w,
        depth_a=depth_a,
        shift_scale=np.zeros((C, 2)),
        reproj_loss=reproj_loss,
        reproj_loss_scale=reproj_loss_scale * kp_std,
        reproj_loss_weight=1.0 / kp_std**2,
        depth_loss=depth_loss,
    ).normalized()
    if return_truth:
        return prob, {"q": q_true, "t": t_true, "X": X_true}
    return prob


def make_config(name: str, seed: int = 0, **kw) -> BundleProblem:
    """Scenes by BASELINE.json config name (see ``CONFIGS``)."""
    C, P, L = CONFIGS[name]
    if name in ("cfg3", "cfg4", "cfg5", "target10m"):
        kw.setdefault("model", CameraModel.SIMPLE_RADIAL)
    if name == "cfg3":
        kw.setdefault("reproj_loss", LossType.CAUCHY)
    return make_scene(C, P, L, seed=seed, **kw)
