"""Residual blocks of the reference's Ceres problem, with local Jacobians.  TEST INFRASTRUCTURE ONLY.

Restates (file:line under /root/reference/mpsfm/sfm/mapper/):

* reprojection blocks built by ``pycolmap.create_default_bundle_adjuster`` (bundle_adjustment.py:104):
  ``r = ImgFromCam(params, q * X + t) - x_obs`` (COLMAP ``ReprojErrorCostFunctor``), loss
  ``ScaledLoss(loss(reproj_loss_scale * kp_std), 1 / kp_std^2)`` (:98-102);
* depth-prior blocks built by the fork-only ``pycolmap.create_depth_bundle_adjuster`` (:163-176) with
  ``logloss=True``: ``r = log(exp(s) d + b) - log((q * X + t)_z)``, one ``ScaledLoss(loss(a_j), w_j)`` per
  observation (:160-161).  The functor's source is not available; this form is SURVEY.md section 8
  row a-4's inference from the call site and ``optimize_prior_shiftscale`` (:236-241) -- PARITY UNPINNED;
* EXTENSION, nothing upstream to compare with ("parity n/a"): point-to-plane normal-prior
  rows ``r = m . (q * X + t) - c`` (``problem.plane_*``; m = sqrt(w_n) n, c = m . prior surface point), loss
  ``ScaledLoss(depth_loss(plane_loss_scale), 1)``.  This file DEFINES them; the kernels restate it;
* local parameterisations (:114-122): tangent of an image = (half-angle rotation vector, translation),
  Jacobians formed the Ceres way, ambient autodiff Jacobian times ``PlusJacobian``;
* ``ceres::internal::Corrector`` applied per residual block; cost = 1/2 sum rho.

Arrays in, arrays out, observation order = the problem's natural order.
"""

from __future__ import annotations

import dataclasses

import numpy as np

from . import camera_models, geometry, losses
from .jet import Jet


@dataclasses.dataclass
class State:
    q: np.ndarray  # [C,4] x,y,z,w
    t: np.ndarray  # [C,3]
    X: np.ndarray  # [P,3]

    @staticmethod
    def of(problem) -> "State":
        q = np.array(problem.img_q, dtype=np.float64)
        q /= np.linalg.norm(q, axis=1, keepdims=True)  # COLMAP normalises before AddResidualBlock
        return State(q, np.array(problem.img_t, dtype=np.float64), np.array(problem.pt_xyz, dtype=np.float64))

    def copy(self) -> "State":
        return State(self.q.copy(), self.t.copy(), self.X.copy())


@dataclasses.dataclass
class Linearization:
    r_reproj: np.ndarray  # [N,2] raw residuals
    r_depth: np.ndarray  # [N_d] raw residuals
    cost: float  # 1/2 sum rho over blocks with at least one non-constant parameter block
    fixed_cost: float  # blocks whose parameter blocks are all constant (Ceres removes them)
    # robustified rows per observation (row 2 = depth prior; zero where the observation has none; with normal-prior
    # rows in the problem there is a row 3 = point-to-plane)
    r: np.ndarray | None = None  # [N,3 or 4]
    Jc: np.ndarray | None = None  # [N,3 or 4,6] w.r.t. image tangent (masked columns zeroed)
    Jp: np.ndarray | None = None  # [N,3 or 4,3] w.r.t. point (zero for constant points)
    r_plane: np.ndarray | None = None  # [N_n] raw point-to-plane residuals (extension)


def free_columns(problem) -> np.ndarray:
    """[C,6] bool: tangent coordinates of each image that are optimised."""
    m = np.asarray(problem.img_free_mask, dtype=np.uint8)
    return ((m[:, None] >> np.arange(6)[None, :]) & 1).astype(bool)


def _cam_point_jets(q, X, t):
    """Ambient Jacobian d(q*X + t)/dq [n,3,4] via forward duals over Eigen's rotation formula."""
    n = q.shape[0]
    qx, qy, qz, qw = Jet.variables(q)

    def cross(a, b):
        return [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]

    u = [qx, qy, qz]
    v = [X[:, 0], X[:, 1], X[:, 2]]
    uv = cross(u, v)
    uv = [c + c for c in uv]
    uuv = cross(u, uv)
    p = [qw * uv[i] + uuv[i] + v[i] + t[:, i] for i in range(3)]
    val = np.stack([c.a for c in p], axis=1)
    jac = np.stack([c.v for c in p], axis=1)  # [n,3,4]
    assert jac.shape == (n, 3, 4)
    return val, jac


def evaluate(problem, state: State | None = None, jacobians: bool = True) -> Linearization:
    st = state or State.of(problem)
    oi, op = problem.obs_img, problem.obs_pt
    N = oi.shape[0]
    q, t, X = st.q[oi], st.t[oi], st.X[op]

    if jacobians:
        pc, dp_dq = _cam_point_jets(q, X, t)
    else:
        pc = geometry.quat_rotate(q, X) + t

    cam = problem.img_camera[oi]
    model, params = problem.cam_model[cam], problem.cam_params[cam]
    if jacobians:
        xy, dxy_dp = camera_models.img_from_cam(model, params, pc)
    else:
        xy = camera_models.img_from_cam_values(model, params, pc)
    r_rep = xy - problem.obs_xy
    # COLMAP >= 3.11 ReprojErrorCostFunctor (upstream src/colmap/estimators/cost_functions.h; the generation of COLMAP
    # that has create_default_bundle_adjuster, bundle_adjustment.py:104): ImgFromCam returns false for a point at
    # z < epsilon in the camera frame and the functor then writes ZERO residuals -- jets included, so the block's
    # Jacobian is zero too.  (Unverifiable here: the fork is absent.  PARITY UNPINNED like the rest of this file.)
    behind = pc[:, 2] < np.finfo(np.float64).eps
    if behind.any():
        r_rep[behind] = 0.0
        if jacobians:
            dxy_dp[behind] = 0.0

    # depth prior rows
    nd = problem.depth_obs.shape[0]
    di = problem.depth_obs
    b, s = problem.shift_scale[oi[di], 0], problem.shift_scale[oi[di], 1]
    z = pc[di, 2]
    r_dep = np.log(np.exp(s) * problem.depth_d + b) - np.log(z)

    # robust losses, one per residual block
    s_rep = np.einsum("ni,ni->n", r_rep, r_rep)
    rho_rep = losses.scaled_rho(int(problem.reproj_loss), problem.reproj_loss_scale, problem.reproj_loss_weight, s_rep)
    s_dep = r_dep * r_dep
    rho_dep = losses.scaled_rho(int(problem.depth_loss), problem.depth_a, problem.depth_w, s_dep)

    img_const = np.asarray(problem.img_free_mask)[oi] == 0
    pt_const = np.asarray(problem.pt_const)[op] != 0
    fixed = img_const & pt_const
    cost = 0.5 * float(np.sum(rho_rep[0][~fixed])) + 0.5 * float(np.sum(rho_dep[0][~fixed[di]]))
    fixed_cost = 0.5 * float(np.sum(rho_rep[0][fixed])) + 0.5 * float(np.sum(rho_dep[0][fixed[di]]))
    # normal-prior rows (extension)
    nn = getattr(problem, "num_plane", 0)
    r_pl = np.zeros(0)
    if nn:
        ni = np.asarray(problem.plane_obs)
        pm = np.asarray(problem.plane_m, dtype=np.float64)
        r_pl = np.einsum("ni,ni->n", pm, pc[ni]) - np.asarray(problem.plane_c, dtype=np.float64)
        s_pl = r_pl * r_pl
        rho_pl = losses.scaled_rho(int(problem.depth_loss), np.full(nn, float(problem.plane_loss_scale)), np.ones(nn), s_pl)
        cost += 0.5 * float(np.sum(rho_pl[0][~fixed[ni]]))
        fixed_cost += 0.5 * float(np.sum(rho_pl[0][fixed[ni]]))
    lin = Linearization(r_rep, r_dep, cost, fixed_cost, r_plane=r_pl)
    if not jacobians:
        return lin

    # ---- local Jacobians: ambient x PlusJacobian (image), plain (point)
    plusJ = geometry.quat_plus_jacobian(q)  # [N,4,3]
    dp_drot = np.einsum("nij,njk->nik", dp_dq, plusJ)  # [N,3,3]
    R = geometry.rotmat(q)  # d(q*X+t)/dX; equals the jet derivative for unit q
    rows = 4 if nn else 3
    Jc = np.zeros((N, rows, 6))
    Jp = np.zeros((N, rows, 3))
    Jc[:, :2, :3] = np.einsum("nij,njk->nik", dxy_dp, dp_drot)
    Jc[:, :2, 3:] = dxy_dp
    Jp[:, :2, :] = np.einsum("nij,njk->nik", dxy_dp, R)
    # depth row: dr/dp = (0, 0, -1/z)
    dz = -1.0 / z
    Jc[di, 2, :3] = dz[:, None] * dp_drot[di, 2, :]
    Jc[di, 2, 5] = dz
    Jp[di, 2, :] = dz[:, None] * R[di, 2, :]

    # ---- corrector (per block)
    r = np.zeros((N, rows))
    sc, alpha_sq, sqrt_r1 = losses.corrector(s_rep, rho_rep[1], rho_rep[2])
    # J <- sqrt(rho') (J - alpha_sq r r^T J)
    rJc = np.einsum("ni,nij->nj", r_rep, Jc[:, :2, :])
    rJp = np.einsum("ni,nij->nj", r_rep, Jp[:, :2, :])
    Jc[:, :2, :] = sqrt_r1[:, None, None] * (Jc[:, :2, :] - alpha_sq[:, None, None] * r_rep[:, :, None] * rJc[:, None, :])
    Jp[:, :2, :] = sqrt_r1[:, None, None] * (Jp[:, :2, :] - alpha_sq[:, None, None] * r_rep[:, :, None] * rJp[:, None, :])
    r[:, :2] = sc[:, None] * r_rep
    scd, alpha_sq_d, sqrt_r1_d = losses.corrector(s_dep, rho_dep[1], rho_dep[2])
    Jc[di, 2, :] = sqrt_r1_d[:, None] * (Jc[di, 2, :] * (1.0 - alpha_sq_d * s_dep)[:, None])
    Jp[di, 2, :] = sqrt_r1_d[:, None] * (Jp[di, 2, :] * (1.0 - alpha_sq_d * s_dep)[:, None])
    r[di, 2] = scd * r_dep
    if nn:
        # dr/dp = m; through the same chain as the other rows
        Jc[ni, 3, :3] = np.einsum("ni,nij->nj", pm, dp_drot[ni])
        Jc[ni, 3, 3:] = pm
        Jp[ni, 3, :] = np.einsum("ni,nij->nj", pm, R[ni])
        scn, alpha_sq_n, sqrt_r1_n = losses.corrector(s_pl, rho_pl[1], rho_pl[2])
        Jc[ni, 3, :] = sqrt_r1_n[:, None] * (Jc[ni, 3, :] * (1.0 - alpha_sq_n * s_pl)[:, None])
        Jp[ni, 3, :] = sqrt_r1_n[:, None] * (Jp[ni, 3, :] * (1.0 - alpha_sq_n * s_pl)[:, None])
        r[ni, 3] = scn * r_pl

    # ---- constant parameter blocks / fixed tangent coordinates carry no columns
    Jc *= free_columns(problem)[oi][:, None, :]
    Jp[pt_const] = 0.0
    # blocks Ceres removed from the program contribute nothing to the normal equations
    r[fixed] = 0.0
    lin.r, lin.Jc, lin.Jp = r, Jc, Jp
    return lin
