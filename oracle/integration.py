"""Depth-map integration with normal priors.  TEST INFRASTRUCTURE ONLY.

CPU restatement (numpy, fp64) of the reference's per-image solver, the step right before every bundle adjustment
(/root/reference/mpsfm/sfm/mapper/base.py:637-644 -> mpsfm/sfm/scene/image/integration.py):

  Integration._integrate          integration.py:383-520   IRLS loop, termination rules, state (integrated, energy_old)
  process_depth_prior             :236-260                 log-depth prior and its precision
  process_normals_prior           :262-279                 (nx, ny, nz) = (n[1], n[0], -n[2]) and their variances
  process_sparse_depth            :281-288, :391-397       sparse log-depths of the triangulated points + scale filter
  init_int_vars                   :314-356                 nz_u, nz_v, normal-constraint precisions
  generate_dx_dy                  :631-680                 the four one-sided difference operators A1..A4
  update_W / calc_Wpm             :358-376                 sigmoid discontinuity weights
  calc_energy                     :139-165
  calc_Amat, b_vec                :167-234, :445-453       normal equations of one IRLS step (5-point stencil)
  cg (scipy.sparse.linalg.cg, the reference's CPU library: mpsfm/utils/integration.py:17-23)  Jacobi-preconditioned
  calculate_hessian + IntegrationUncertainty.solve  :522-574, :51-79   variance propagation

The reference builds scipy/cupy CSR matrices; on a full H x W grid (normal_mask is all ones, camera.py:55) they are
5-point stencils, which is how they are written here (and in the CUDA kernels):
  (A1 z)_i = nz_u_i (z_{i-W} - z_i)  rows r >= 1      "top"      (D_vertical_pos)
  (A2 z)_i = nz_u_i (z_i - z_{i+W})  rows r <= H-2    "bottom"   (D_vertical_neg)
  (A3 z)_i = nz_v_i (z_{i+1} - z_i)  cols c <= W-2    "right"    (D_horizontal_pos)
  (A4 z)_i = nz_v_i (z_i - z_{i-1})  cols c >= 1      "left"     (D_horizontal_neg)
PINNED: tests/golden/integration.npz is written by tests/golden/make_golden_integration.py, which imports the
reference's integration.py UNMODIFIED (its own CPU mode: scipy, because no CUDA device is visible there) and records
inputs and outputs; tests/test_integration_oracle.py holds this file to them.
"""

from __future__ import annotations

import dataclasses

import numpy as np


@dataclasses.dataclass
class IntegrationConf:
    """The integration keys of Image.default_conf (mpsfm/sfm/scene/image/base.py:36-53)."""

    large_number: float = 1e6
    max_iter: int = 10
    tol: float = 5e-2
    step_size: float = 1
    cg_max_iter: int = 5000
    cg_tol: float = 1e-3
    lambda1: float = 1
    lambda2: float = 1
    k: float = 1
    depth_magnitude_multiplier: float = 1
    normals_magnitude_multiplier: float = 1
    scale_filter: bool = True
    scale_filter_factor: float = 1.5


@dataclasses.dataclass
class IntegrationState:
    """What the reference keeps on the Image between calls (IntVars, integration.py:18-29)."""

    integrated: bool = False
    energy_old: float = 0.0
    wu: np.ndarray | None = None
    wv: np.ndarray | None = None


def sigmoid(x, k):
    """mpsfm/utils/integration.py:52-55"""
    return 1.0 / (1.0 + np.exp(np.clip(-k * x, -709, 709)))


class Stencil:
    """The four difference operators on an H x W grid (generate_dx_dy) and their transposes."""

    def __init__(self, H, W, nz_u, nz_v, step_size=1):
        self.H, self.W = H, W
        self.nu = (nz_u / step_size).reshape(H, W)
        self.nv = (nz_v / step_size).reshape(H, W)

    def apply(self, z):
        H, W = self.H, self.W
        z = z.reshape(H, W)
        a1, a2, a3, a4 = (np.zeros((H, W)) for _ in range(4))
        a1[1:, :] = self.nu[1:, :] * z[:-1, :] + (-self.nu[1:, :]) * z[1:, :]       # CSR order: (self, top) -> -nz z_i + nz z_top
        a2[:-1, :] = (-self.nu[:-1, :]) * z[1:, :] + self.nu[:-1, :] * z[:-1, :]
        a3[:, :-1] = (-self.nv[:, :-1]) * z[:, :-1] + self.nv[:, :-1] * z[:, 1:]
        a4[:, 1:] = (-self.nv[:, 1:]) * z[:, :-1] + self.nv[:, 1:] * z[:, 1:]
        return a1.reshape(-1), a2.reshape(-1), a3.reshape(-1), a4.reshape(-1)

    def transpose_sum(self, v1, v2, v3, v4):
        """A1^T v1 + A2^T v2 + A3^T v3 + A4^T v4"""
        H, W = self.H, self.W
        v1, v2, v3, v4 = (v.reshape(H, W) for v in (v1, v2, v3, v4))
        out = np.zeros((H, W))
        t = self.nu * v1
        out[1:, :] -= t[1:, :]
        out[:-1, :] += t[1:, :]
        t = self.nu * v2
        out[:-1, :] += t[:-1, :]
        out[1:, :] -= t[:-1, :]
        t = self.nv * v3
        out[:, :-1] -= t[:, :-1]
        out[:, 1:] += t[:, :-1]
        t = self.nv * v4
        out[:, 1:] += t[:, 1:]
        out[:, :-1] -= t[:, 1:]
        return out.reshape(-1)


def normal_equations(H, W, nz_u, nz_v, wu_plus, wu_minus, wv_plus, wv_minus, depth_precision, lambda1, sp_ids, sp_prec,
                     lambda2, with_sparse=True):
    """calc_Amat (:167-234) as stencil coefficients: diag [n], e_r [n] = A(i, i+1), e_d [n] = A(i, i+W)."""
    sh = (H, W)
    nu2, nv2 = (nz_u ** 2).reshape(sh), (nz_v ** 2).reshape(sh)
    top = np.zeros(sh)
    bottom = np.zeros(sh)
    left = np.zeros(sh)
    right = np.zeros(sh)
    top[1:, :] = wu_plus.reshape(sh)[1:, :] * nu2[1:, :]
    bottom[:-1, :] = wu_minus.reshape(sh)[:-1, :] * nu2[:-1, :]
    left[:, 1:] = wv_minus.reshape(sh)[:, 1:] * nv2[:, 1:]
    right[:, :-1] = wv_plus.reshape(sh)[:, :-1] * nv2[:, :-1]
    diag = np.zeros(sh)
    # the reference's order of the eight += (:191-198)
    diag[:, 1:] += left[:, 1:]
    diag[:, :-1] += left[:, 1:]
    diag[:, :-1] += right[:, :-1]
    diag[:, 1:] += right[:, :-1]
    diag[1:, :] += top[1:, :]
    diag[:-1, :] += top[1:, :]
    diag[:-1, :] += bottom[:-1, :]
    diag[1:, :] += bottom[:-1, :]
    diag = diag.reshape(-1)
    diag = diag + lambda1 * depth_precision
    if with_sparse and len(sp_ids) > 0:
        diag[sp_ids] += lambda2 * sp_prec  # numpy fancy +=: a pixel hit twice keeps the LAST contribution only
    e_r = np.zeros(sh)
    e_d = np.zeros(sh)
    e_r[:, :-1] = (-right[:, :-1]) + (-left[:, 1:])    # A_odu = top + bottom + right + left (:229)
    e_d[:-1, :] = (-top[1:, :]) + (-bottom[:-1, :])
    return diag, e_r.reshape(-1), e_d.reshape(-1)


def stencil_matvec(H, W, diag, e_r, e_d, p):
    """A p in scipy's CSR order of one row: columns i-W, i-1, i, i+1, i+W."""
    sh = (H, W)
    P, D, ER, ED = p.reshape(sh), diag.reshape(sh), e_r.reshape(sh), e_d.reshape(sh)
    q = np.zeros(sh)
    q[1:, :] += ED[:-1, :] * P[:-1, :]
    q[:, 1:] += ER[:, :-1] * P[:, :-1]
    q += D * P
    q[:, :-1] += ER[:, :-1] * P[:, 1:]
    q[:-1, :] += ED[:-1, :] * P[1:, :]
    return q.reshape(-1)


def cg(H, W, diag, e_r, e_d, b, x0, minv, maxiter, rtol):
    """scipy.sparse.linalg.cg (1.12+ signature: rtol, atol = 0) with M = diag(minv), warm start x0.
    Returns (x, iterations)."""
    x = x0.copy()
    bnrm2 = np.linalg.norm(b)
    atol = max(0.0, rtol * bnrm2)
    if bnrm2 == 0:
        return b.copy(), 0
    r = b - stencil_matvec(H, W, diag, e_r, e_d, x) if x.any() else b.copy()
    rho_prev, p = None, None
    for it in range(maxiter):
        if np.linalg.norm(r) < atol:
            return x, it
        z = minv * r
        rho = float(np.dot(r, z))
        if it > 0:
            p = p * (rho / rho_prev) + z
        else:
            p = z.copy()
        q = stencil_matvec(H, W, diag, e_r, e_d, p)
        alpha = rho / float(np.dot(p, q))
        x = x + alpha * p
        r = r - alpha * q
        rho_prev = rho
    return x, maxiter


@dataclasses.dataclass
class Prepared:
    H: int
    W: int
    z0: np.ndarray
    z_prior: np.ndarray
    depth_precision: np.ndarray
    nx: np.ndarray
    ny: np.ndarray
    nz_u: np.ndarray
    nz_v: np.ndarray
    Nu_precision: np.ndarray
    Nv_precision: np.ndarray
    sp_ids: np.ndarray
    sp_prec: np.ndarray
    sp_depth: np.ndarray


def prepare(depth_prior, uncertainty, valid, normals, normals_uncertainty, depth_init, kps, depth3d, zvars3d, K,
            conf: IntegrationConf) -> Prepared:
    """Everything of _integrate that does not depend on z (:386-412).  K = the four numbers the reference passes
    (:124-129: [K11 sy, K00 sx, K12 sy, K02 sx], named fx, fy, cx, cy in :384)."""
    H, W = depth_prior.shape
    k0, k1, k2, k3 = (float(v) for v in K)
    depth_prior = np.asarray(depth_prior, dtype=np.float64)
    depth_precision = conf.depth_magnitude_multiplier * np.asarray(1 / (uncertainty + 1e-6), dtype=np.float64)
    depth_precision = (depth_precision * depth_prior ** 2).reshape(-1)
    z_prior = np.log(depth_prior).reshape(-1)
    normals = np.asarray(normals, dtype=np.float64)
    nx, ny, nz = normals[..., 1].reshape(-1), normals[..., 0].reshape(-1), -normals[..., 2].reshape(-1)
    nu = np.array(normals_uncertainty, dtype=np.float64)
    nu[~np.asarray(valid, dtype=bool)] = conf.large_number  # every entry of the 3x3 block (:275)
    inv_m = 1 / conf.normals_magnitude_multiplier
    Vnx, Vny, Vnz = inv_m * nu[..., 1, 1].reshape(-1), inv_m * nu[..., 0, 0].reshape(-1), inv_m * nu[..., 2, 2].reshape(-1)
    z0 = np.log(np.asarray(depth_init, dtype=np.float64).reshape(-1))
    kps = np.asarray(kps).reshape(-1, 2)
    if len(kps):
        sp_ids = np.ravel_multi_index((kps[:, 1], kps[:, 0]), (H, W))
        depth3d = np.asarray(depth3d, dtype=np.float64)
        sp_prec = (np.asarray(1 / np.asarray(zvars3d, dtype=np.float64)) * depth3d ** 2).reshape(-1)
        sp_depth = np.log(depth3d).reshape(-1)
        if conf.scale_filter:
            div = np.exp(sp_depth) / np.exp(z_prior[sp_ids])
            ok = (div < conf.scale_filter_factor) * (div > 1 / conf.scale_filter_factor)
            sp_ids, sp_prec, sp_depth = sp_ids[ok], sp_prec[ok], sp_depth[ok]
    else:
        sp_ids, sp_prec, sp_depth = np.zeros(0, np.int64), np.zeros(0), np.zeros(0)
    rr, cc = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    uu = (H - 1 - rr).reshape(-1) - k2  # xx flipped along axis 0 (:321-323)
    vv = cc.reshape(-1) - k3
    nz_u = uu * nx + vv * ny + k0 * nz
    nz_v = uu * nx + vv * ny + k1 * nz
    Duz, Dvz = -nx / nz_u, -ny / nz_v
    one = np.ones_like(z0)
    Nu = 1 / (Vnx * ((uu * Duz + one) ** 2) + Vny * (vv * Duz) ** 2 + k0 ** 2 * Vnz * Duz ** 2)
    Nv = 1 / (Vnx * (uu * Dvz) ** 2 + Vny * (vv * Dvz + one) ** 2 + k1 ** 2 * Vnz * Dvz ** 2)
    return Prepared(H, W, z0, z_prior, depth_precision, nx, ny, nz_u, nz_v, Nu, Nv, sp_ids, sp_prec, sp_depth)


def weights(st: Stencil, z, k):
    """update_W (:358-364)"""
    a1, a2, a3, a4 = st.apply(z)
    return sigmoid(a2 ** 2 - a1 ** 2, k), sigmoid(a4 ** 2 - a3 ** 2, k)


def energy_of(st: Stencil, pr: Prepared, conf, z, wu, wv):
    """calc_Wpm + calc_energy (:366-376, :139-165)"""
    wu_plus, wu_minus = wu * pr.Nu_precision, (1 - wu) * pr.Nu_precision
    wv_plus, wv_minus = wv * pr.Nv_precision, (1 - wv) * pr.Nv_precision
    a1, a2, a3, a4 = st.apply(z)
    e = np.sum(wu_plus * (a1 + pr.nx) ** 2 + wu_minus * (a2 + pr.nx) ** 2 + wv_plus * (a3 + pr.ny) ** 2
               + wv_minus * (a4 + pr.ny) ** 2)
    e += np.sum(conf.lambda1 * pr.depth_precision * (pr.z_prior - z) ** 2)
    if len(pr.sp_ids) > 0:
        e += np.sum(conf.lambda2 * pr.sp_prec * (pr.sp_depth - z[pr.sp_ids]) ** 2)
    return float(e), (wu_plus, wu_minus, wv_plus, wv_minus)


def integrate(pr: Prepared, conf: IntegrationConf, state: IntegrationState | None = None, trace: dict | None = None):
    """Integration._integrate (:383-520).  Returns (updated: bool, depth [H,W] or None, new state).
    trace (optional dict) receives energies, cg_iterations and z after every IRLS step."""
    state = state or IntegrationState()
    H, W = pr.H, pr.W
    st = Stencil(H, W, pr.nz_u, pr.nz_v, conf.step_size)
    z = pr.z0.copy()
    if state.integrated:
        wu, wv = state.wu.copy(), state.wv.copy()  # init and self.integrated: cached weights (:338-340)
    else:
        wu, wv = weights(st, z, conf.k)
    energy, (wu_plus, wu_minus, wv_plus, wv_minus) = energy_of(st, pr, conf, z, wu, wv)
    if trace is not None:
        trace.update(energies=[energy], cg_iterations=[], z_steps=[])
    if state.integrated and not (abs(energy - state.energy_old) / state.energy_old > conf.tol):
        return False, None, IntegrationState(True, state.energy_old, wu, wv)
    energy_0 = min_energy = energy
    nu_, nv_ = pr.nz_u / conf.step_size, pr.nz_v / conf.step_size
    for _ in range(conf.max_iter):
        diag, e_r, e_d = normal_equations(H, W, pr.nz_u, pr.nz_v, wu_plus, wu_minus, wv_plus, wv_minus, pr.depth_precision,
                                          conf.lambda1, pr.sp_ids, pr.sp_prec, conf.lambda2)
        b = st.transpose_sum(wu_plus * (-pr.nx), wu_minus * (-pr.nx), wv_plus * (-pr.ny), wv_minus * (-pr.ny))
        b = b + conf.lambda1 * pr.depth_precision * pr.z_prior
        if len(pr.sp_ids) > 0:
            b[pr.sp_ids] += conf.lambda2 * pr.sp_prec * pr.sp_depth  # last duplicate wins, like the diagonal
        minv = 1 / np.clip(diag, 1e-5, None)
        z, its = cg(H, W, diag, e_r, e_d, b, z, minv, conf.cg_max_iter, conf.cg_tol)
        wu, wv = weights(st, z, conf.k)
        energy_old = energy
        min_energy = min(energy, min_energy)
        energy, (wu_plus, wu_minus, wv_plus, wv_minus) = energy_of(st, pr, conf, z, wu, wv)
        if trace is not None:
            trace["energies"].append(energy)
            trace["cg_iterations"].append(its)
            trace["z_steps"].append(z.copy())
        rel = abs(energy - energy_old) / energy_old
        rel_min = abs(energy - min_energy) / min_energy
        if ((rel < conf.tol and (energy_old - energy) > 0) or (rel_min < conf.tol and (min_energy - energy) > 0)) \
                and energy < energy_0:
            break
        if energy > energy_0:
            return False, None, IntegrationState(True, energy_0, wu, wv)
    del nu_, nv_
    return True, np.exp(z.reshape(H, W)), IntegrationState(True, energy, wu, wv)


def hessian(pr: Prepared, conf: IntegrationConf, ignore_depths=True):
    """calculate_hessian (:522-574): the normal-equation matrix at the CURRENT depth (pr.z0 = log depth.data) with
    freshly computed weights; sparse term left out when ignore_depths.  NOTE: calculate_hessian does NOT apply the
    scale filter to the sparse depths (:542 has no counterpart of :391-397) -- build `pr` with
    conf.scale_filter = False when ignore_depths is False."""
    st = Stencil(pr.H, pr.W, pr.nz_u, pr.nz_v, conf.step_size)
    wu, wv = weights(st, pr.z0, conf.k)
    wu_plus, wu_minus = wu * pr.Nu_precision, (1 - wu) * pr.Nu_precision
    wv_plus, wv_minus = wv * pr.Nv_precision, (1 - wv) * pr.Nv_precision
    return normal_equations(pr.H, pr.W, pr.nz_u, pr.nz_v, wu_plus, wu_minus, wv_plus, wv_minus, pr.depth_precision,
                            conf.lambda1, pr.sp_ids, pr.sp_prec, conf.lambda2, with_sparse=not ignore_depths)


def variances_at(H, W, diag, e_r, e_d, xy):
    """IntegrationUncertainty.solve (:63-79): for every query pixel, solve H x = e_pixel and return x.sum() -- the
    COLUMN SUM of H^-1, not its diagonal (that is what `(x).sum(0)` computes).  H is symmetric, so all of them are
    entries of the single vector y = H^-1 1: variance(pixel) = y[pixel].  Exact sparse solve here (the reference
    factorises with cholespy in float32)."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl

    n = H * W
    idx = np.arange(n)
    c = idx % W
    rows = [idx, idx[c < W - 1], idx[c < W - 1] + 1, idx[: n - W], idx[: n - W] + W]
    cols = [idx, idx[c < W - 1] + 1, idx[c < W - 1], idx[: n - W] + W, idx[: n - W]]
    vals = [diag, e_r[c < W - 1], e_r[c < W - 1], e_d[: n - W], e_d[: n - W]]
    A = sp.csc_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n, n))
    y = spl.splu(A).solve(np.ones(n))
    xy = np.round(np.asarray(xy)).astype(int)
    return y[np.ravel_multi_index((xy[:, 1], xy[:, 0]), (H, W))]
