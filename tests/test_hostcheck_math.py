"""The product's per-observation arithmetic (ba_math.cuh built with g++) against the numpy oracle."""

import ctypes as C

import numpy as np
import pytest

from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene
from oracle import camera_models, geometry, lm
from oracle.residuals import State, evaluate, free_columns

import kernel_emulation as ke

MODELS5 = [CameraModel.SIMPLE_PINHOLE, CameraModel.PINHOLE, CameraModel.SIMPLE_RADIAL, CameraModel.RADIAL,
           CameraModel.OPENCV]


def rel(a, b):
    return np.max(np.abs(a - b) / (np.abs(b) + 1e-300 + 1e-12 * np.max(np.abs(b))))


@pytest.mark.parametrize("model", MODELS5)
@pytest.mark.parametrize("losses", [(LossType.SOFT_L1, LossType.CAUCHY), (LossType.CAUCHY, LossType.TRIVIAL),
                                    (LossType.HUBER, LossType.HUBER), (LossType.TRIVIAL, LossType.SOFT_L1)])
def test_observation_blocks_match_oracle(hostcheck, model, losses):
    prob = make_scene(8, 120, 4, model=model, seed=3, reproj_loss=losses[0], depth_loss=losses[1], depth_frac=0.7)
    prob.img_free_mask[:] = 0b111111  # compare unmasked blocks
    st = State.of(prob)
    lin = evaluate(prob, st, jacobians=True)
    q = ke.obs_quantities(hostcheck, prob, st.q, st.t, st.X)
    # raw residuals: 1e-9 relative gate of BASELINE.json
    assert rel(q["r"][:, :2], lin.r_reproj) < 1e-9
    assert rel(q["r"][prob.depth_obs, 2], lin.r_depth) < 1e-9
    assert abs(0.5 * q["rho"].sum() - (lin.cost + lin.fixed_cost)) <= 1e-12 * lin.cost
    Hcc = np.einsum("nki,nkj->nij", lin.Jc, lin.Jc)
    gc = np.einsum("nki,nk->ni", lin.Jc, lin.r)
    assert rel(q["Hcc"], Hcc) < 1e-9
    assert rel(q["gc"], gc) < 1e-9
    R = geometry.rotmat(st.q[prob.obs_img])
    Hpp = np.einsum("nki,nkl,nlj->nij", R, q["N"], R)
    assert rel(Hpp, np.einsum("nki,nkj->nij", lin.Jp, lin.Jp)) < 1e-9
    assert rel(np.einsum("nki,nk->ni", R, q["h"]), np.einsum("nki,nk->ni", lin.Jp, lin.r)) < 1e-9
    G = ke.g_matrix(q["prot"])
    W = np.einsum("nki,nkl,nlj->nij", G, q["N"], R)
    assert rel(W, np.einsum("nki,nkj->nij", lin.Jc, lin.Jp)) < 1e-9


def test_all_camera_models_projection_and_jacobian(hostcheck):
    rng = np.random.default_rng(0)
    params = {
        0: [900.0, 500.0, 400.0], 1: [900.0, 950.0, 500.0, 400.0], 2: [900.0, 500.0, 400.0, 0.05],
        3: [900.0, 500.0, 400.0, 0.05, -0.01], 4: [900.0, 950.0, 500.0, 400.0, 0.05, -0.01, 1e-3, -2e-3],
        5: [900.0, 950.0, 500.0, 400.0, 0.02, -0.01, 3e-3, -1e-3],
        6: [900.0, 950.0, 500.0, 400.0, 0.05, -0.01, 1e-3, -2e-3, 3e-3, 0.02, -4e-3, 1e-3],
        7: [900.0, 950.0, 500.0, 400.0, 0.9], 8: [900.0, 500.0, 400.0, 0.03], 9: [900.0, 500.0, 400.0, 0.03, -0.004],
        10: [900.0, 950.0, 500.0, 400.0, 0.02, -0.01, 1e-3, -2e-3, 3e-3, -1e-3, 2e-3, -3e-3],
    }
    for model, p in params.items():
        prm = np.zeros((1, 12))
        prm[0, : len(p)] = p
        for _ in range(20):
            u, v = rng.uniform(-0.6, 0.6, 2)
            pc = np.array([[u * 2.5, v * 2.5, 2.5]])
            xy_o, jac_o = camera_models.img_from_cam(np.array([model]), prm, pc)
            out = np.zeros(4)
            D = np.zeros(4)
            hostcheck.hc_project(C.c_int(model), prm.ctypes.data_as(C.c_void_p), C.c_double(u), C.c_double(v),
                                 out.ctypes.data_as(C.c_void_p), D.ctypes.data_as(C.c_void_p))
            assert np.allclose(out[:2], xy_o[0], rtol=1e-12, atol=1e-9), model
            assert np.allclose(out[2:], xy_o[0], rtol=1e-12, atol=1e-9), model
            # oracle Jacobian is w.r.t. the camera-frame point: d/dp = D [[1/z,0,-u/z],[0,1/z,-v/z]]
            z = 2.5
            A = np.array([[1 / z, 0, -u / z], [0, 1 / z, -v / z]])
            assert np.allclose(D.reshape(2, 2) @ A, jac_o[0], rtol=1e-9, atol=1e-9), model
    # FOV small-omega and small-radius branches
    for omega, uv in [(5e-3, (0.3, 0.2)), (0.9, (1e-3, 2e-3))]:
        prm = np.zeros((1, 12))
        prm[0, :5] = [900.0, 950.0, 500.0, 400.0, omega]
        pc = np.array([[uv[0], uv[1], 1.0]])
        xy_o, jac_o = camera_models.img_from_cam(np.array([7]), prm, pc)
        out, D = np.zeros(4), np.zeros(4)
        hostcheck.hc_project(C.c_int(7), prm.ctypes.data_as(C.c_void_p), C.c_double(uv[0]), C.c_double(uv[1]),
                             out.ctypes.data_as(C.c_void_p), D.ctypes.data_as(C.c_void_p))
        assert np.allclose(out[:2], xy_o[0], rtol=1e-12)
        A = np.array([[1, 0, -uv[0]], [0, 1, -uv[1]]])
        assert np.allclose(D.reshape(2, 2) @ A, jac_o[0], rtol=1e-9, atol=1e-9)


def test_small_helpers(hostcheck):
    rng = np.random.default_rng(1)
    for _ in range(10):
        A = rng.normal(size=(3, 3))
        H = A @ A.T + 0.1 * np.eye(3)
        h6 = np.array([H[0, 0], H[0, 1], H[0, 2], H[1, 1], H[1, 2], H[2, 2]])
        out = np.zeros(6)
        assert hostcheck.hc_sym3_inverse(h6.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)) == 1
        Hi = np.linalg.inv(H)
        assert np.allclose(out, [Hi[0, 0], Hi[0, 1], Hi[0, 2], Hi[1, 1], Hi[1, 2], Hi[2, 2]], rtol=1e-10)
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        d = rng.normal(size=3) * 0.1
        o = np.zeros(4)
        hostcheck.hc_quat_plus(q.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
        assert np.allclose(o, geometry.quat_plus(q[None], d[None])[0], rtol=1e-13, atol=1e-15)
        pa, pb = rng.normal(size=3), rng.normal(size=3)
        K = rng.normal(size=(3, 3))
        M = np.zeros(36)
        hostcheck.hc_gt_k_g(pa.ctypes.data_as(C.c_void_p), K.ctypes.data_as(C.c_void_p), pb.ctypes.data_as(C.c_void_p),
                            M.ctypes.data_as(C.c_void_p))
        Ga, Gb = ke.g_matrix(pa[None])[0], ke.g_matrix(pb[None])[0]
        assert np.allclose(M.reshape(6, 6), Ga.T @ K @ Gb, rtol=1e-13, atol=1e-13)
    neg = np.array([1.0, 2.0, 0.0, 1.0, 0.0, 1.0])  # not positive definite
    assert hostcheck.hc_sym3_inverse(neg.ctypes.data_as(C.c_void_p), np.zeros(6).ctypes.data_as(C.c_void_p)) == 0
    hostcheck.hc_lm_damping.restype = C.c_double
    val = hostcheck.hc_lm_damping(C.c_double(4.0), C.c_double(0.5), C.c_double(100.0), C.c_double(1e-6), C.c_double(1e32))
    assert np.isclose(val, (0.25 * 4.0) / (100.0 * 0.25))


def test_emulated_lm_step_matches_oracle(hostcheck):
    """One damped Newton step assembled the way the kernels do it == the oracle's exact Schur solve."""
    prob = make_scene(6, 60, 4, seed=5, depth_frac=0.8, const_point_frac=0.1)
    st = State.of(prob)
    lin = evaluate(prob, st, jacobians=True)
    oi, op = prob.obs_img, prob.obs_pt
    Cn, P = prob.num_images, prob.num_points
    q = ke.obs_quantities(hostcheck, prob, st.q, st.t, st.X)
    freec = free_columns(prob).astype(float)
    freep = (prob.pt_const == 0).astype(float)
    R = geometry.rotmat(st.q[oi])
    G = ke.g_matrix(q["prot"])
    Hcc = lm.segment_sum(oi, q["Hcc"], Cn) * freec[:, :, None] * freec[:, None, :]
    gc = lm.segment_sum(oi, q["gc"], Cn) * freec
    Hpp = lm.segment_sum(op, np.einsum("nki,nkl,nlj->nij", R, q["N"], R), P) * freep[:, None, None]
    gp = lm.segment_sum(op, np.einsum("nki,nk->ni", R, q["h"]), P) * freep[:, None]
    # Jacobi scaling + LM diagonal exactly as lm.solve's first iteration
    c_img = np.where(freec > 0, 1.0 / (1.0 + np.sqrt(np.einsum("cii->ci", Hcc))), 0.0)
    c_pt = np.where(freep[:, None] > 0, 1.0 / (1.0 + np.sqrt(np.einsum("pii->pi", Hpp))), 0.0)
    radius = 1e4
    ne = lm.NormalEquations(prob, lin, c_img, c_pt)
    dc, dp = ne.column_sq_norms()
    dc, dp = np.clip(dc, 1e-6, 1e32), np.clip(dp, 1e-6, 1e32)
    yc_o, yp_o = ne.solve(dc / radius, dp / radius)
    step_c_o, step_p_o = -yc_o * c_img, -yp_o * c_pt
    m = np.einsum("nki,ni->nk", ne.Jc, yc_o[oi]) + np.einsum("nki,ni->nk", ne.Jp, yp_o[op])
    mcc_o = float(np.sum(m * (ne.r - 0.5 * m)))

    # ---- kernel algebra in unscaled variables
    def damp(hjj, c):
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.where(c > 0, np.clip(c * c * hjj, 1e-6, 1e32) / (radius * c * c), 0.0)

    lam_c = damp(np.einsum("cii->ci", Hcc), c_img)
    lam_p = damp(np.einsum("pii->pi", Hpp), c_pt)
    act = (freep > 0) & (np.bincount(op, minlength=P) > 0)
    Hd = Hpp + np.einsum("pi,ij->pij", lam_p, np.eye(3))
    Hinv = np.zeros_like(Hd)
    Hinv[act] = np.linalg.inv(Hd[act])
    Q = np.einsum("nij,njk->nik", q["N"], R)
    W = np.einsum("nki,nkj->nij", G, Q) * freec[oi][:, :, None] * act[op][:, None, None]
    S = np.zeros((6 * Cn, 6 * Cn))
    rhs = gc.copy()
    for c in range(Cn):
        S[6 * c:6 * c + 6, 6 * c:6 * c + 6] = Hcc[c] + np.diag(lam_c[c])
    for p in np.flatnonzero(act):
        ks = np.flatnonzero(op == p)
        e = Hinv[p] @ gp[p]
        for a in ks:
            rhs[oi[a]] -= W[a] @ e
            for b in ks:
                S[6 * oi[a]:6 * oi[a] + 6, 6 * oi[b]:6 * oi[b] + 6] -= W[a] @ Hinv[p] @ W[b].T
    fixed = np.flatnonzero(freec.reshape(-1) == 0)
    S[fixed, :] = 0
    S[:, fixed] = 0
    S[fixed, fixed] = 1.0
    rhs = rhs.reshape(-1)
    rhs[fixed] = 0
    yc = np.linalg.solve(S, rhs).reshape(Cn, 6)
    ga = np.einsum("nki,ni->nk", G, yc[oi])
    Ng = np.einsum("nij,nj->ni", q["N"], ga)
    wty = lm.segment_sum(op, np.einsum("nki,nk->ni", R, Ng), P)
    yp = np.einsum("pij,pj->pi", Hinv, gp - wty)
    mcc = float(np.sum(np.einsum("ni,ni->n", ga, q["h"]) - 0.5 * np.einsum("ni,ni->n", ga, Ng)))
    mcc += float(np.sum(np.einsum("pi,pi->p", yp, gp - wty) - 0.5 * np.einsum("pi,pij,pj->p", yp, Hpp, yp)))
    assert np.allclose(-yc, step_c_o, rtol=1e-7, atol=1e-12)
    assert np.allclose(-yp, step_p_o, rtol=1e-7, atol=1e-12)
    assert abs(mcc - mcc_o) <= 1e-9 * abs(mcc_o)

    # ---- what k_points / k_back_substitute / k_cam_update do since the Schur factors are stored:
    #   Hpp + D = L L^T,  Z_a = N_a R_a L^-T,  e = L^-1 gp,  y_p = L^-T (e - sum_a Z_a^T G_a y_c),
    #   model cost change = (g'y + y'Dy) / 2  (the normal equations the step solves)
    Linv = np.zeros_like(Hd)
    Linv[act] = np.linalg.inv(np.linalg.cholesky(Hd[act]))
    Z = np.einsum("nij,nkj->nik", Q, Linv[op])  # Q L^-T
    e = np.einsum("pij,pj->pi", Linv, gp)
    ztg = lm.segment_sum(op, np.einsum("nki,nk->ni", Z, ga) * act[op][:, None], P)
    yp_f = np.einsum("pji,pj->pi", Linv, e - ztg)
    assert np.allclose(yp_f, yp, rtol=1e-9, atol=1e-13)
    mcc_f = 0.5 * (np.sum(gp * yp_f) + np.sum(lam_p * yp_f**2) + np.sum(gc * yc) + np.sum(lam_c * freec * yc**2))
    assert abs(mcc_f - mcc_o) <= 1e-9 * abs(mcc_o)
