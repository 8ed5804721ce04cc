"""The normal-prior rows (EXTENSION -- nothing upstream to hold them against, "parity n/a").

oracle/residuals.py DEFINES the rows; here (a) the definition is checked against finite differences of its own cost,
(b) the product's arithmetic (ba_math.cuh built with g++, tests/hostcheck) is held against that definition at the
1e-9 gate of the other rows, (c) the host side (undistortion, plane construction) against the forward camera models."""

import numpy as np
import pytest

from mpsfm_b200 import normal_prior
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import DEFAULT_PARAMS, _distort, make_scene
from oracle import geometry, lm
from oracle.residuals import State, evaluate

import kernel_emulation as ke


def rel(a, b):
    return np.max(np.abs(a - b) / (np.abs(b) + 1e-300 + 1e-12 * np.max(np.abs(b))))


def _scene(model=CameraModel.SIMPLE_RADIAL, depth_loss=LossType.CAUCHY, **kw):
    prob = make_scene(8, 150, 4, model=model, seed=4, depth_loss=depth_loss, depth_frac=0.8, **kw)
    return normal_prior.synthetic_normal_priors(prob, frac=0.6, seed=1)


@pytest.mark.parametrize("model", sorted(DEFAULT_PARAMS))
def test_cam_from_img_inverts_the_forward_model(model):
    rng = np.random.default_rng(int(model))
    uv = rng.uniform(-0.5, 0.5, size=(500, 2))
    uv[0] = 0.0
    p = np.asarray(DEFAULT_PARAMS[model])
    with np.errstate(invalid="ignore", divide="ignore"):
        x, y = _distort(model, p, np.where(uv[:, 0] == 0, 1e-13, uv[:, 0]), uv[:, 1])
    back = normal_prior.cam_from_img(model, p, np.stack([x, y], axis=1))
    assert np.max(np.abs(back - uv)) < 1e-11


def test_plane_rows_vanish_on_the_prior_surface():
    prob = _scene()
    assert prob.num_plane > 100 and np.all(np.isin(prob.plane_obs, prob.depth_obs))
    # a point at q = depth * (u, v, 1) of its own observation has residual 0 and lies at the prior depth
    o = prob.plane_obs
    k = np.searchsorted(prob.depth_obs, o)
    cam = prob.img_camera[prob.obs_img[o]]
    uv = normal_prior.cam_from_img(CameraModel(int(prob.cam_model[0])), prob.cam_params[0], prob.obs_xy[o])
    q = prob.depth_d[k][:, None] * np.concatenate([uv, np.ones((o.size, 1))], axis=1)
    assert np.all(cam == 0)
    assert np.max(np.abs(np.einsum("ni,ni->n", prob.plane_m, q) - prob.plane_c)) < 1e-9
    assert np.allclose(np.linalg.norm(prob.plane_m, axis=1), 1.0 / (0.05 * prob.depth_d[k]))
    with pytest.raises(ValueError):
        no_depth = np.setdiff1d(np.arange(prob.num_obs), prob.depth_obs)[:3]
        normal_prior.plane_rows(prob, no_depth, np.tile([0, 0, -1.0], (3, 1)), 1.0)


@pytest.mark.parametrize("depth_loss", [LossType.CAUCHY, LossType.TRIVIAL, LossType.HUBER])
def test_oracle_gradient_is_the_derivative_of_its_cost(depth_loss):
    prob = _scene(depth_loss=depth_loss)
    st = State.of(prob)
    lin = evaluate(prob, st, jacobians=True)
    base = evaluate(dataclass_without_planes(prob), st, jacobians=True)
    assert lin.r.shape[1] == 4 and base.r.shape[1] == 3
    assert lin.cost > base.cost and lin.r_plane.shape == (prob.num_plane,)
    assert np.array_equal(lin.r[:, :3], base.r[:, :3]) and np.array_equal(lin.Jc[:, :3], base.Jc[:, :3])
    # gradient of 0.5 sum rho w.r.t. one point / one image tangent by central differences
    oi, op = prob.obs_img, prob.obs_pt
    gp = lm.segment_sum(op, np.einsum("nki,nk->ni", lin.Jp, lin.r), prob.num_points)
    gc = lm.segment_sum(oi, np.einsum("nki,nk->ni", lin.Jc, lin.r), prob.num_images)
    h, tol = 1e-4, 2e-4  # the cost is ~1e6 with the trivial loss: 1e-16 * 1e6 / h of rounding
    for p in prob.obs_pt[prob.plane_obs[:4]]:
        for a in range(3):
            c = []
            for sgn in (1, -1):
                X = st.X.copy()
                X[p, a] += sgn * h
                c.append(evaluate(prob, State(st.q, st.t, X), jacobians=False).cost)
            assert abs((c[0] - c[1]) / (2 * h) - gp[p, a]) < tol * max(1.0, abs(gp[p, a]))
    img = int(prob.obs_img[prob.plane_obs[-1]])
    assert prob.img_free_mask[img] == 0b111111
    for a in range(6):
        c = []
        for sgn in (1, -1):
            d = np.zeros(6)
            d[a] = sgn * h
            q, t = st.q.copy(), st.t.copy()
            q[img] = geometry.quat_plus(st.q[img][None], d[None, :3])[0]
            t[img] = st.t[img] + d[3:]
            c.append(evaluate(prob, State(q, t, st.X), jacobians=False).cost)
        assert abs((c[0] - c[1]) / (2 * h) - gc[img, a]) < tol * max(1.0, abs(gc[img, a]))


def dataclass_without_planes(prob):
    import dataclasses

    return dataclasses.replace(prob, plane_obs=None, plane_m=None, plane_c=None)


@pytest.mark.parametrize("model", [CameraModel.PINHOLE, CameraModel.SIMPLE_RADIAL, CameraModel.OPENCV])
@pytest.mark.parametrize("depth_loss", [LossType.CAUCHY, LossType.SOFT_L1, LossType.TRIVIAL])
def test_product_arithmetic_matches_the_definition(hostcheck, model, depth_loss):
    prob = _scene(model=model, depth_loss=depth_loss)
    prob.img_free_mask[:] = 0b111111
    st = State.of(prob)
    lin = evaluate(prob, st, jacobians=True)
    q = ke.obs_quantities(hostcheck, prob, st.q, st.t, st.X)
    assert rel(q["rn"][prob.plane_obs], lin.r_plane) < 1e-9
    others = np.setdiff1d(np.arange(prob.num_obs), prob.plane_obs)
    assert np.all(q["rn"][others] == 0.0)
    assert abs(0.5 * q["rho"].sum() - (lin.cost + lin.fixed_cost)) <= 1e-12 * lin.cost
    assert rel(q["Hcc"], np.einsum("nki,nkj->nij", lin.Jc, lin.Jc)) < 1e-9
    assert rel(q["gc"], np.einsum("nki,nk->ni", lin.Jc, lin.r)) < 1e-9
    R = geometry.rotmat(st.q[prob.obs_img])
    assert rel(np.einsum("nki,nkl,nlj->nij", R, q["N"], R), np.einsum("nki,nkj->nij", lin.Jp, lin.Jp)) < 1e-9
    assert rel(np.einsum("nki,nk->ni", R, q["h"]), np.einsum("nki,nk->ni", lin.Jp, lin.r)) < 1e-9
    G = ke.g_matrix(q["prot"])
    assert rel(np.einsum("nki,nkl,nlj->nij", G, q["N"], R), np.einsum("nki,nkj->nij", lin.Jc, lin.Jp)) < 1e-9
    # without Jacobians: same cost
    q0 = ke.obs_quantities(hostcheck, prob, st.q, st.t, st.X, want_jac=False)
    assert np.array_equal(q0["rho"], q["rho"])


def test_layout_carries_the_plane_records_in_both_orders():
    from mpsfm_b200 import engine

    prob = _scene()
    for lay in (engine.build_layout(prob), engine.build_layout_numpy(prob)):
        full = np.zeros((prob.num_obs, 4))
        full[prob.plane_obs, :3], full[prob.plane_obs, 3] = prob.plane_m, prob.plane_c
        assert np.array_equal(lay.pm_plane, full[lay.pm_src]) and np.array_equal(lay.cm_plane, full[lay.cm_src])
        assert lay.plane_loss_scale == prob.plane_loss_scale
    plain = engine.build_layout(dataclass_without_planes(prob))
    assert plain.pm_plane is None and plain.cm_plane is None
    # shards carry their own slots
    sh, spec = engine.shard_layout(engine.build_layout(prob), 1, 2)
    lay = engine.build_layout(prob)
    s0, s1 = int(lay.pt_ptr[spec.pt_begin]), int(lay.pt_ptr[spec.pt_end])
    c0, c1 = int(lay.img_ptr[spec.img_begin]), int(lay.img_ptr[spec.img_end])
    assert np.array_equal(sh.pm_plane[: s1 - s0], lay.pm_plane[s0:s1])
    assert np.array_equal(sh.cm_plane[: c1 - c0], lay.cm_plane[c0:c1])


def test_cpu_port_with_plane_rows_follows_the_oracle_trajectory():
    """The CPU port runs the product's linearize_obs: the whole LM trajectory with the extra rows against the oracle's."""
    from mpsfm_b200 import engine
    from oracle import cpu_port

    prob = normal_prior.synthetic_normal_priors(make_scene(14, 700, 5, seed=9, model=CameraModel.SIMPLE_RADIAL,
                                                           depth_frac=0.7), frac=0.6, seed=2)
    osum = lm.solve(prob)
    base = lm.solve(dataclass_without_planes(prob))
    summ, q, t, X = cpu_port.solve(engine.build_layout(prob))
    assert summ["num_iterations"] == len(osum.iterations) - 1
    oc = np.array([it.cost for it in osum.iterations])
    assert np.max(np.abs(np.array(summ["costs"]) - oc) / oc) <= 1e-6
    assert np.max(np.abs(t - osum.state.t)) < 1e-6 and np.max(np.abs(X - osum.state.X)) < 1e-6
    # the rows do something: another optimum than without them
    assert np.max(np.abs(osum.state.X - base.state.X)) > 1e-4


def test_optimizer_option_adds_rows_for_images_with_normals():
    """``gpu.normal_prior.enabled`` (off by default): the assembled problem gains one row per depth-prior observation of
    the images that carry a normal map; everything else of the assembly is unchanged."""
    from mpsfm_b200.optimizer import Optimizer
    from scene_factory import attach_normals, build_scene

    rec = attach_normals(build_scene(seed=11, num_images=8, num_points=260))
    rec.images[3].normals = None  # one image without a normal map
    bundle = {"optim_ids": [2, 1, 3, 4, 5, 6, 7, 8], "pts3D": set(rec.points3D)}
    off = Optimizer({}, rec, None)
    off.update_truncation_multiplier(rec.reg_image_ids())
    on = Optimizer({"gpu": {"normal_prior": {"enabled": True, "rel_sigma": 0.05}}}, rec, None)
    on.truncation_multiplier = off.truncation_multiplier
    a0 = off.assemble(bundle, fix_pose=False, mode="global", allow_scale_filter=True).problem
    a1 = on.assemble(bundle, fix_pose=False, mode="global", allow_scale_filter=True).problem
    assert a0.num_plane == 0 and a1.num_plane > 0
    for f in ("obs_img", "obs_pt", "obs_xy", "depth_obs", "depth_d", "depth_w", "depth_a", "img_q", "img_t", "pt_xyz"):
        assert np.array_equal(getattr(a0, f), getattr(a1, f)), f
    row3 = [2, 1, 3, 4, 5, 6, 7, 8].index(3)
    with_map = a1.depth_obs[a1.obs_img[a1.depth_obs] != row3]
    assert np.array_equal(np.sort(a1.plane_obs), np.sort(with_map))
    k = np.searchsorted(a1.depth_obs, a1.plane_obs)
    assert np.allclose(np.linalg.norm(a1.plane_m, axis=1), 1.0 / (0.05 * a1.depth_d[k]))
    assert np.all(a1.plane_m[:, 2] < 0)  # facing the camera
    # the oracle solves it, and the planes pull: another optimum, points closer to the prior surface
    s0, s1 = lm.solve(a0), lm.solve(a1)
    assert s1.termination == "CONVERGENCE"

    def plane_rms(sol):
        pc = geometry.quat_rotate(sol.state.q[a1.obs_img], sol.state.X[a1.obs_pt]) + sol.state.t[a1.obs_img]
        return float(np.sqrt(np.mean((np.einsum("ni,ni->n", a1.plane_m, pc[a1.plane_obs]) - a1.plane_c) ** 2)))

    assert plane_rms(s1) < plane_rms(s0)
