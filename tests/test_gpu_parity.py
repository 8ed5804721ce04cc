"""GPU parity: every stage of the CUDA path, through the C ABI, against the CPU oracle.

Tolerances are BASELINE.json's: per-residual <= 1e-9 relative (fp64); final cost <= 1e-6 relative after
the same LM iteration count; poses <= 1e-6 rad / 1e-6 scene units; indexing bit-exact."""

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene
from oracle import geometry, lm
from oracle.residuals import State, evaluate, free_columns

pytestmark = pytest.mark.gpu

SCENES = {
    "pinhole_softl1_cauchy": dict(model=CameraModel.PINHOLE),
    "simple_radial_cauchy": dict(model=CameraModel.SIMPLE_RADIAL, reproj_loss=LossType.CAUCHY),
    "opencv_huber": dict(model=CameraModel.OPENCV, reproj_loss=LossType.HUBER, depth_loss=LossType.HUBER),
    "radial_trivial_partial_depth": dict(model=CameraModel.RADIAL, reproj_loss=LossType.TRIVIAL,
                                         depth_loss=LossType.SOFT_L1, depth_frac=0.5, const_point_frac=0.1),
    "simple_pinhole_nodepth": dict(model=CameraModel.SIMPLE_PINHOLE, depth_prior=False),
    # the six models whose Jacobians come from forward duals on the device: nvcc's FMA
    # contraction and the device atan / tan are only visible on the CUDA path
    "opencv_fisheye_softl1": dict(model=CameraModel.OPENCV_FISHEYE),
    "full_opencv_cauchy": dict(model=CameraModel.FULL_OPENCV, reproj_loss=LossType.CAUCHY),
    "fov_huber": dict(model=CameraModel.FOV, reproj_loss=LossType.HUBER),
    "simple_radial_fisheye_softl1": dict(model=CameraModel.SIMPLE_RADIAL_FISHEYE, depth_frac=0.6),
    "radial_fisheye_trivial": dict(model=CameraModel.RADIAL_FISHEYE, reproj_loss=LossType.TRIVIAL),
    "thin_prism_fisheye_cauchy": dict(model=CameraModel.THIN_PRISM_FISHEYE, reproj_loss=LossType.CAUCHY,
                                      const_point_frac=0.05),
}


def relerr(a, b):
    scale = np.max(np.abs(b)) if b.size else 1.0
    return float(np.max(np.abs(a - b)) / max(scale, 1e-300)) if b.size else 0.0


@pytest.fixture(scope="module", params=sorted(SCENES))
def scene(request):
    kw = SCENES[request.param]
    prob = make_scene(14, 900, 5, seed=11, **kw)
    return prob, engine.BundleAdjuster(prob, device=0)


def test_residuals_and_cost(scene):
    prob, ba = scene
    ba.reset_state()
    rr, rd, cost, fixed = ba.evaluate()
    ref = evaluate(prob, State.of(prob), jacobians=False)
    assert np.max(np.abs(rr - ref.r_reproj) / (np.abs(ref.r_reproj) + 1e-6)) < 1e-9
    has = ~np.isnan(rd)
    assert np.array_equal(np.flatnonzero(has), np.sort(prob.depth_obs))  # indexing bit-exact
    assert np.max(np.abs(rd[prob.depth_obs] - ref.r_depth) / (np.abs(ref.r_depth) + 1e-9), initial=0.0) < 1e-9
    assert abs(cost - ref.cost) <= 1e-12 * max(ref.cost, 1.0)
    assert abs(fixed - ref.fixed_cost) <= 1e-12 * max(ref.fixed_cost, 1.0)


def test_normal_equation_blocks(scene):
    prob, ba = scene
    ba.reset_state()
    Hcc, gc, Hpp, gp, cost, fixed = ba.linearize()
    lin = evaluate(prob, State.of(prob), jacobians=True)
    freec = free_columns(prob).astype(float)
    freep = np.repeat((prob.pt_const == 0)[:, None], 3, axis=1).astype(float)
    ne = lm.NormalEquations(prob, lin, freec, freep)
    assert relerr(Hcc, ne.Hcc) < 1e-11
    assert relerr(gc, ne.gc) < 1e-11
    assert relerr(Hpp, ne.Hpp) < 1e-11
    assert relerr(gp, ne.gp) < 1e-11
    assert abs(cost - lin.cost) <= 1e-12 * lin.cost


def _oracle_first_system(prob, radius):
    st = State.of(prob)
    lin = evaluate(prob, st, jacobians=True)
    freec = free_columns(prob)
    freep = np.repeat((prob.pt_const == 0)[:, None], 3, axis=1)
    ne0 = lm.NormalEquations(prob, lin, np.ones((prob.num_images, 6)), np.ones((prob.num_points, 3)))
    n2c, n2p = ne0.column_sq_norms()
    c_img = np.where(freec, 1.0 / (1.0 + np.sqrt(n2c)), 0.0)
    c_pt = np.where(freep, 1.0 / (1.0 + np.sqrt(n2p)), 0.0)
    ne = lm.NormalEquations(prob, lin, c_img, c_pt)
    dc, dp = ne.column_sq_norms()
    dc, dp = np.clip(dc, 1e-6, 1e32), np.clip(dp, 1e-6, 1e32)
    return ne, dc / radius, dp / radius, c_img, c_pt


def test_reduced_camera_system_and_pcg(scene):
    prob, ba = scene
    ba.reset_state()
    radius = 1e4
    S_blocks, rhs, y, iters = ba.reduced_system(radius, engine.default_options(pcg_tolerance=1e-12))
    ne, Dc2, Dp2, c_img, c_pt = _oracle_first_system(prob, radius)
    S_o, rhs_o, _ = ne.schur_system(Dc2, Dp2)
    # oracle works in Jacobi-scaled variables: S_unscaled = diag(1/c) S_scaled diag(1/c) on free coordinates
    free = np.flatnonzero(free_columns(prob).reshape(-1))
    cs = c_img.reshape(-1)[free]
    S_oo = S_o.toarray()[np.ix_(free, free)] / np.outer(cs, cs)
    S = engine.dense_reduced_system(ba.layout, S_blocks)
    assert relerr(S[np.ix_(free, free)], S_oo) < 1e-9
    fixed = np.setdiff1d(np.arange(S.shape[0]), free)
    assert np.array_equal(S[np.ix_(fixed, fixed)], np.eye(fixed.size))
    assert relerr(rhs.reshape(-1)[free], rhs_o[free] / cs) < 1e-9
    yc_o, _ = ne.solve(Dc2, Dp2)
    y_o = (yc_o * c_img).reshape(-1)
    assert np.max(np.abs(y.reshape(-1) - y_o)) <= 1e-7 * np.max(np.abs(y_o))
    assert iters == 0  # 14 images: the dense direct solve (dense_solve.cu), no conjugate gradients
    # the iterative path on the same system (what larger bundles run)
    S2, rhs2, y2, iters2 = ba.reduced_system(radius, engine.default_options(pcg_kernel=2, pcg_tolerance=1e-12))
    assert np.array_equal(S2, S_blocks) and np.array_equal(rhs2, rhs)
    assert np.max(np.abs(y2.reshape(-1) - y_o)) <= 1e-7 * np.max(np.abs(y_o))
    assert 0 < iters2 < 2000


def _angle(qa, qb):
    d = np.abs(np.sum(qa * qb, axis=1)).clip(max=1.0)
    return 2.0 * np.arccos(d)


def test_full_solve_matches_oracle(scene):
    prob, ba = scene
    ba.reset_state()
    summ = ba.solve()
    osum = lm.solve(prob)
    assert summ.termination == osum.termination, (summ.message, osum.message)
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.initial_cost - osum.initial_cost) <= 1e-12 * osum.initial_cost
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    for a, b in zip(summ.iterations, osum.iterations):
        assert abs(a["cost"] - b.cost) <= 1e-6 * b.cost
        assert bool(a["step_is_successful"]) == bool(b.step_is_successful)
    q, t, X = ba.state()
    assert np.max(_angle(q, osum.state.q)) < 1e-6
    assert np.max(np.abs(t - osum.state.t)) < 1e-6
    assert np.max(np.abs(X - osum.state.X)) < 1e-6
    # constant blocks are untouched (bundle_adjustment.py:114-121)
    q0 = State.of(prob)
    assert np.array_equal(q[0], q0.q[0]) and np.array_equal(t[0], q0.t[0])
    assert t[1, 0] == q0.t[1, 0]
    const = prob.pt_const != 0
    assert np.array_equal(X[const], prob.pt_xyz[const])


@pytest.mark.parametrize("case", ["fov_small_omega", "fov_small_radius", "fisheye_on_axis"])
def test_camera_model_branches_on_device(case):
    """The series branches of FOV (omega^2 < 1e-4; r^2 < 1e-4) and the r <= eps branch of the fisheye family, on the
    CUDA path: a scene whose points sit (nearly) on the optical axes, residuals and normal-equation blocks."""
    if case == "fov_small_omega":
        prob = make_scene(8, 300, 4, model=CameraModel.FOV, seed=4)
        prob.cam_params[0, 4] = 5e-3
    elif case == "fov_small_radius":
        prob = make_scene(8, 300, 4, model=CameraModel.FOV, seed=4)
    else:
        prob = make_scene(8, 300, 4, model=CameraModel.RADIAL_FISHEYE, seed=4)
    if case != "fov_small_omega":
        # move a third of the points onto the optical axis of their first observing image (u, v ~ 0 or exactly 0)
        R = geometry.rotmat(prob.img_q)
        first = {}
        for o in range(prob.num_obs):
            first.setdefault(int(prob.obs_pt[o]), int(prob.obs_img[o]))
        for p_, i in list(first.items())[::3]:
            centre = -R[i].T @ prob.img_t[i]
            depth = 35.0 + (p_ % 7)
            off = np.array([1e-3, 2e-3, 0.0]) * (p_ % 2) if case == "fov_small_radius" else np.zeros(3)
            prob.pt_xyz[p_] = centre + R[i].T @ (np.array([0.0, 0.0, depth]) + off * depth)
    ba = engine.BundleAdjuster(prob, device=0)
    rr, rd, cost, fixed = ba.evaluate()
    ref = evaluate(prob, State.of(prob), jacobians=False)
    assert np.max(np.abs(rr - ref.r_reproj) / (np.abs(ref.r_reproj) + 1e-6)) < 1e-9
    assert abs(cost - ref.cost) <= 1e-12 * max(ref.cost, 1.0)
    Hcc, gc, Hpp, gp, _, _ = ba.linearize()
    lin = evaluate(prob, State.of(prob), jacobians=True)
    freec = free_columns(prob).astype(float)
    freep = np.repeat((prob.pt_const == 0)[:, None], 3, axis=1).astype(float)
    ne = lm.NormalEquations(prob, lin, freec, freep)
    assert relerr(Hcc, ne.Hcc) < 1e-11 and relerr(gc, ne.gc) < 1e-11
    assert relerr(Hpp, ne.Hpp) < 1e-11 and relerr(gp, ne.gp) < 1e-11


def test_structure_only_refine_3d_points():
    """refine_3d_points: all poses constant, trivial depth loss (bundle_adjustment.py:276-283)."""
    prob = make_scene(10, 500, 5, seed=2, depth_loss=LossType.TRIVIAL)
    prob.img_free_mask[:] = 0
    ba = engine.BundleAdjuster(prob, device=0)
    summ = ba.solve()
    osum = lm.solve(prob)
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    q, t, X = ba.state()
    assert np.array_equal(t, State.of(prob).t)
    assert np.max(np.abs(X - osum.state.X)) < 1e-6


def test_size_independent_properties_at_scale():
    """cfg2-sized scene (100 images / 50k points / 300k observations): properties that need no oracle run."""
    from mpsfm_b200.synthetic import make_config

    prob = make_config("cfg2")
    ba = engine.BundleAdjuster(prob, device=0)
    rr, rd, cost0, _ = ba.evaluate()
    # the cost is the checksum of the residuals
    from oracle import losses

    rho_r = losses.scaled_rho(int(prob.reproj_loss), prob.reproj_loss_scale, prob.reproj_loss_weight, (rr * rr).sum(1))[0]
    rho_d = losses.scaled_rho(int(prob.depth_loss), prob.depth_a, prob.depth_w, rd[prob.depth_obs] ** 2)[0]
    assert abs(0.5 * (rho_r.sum() + rho_d.sum()) - cost0) <= 1e-10 * cost0
    summ = ba.solve()
    costs = [it["cost"] for it in summ.iterations]
    assert all(b <= a * (1 + 1e-15) for a, b in zip(costs, costs[1:]))  # monotone
    assert summ.final_cost < 0.5 * summ.initial_cost
    # idempotence: solving again from the solution terminates at once with the same cost
    again = ba.solve()
    assert again.num_iterations <= 2
    assert abs(again.final_cost - summ.final_cost) <= 1e-5 * summ.final_cost
    # determinism of the reductions that have a fixed order
    ba.reset_state()
    a = ba.evaluate()[2]
    b = ba.evaluate()[2]
    assert a == b


def test_no_cpu_fallback_errors_are_loud():
    prob = make_scene(6, 50, 3, seed=0)
    ba = engine.BundleAdjuster(prob, device=0)
    ws = ba.workspace
    ba.workspace = ws[:16]
    with pytest.raises(engine.EngineError):
        ba.solve()
    ba.workspace = ws


@pytest.mark.parametrize("shape", [(2, 60, 2), (6, 80, 4), (14, 900, 5), (21, 1500, 6)])
@pytest.mark.parametrize("radius", [1e4, 1e12])
def test_dense_direct_solve_of_small_systems(shape, radius):
    """Up to 21 images the reduced camera system is solved directly by one CTA (dense_solve.cu; the default,
    pcg_kernel = 0): against numpy's dense solve of the same S, and against the cluster PCG."""
    prob = make_scene(*shape, seed=7)
    ba = engine.BundleAdjuster(prob, device=0)
    S0, rhs0, y0, it0 = ba.reduced_system(radius, engine.default_options())
    S2, rhs2, y2, it2 = ba.reduced_system(radius, engine.default_options(pcg_kernel=2, pcg_tolerance=1e-13))
    assert it0 == 0 and it2 > 0 and np.array_equal(S0, S2) and np.array_equal(rhs0, rhs2)
    S = engine.dense_reduced_system(ba.layout, S0)
    y_ref = np.linalg.solve(S, rhs0.reshape(-1))
    scale = np.max(np.abs(y_ref))
    assert np.max(np.abs(y0.reshape(-1) - y_ref)) <= 1e-9 * scale
    # no worse than the iterative solve it replaces
    assert np.max(np.abs(y0.reshape(-1) - y_ref)) <= max(np.max(np.abs(y2.reshape(-1) - y_ref)), 1e-12 * scale)
    _, _, y0b, _ = ba.reduced_system(radius, engine.default_options())
    assert np.array_equal(y0, y0b)


@pytest.mark.parametrize("shape", [(6, 80, 4), (40, 3000, 5), (260, 12000, 5)])
def test_cluster_and_grid_pcg_agree(shape):
    """The cluster-resident PCG (pcg_cluster.cu) and the grid-wide PCG solve the same system to the same
    answer; with the same (block-Jacobi) preconditioner the iteration counts agree (+-2: the single-reduction
    recurrence rounds differently); the two-level preconditioner (coarse.cu) needs fewer iterations."""
    prob = make_scene(*shape, seed=5)
    ba = engine.BundleAdjuster(prob, device=0)
    S1, rhs1, y1, it1 = ba.reduced_system(1e4, engine.default_options(pcg_kernel=1, pcg_coarse=1, pcg_tolerance=1e-12))
    S3, rhs3, y3, it3 = ba.reduced_system(1e4, engine.default_options(pcg_kernel=1, pcg_tolerance=1e-12))  # grid, two-level
    S0, rhs0, y0, it0 = ba.reduced_system(1e4, engine.default_options(pcg_kernel=2, pcg_coarse=1, pcg_tolerance=1e-12))
    S2, rhs2, y2, it2 = ba.reduced_system(1e4, engine.default_options(pcg_kernel=2, pcg_tolerance=1e-12))
    assert np.array_equal(S0, S1) and np.array_equal(rhs0, rhs1) and np.array_equal(S0, S2)
    assert abs(it0 - it1) <= 2, (it0, it1)
    assert np.max(np.abs(y0 - y1)) <= 1e-9 * np.max(np.abs(y1))
    # exact check against a dense solve
    S = engine.dense_reduced_system(ba.layout, S0)
    y_ref = np.linalg.solve(S, rhs0.reshape(-1))
    assert np.max(np.abs(y0.reshape(-1) - y_ref)) <= 1e-8 * np.max(np.abs(y_ref))
    assert np.max(np.abs(y2.reshape(-1) - y_ref)) <= 1e-8 * np.max(np.abs(y_ref))
    assert it2 <= it0
    assert np.max(np.abs(y3.reshape(-1) - y_ref)) <= 1e-8 * np.max(np.abs(y_ref))
    assert it3 <= it1
    if shape[0] >= 40:
        assert it2 < 0.6 * it0, (it2, it0)
        assert it3 < 0.7 * it1, (it3, it1)
    # deterministic: bitwise identical on a second run
    _, _, y2b, it2b = ba.reduced_system(1e4, engine.default_options(pcg_kernel=2, pcg_tolerance=1e-12))
    assert it2b == it2 and np.array_equal(y2, y2b)
