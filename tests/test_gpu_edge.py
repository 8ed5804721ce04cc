"""Edge cases of the C ABI on the GPU: degenerate problems must solve (or fail loudly), never crash or hang."""

import dataclasses

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene
from oracle import geometry, lm
from oracle.residuals import State, evaluate

pytestmark = pytest.mark.gpu


def _subset(prob, keep):
    new_index = np.cumsum(keep) - 1
    dkeep = keep[prob.depth_obs]
    return dataclasses.replace(
        prob, obs_img=prob.obs_img[keep], obs_pt=prob.obs_pt[keep], obs_xy=prob.obs_xy[keep],
        depth_obs=new_index[prob.depth_obs[dkeep]].astype(np.int32), depth_d=prob.depth_d[dkeep],
        depth_w=prob.depth_w[dkeep], depth_a=prob.depth_a[dkeep])


def test_all_points_constant_only_poses_move():
    prob = make_scene(8, 300, 4, seed=1, point_noise=0.0)
    prob.pt_const[:] = 1
    ba = engine.BundleAdjuster(prob, device=0)
    summ = ba.solve()
    osum = lm.solve(prob)
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    assert np.array_equal(ba.state()[2], prob.pt_xyz)


def test_image_without_observations_and_unobserved_points():
    prob = make_scene(9, 200, 3, seed=4)
    keep = (prob.obs_img != 5) & (prob.obs_pt % 7 != 0)  # image 5 sees nothing, every 7th point is unobserved
    prob = _subset(prob, keep)
    ba = engine.BundleAdjuster(prob, device=0)
    summ = ba.solve()
    osum = lm.solve(prob)
    assert summ.termination == osum.termination
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    q, t, X = ba.state()
    assert np.array_equal(X[::7], prob.pt_xyz[::7])
    # the unobserved image only follows the gauge normalisation, as in the oracle
    assert np.max(np.abs(t[5] - osum.state.t[5])) < 1e-7


def test_no_observations_at_all():
    prob = make_scene(4, 30, 3, seed=0)
    prob = _subset(prob, np.zeros(prob.num_obs, dtype=bool))
    ba = engine.BundleAdjuster(prob, device=0)
    summ = ba.solve()
    assert summ.final_cost == 0.0 and summ.num_iterations <= 1
    for a, b in zip(ba.state()[1:], (prob.img_t, prob.pt_xyz)):
        assert np.array_equal(a, b)


def _behind_scene(depth_prior):
    prob = make_scene(8, 200, 4, seed=2, depth_prior=depth_prior)
    # mirror three points through the ring: behind every camera that observes them
    # (constant points: a free point behind a camera drifts towards z = 0, where 1 / z makes any comparison fragile)
    for p in (0, 17, 90):
        prob.pt_xyz[p] = prob.pt_xyz[p] * -12.0
        prob.pt_const[p] = 1
    return prob


def test_point_behind_the_camera_is_a_zero_block_like_colmap():
    """Default policy (mpba_options.behind_camera = 0): COLMAP >= 3.11's cost functor writes a zero residual when
    ImgFromCam fails (z < eps); the solve carries on -- the reference never fails here (bundle_adjustment.py:266,
    :283).  Residuals, normal equations and the whole trajectory against the oracle, which states the same rule."""
    prob = _behind_scene(depth_prior=False)
    ba = engine.BundleAdjuster(prob, device=0)
    rr, rd, cost, fixed = ba.evaluate()
    ref = evaluate(prob, State.of(prob), jacobians=False)
    pz = (geometry.quat_rotate(State.of(prob).q[prob.obs_img], prob.pt_xyz[prob.obs_pt]) + prob.img_t[prob.obs_img])[:, 2]
    behind = np.flatnonzero(pz < np.finfo(float).eps)
    assert behind.size >= 3 and np.all(rr[behind] == 0.0) and np.all(ref.r_reproj[behind] == 0.0)
    assert np.max(np.abs(rr - ref.r_reproj) / (np.abs(ref.r_reproj) + 1e-6)) < 1e-9
    assert abs(cost - ref.cost) <= 1e-12 * ref.cost
    summ = ba.solve()
    osum = lm.solve(prob)
    assert summ.termination == osum.termination and summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    q, t, X = ba.state()
    assert np.max(np.abs(t - osum.state.t)) < 1e-6
    pz1 = (geometry.quat_rotate(osum.state.q[prob.obs_img], osum.state.X[prob.obs_pt]) + osum.state.t[prob.obs_img])[:, 2]
    assert summ.num_behind_camera == int(np.sum(pz1 < np.finfo(float).eps)) > 0


def test_non_finite_initial_cost_is_a_failed_solve_not_an_exception():
    """A depth prior on a point with z <= 0 gives log(negative) = nan: Ceres' initial evaluation fails, pyceres.solve
    returns FAILURE without raising and the parameters stay put (bundle_adjustment.py:285-293 ignores the summary).
    Same here: MPBA_OK, termination FAILURE, state bitwise untouched."""
    prob = _behind_scene(depth_prior=True)
    ba = engine.BundleAdjuster(prob, device=0)
    before = [a.copy() for a in ba.state()]
    summ = ba.solve()
    assert summ.termination == "FAILURE" and summ.num_iterations == 0
    assert not np.isfinite(summ.initial_cost)
    for a, b in zip(ba.state(), before):
        assert np.array_equal(a, b)
    # the handle stays usable
    prob2 = make_scene(6, 60, 3, seed=2)
    ba2 = engine.BundleAdjuster(prob2, device=0)
    assert ba2.solve().termination == "CONVERGENCE"


def test_point_behind_the_camera_strict_mode_is_loud():
    prob = _behind_scene(depth_prior=True)
    ba = engine.BundleAdjuster(prob, device=0)
    with pytest.raises(engine.EngineError):
        ba.solve(engine.default_options(behind_camera=1))


def test_two_images_minimal_gauge():
    prob = make_scene(2, 40, 2, seed=3, model=CameraModel.PINHOLE, reproj_loss=LossType.TRIVIAL)
    ba = engine.BundleAdjuster(prob, device=0)
    summ = ba.solve()
    osum = lm.solve(prob)
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * max(osum.final_cost, 1e-12)


def test_upload_derives_the_image_order_arrays_on_the_device():
    """pinned staging: upload() sends the track order + a permutation and mpba_gather_observations rebuilds
    cm_xy / cm_depth; the result must be the host layout's arrays bit for bit, and the solve unchanged."""
    prob = make_scene(12, 900, 5, seed=8, depth_prior=True)
    plain = engine.BundleAdjuster(prob, device=0)
    ref = plain.solve()
    ba = engine.BundleAdjuster(prob, device=0, pinned=True)
    assert "cm_xy" not in ba.host and "cm_pm" in ba.host
    ba.buf["cm_xy"].fill_(float("nan"))
    ba.buf["cm_depth"].fill_(float("nan"))
    nbytes = ba.upload()
    assert nbytes < plain.h2d_bytes
    assert np.array_equal(ba.buf["cm_xy"].cpu().numpy(), ba.layout.cm_xy)
    assert np.array_equal(ba.buf["cm_depth"].cpu().numpy(), ba.layout.cm_depth)
    summ = ba.solve()
    assert summ.num_iterations == ref.num_iterations and summ.final_cost == ref.final_cost
    for a, b in zip(ba.state(), plain.state()):
        assert np.array_equal(a, b)


def test_large_coarse_space_cluster_inverse_matches_block_jacobi_and_oracle():
    """>= 600 images: 32 coarse groups (224 modes), the coarse matrix is inverted by the four-CTA cluster kernel.
    Both PCG kernels with that preconditioner must land on the block-Jacobi-only solve and on the exact-solve
    oracle (same LM iteration count, cost <= 1e-6, poses <= 1e-6)."""
    prob = make_scene(640, 6000, 4, seed=3)
    osum = lm.solve(prob)
    results = {}
    for name, kw in (("cluster", dict(pcg_kernel=0, pcg_coarse=0)), ("grid", dict(pcg_kernel=1, pcg_coarse=0)),
                     ("jacobi", dict(pcg_kernel=0, pcg_coarse=1))):
        ba = engine.BundleAdjuster(prob, device=0)
        summ = ba.solve(engine.default_options(pcg_max_iterations=6000, **kw))
        results[name] = (summ, ba.state())
        ba.close()
        assert summ.num_iterations == len(osum.iterations) - 1, name
        assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost, name
        assert np.max(np.abs(results[name][1][1] - osum.state.t)) < 1e-6, name
    # the coarse space is what it is for: far fewer PCG iterations than block-Jacobi alone
    assert results["cluster"][0].total_pcg_iterations < 0.5 * results["jacobi"][0].total_pcg_iterations
    assert results["grid"][0].total_pcg_iterations < 0.5 * results["jacobi"][0].total_pcg_iterations
