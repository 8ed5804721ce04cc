"""Edge cases of the C ABI on the GPU: degenerate problems must solve (or fail loudly), never crash or hang."""

import dataclasses

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene
from oracle import lm

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


def test_point_behind_the_camera_fails_loudly():
    prob = make_scene(6, 60, 3, seed=2, depth_prior=True)
    prob.pt_xyz[0] = prob.pt_xyz[0] * -50.0  # far behind its cameras: log of a negative depth ratio
    ba = engine.BundleAdjuster(prob, device=0)
    with pytest.raises(engine.EngineError):
        ba.solve()


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
