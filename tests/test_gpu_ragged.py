"""Ragged tracks: empty tracks, single observations, tracks longer than one warp (33+) -- the rarely taken
paths of the track-order kernels (warp path of k_points, round loop of k_back_substitute) against the oracle."""

import dataclasses

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene
from oracle import lm
from oracle.residuals import State, evaluate

pytestmark = pytest.mark.gpu


def ragged_scene(seed=0):
    """48 images; track lengths 0, 1, 2, 5, 31, 32, 33, 40 (cycled over the points), 8 % constant points."""
    prob = make_scene(48, 120, 40, seed=seed, model=CameraModel.SIMPLE_RADIAL, reproj_loss=LossType.CAUCHY,
                      const_point_frac=0.08, outlier_frac=0.02)
    want = np.array([0, 1, 2, 5, 31, 32, 33, 40])[np.arange(prob.num_points) % 8]
    # rank of every observation inside its track (natural order is image-major)
    order = np.argsort(prob.obs_pt, kind="stable")
    rank = np.empty(prob.num_obs, dtype=np.int64)
    start = np.searchsorted(prob.obs_pt[order], np.arange(prob.num_points))
    rank[order] = np.arange(prob.num_obs) - start[prob.obs_pt[order]]
    keep = rank < want[prob.obs_pt]
    new_index = np.cumsum(keep) - 1
    dkeep = keep[prob.depth_obs]
    return dataclasses.replace(
        prob, obs_img=prob.obs_img[keep], obs_pt=prob.obs_pt[keep], obs_xy=prob.obs_xy[keep],
        depth_obs=new_index[prob.depth_obs[dkeep]].astype(np.int32), depth_d=prob.depth_d[dkeep],
        depth_w=prob.depth_w[dkeep], depth_a=prob.depth_a[dkeep])


@pytest.fixture(scope="module")
def ragged():
    prob = ragged_scene(3)
    lay = engine.build_layout(prob)
    lengths = np.diff(lay.pt_ptr)
    assert lengths.min() == 0 and lengths.max() == 40 and np.any(lengths == 33)
    groups = np.diff(lay.track_warp_first)
    assert groups.min() >= 1 and np.all(np.add.reduceat(lengths, lay.track_warp_first[:-1])[groups > 1] <= 32)
    return prob, lay, engine.BundleAdjuster(prob, device=0, layout=lay)


def test_blocks_match_the_oracle(ragged):
    prob, lay, ba = ragged
    ba.reset_state()
    rr, rd, cost, fixed = ba.evaluate()
    ref = evaluate(prob, State.of(prob), jacobians=False)
    assert np.max(np.abs(rr - ref.r_reproj) / (np.abs(ref.r_reproj) + 1e-6)) < 1e-9
    assert abs(cost - ref.cost) <= 1e-12 * ref.cost
    Hcc, gc, Hpp, gp, _, _ = ba.linearize()
    lin = evaluate(prob, State.of(prob), jacobians=True)
    ne = lm.NormalEquations(prob, lin, np.ones((prob.num_images, 6)), np.ones((prob.num_points, 3)))
    scale = lambda a: max(float(np.max(np.abs(a))), 1e-300)  # noqa: E731
    assert np.max(np.abs(Hpp - ne.Hpp)) / scale(ne.Hpp) < 1e-11
    assert np.max(np.abs(gp - ne.gp)) / scale(ne.gp) < 1e-11
    assert np.max(np.abs(Hcc - ne.Hcc)) / scale(ne.Hcc) < 1e-11


def test_full_solve_matches_the_oracle(ragged):
    prob, lay, ba = ragged
    ba.reset_state()
    summ = ba.solve()
    osum = lm.solve(prob)
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    for a, b in zip(summ.iterations, osum.iterations):
        assert abs(a["cost"] - b.cost) <= 1e-6 * b.cost
        assert bool(a["step_is_successful"]) == bool(b.step_is_successful)
    q, t, X = ba.state()
    assert np.max(np.abs(t - osum.state.t)) < 1e-6 and np.max(np.abs(X - osum.state.X)) < 1e-6
    # points without observations and constant points keep their coordinates bit for bit
    untouched = (np.diff(lay.pt_ptr) == 0) | (prob.pt_const != 0)
    assert np.array_equal(X[untouched], prob.pt_xyz[untouched])


def test_sharded_code_path_and_covariances(ragged):
    prob, lay, ba = ragged
    ba.reset_state()
    s1 = ba.solve()
    sh = engine.BundleAdjuster(prob, device=0, layout=lay, shard=(0, 1))
    s2 = sh.solve()
    assert s1.num_iterations == s2.num_iterations and abs(s1.final_cost - s2.final_cost) <= 1e-12 * s1.final_cost
    for x, y in zip(ba.state(), sh.state()):
        assert np.max(np.abs(x - y)) <= 1e-9
    cov = ba.point_covariances()
    active = (np.diff(lay.pt_ptr) > 0) & (prob.pt_const == 0)
    assert np.all(cov[~active] == 0) and np.all(np.linalg.eigvalsh(cov[active]) > 0)
