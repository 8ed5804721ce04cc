"""The C++/OpenMP CPU port used as bench.py's timing baseline (oracle/cpu_port) against the numpy oracle:
same LM trajectory.  Runs anywhere (no GPU)."""

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene
from oracle import cpu_port, lm


@pytest.mark.parametrize("kw", [
    dict(model=CameraModel.PINHOLE),
    dict(model=CameraModel.SIMPLE_RADIAL, reproj_loss=LossType.CAUCHY, const_point_frac=0.1, depth_frac=0.6),
    dict(model=CameraModel.OPENCV, reproj_loss=LossType.HUBER, depth_prior=False),
])
def test_cpu_port_follows_the_oracle_trajectory(kw):
    prob = make_scene(14, 700, 5, seed=9, **kw)
    osum = lm.solve(prob)
    summ, q, t, X = cpu_port.solve(engine.build_layout(prob))
    assert summ["num_iterations"] == len(osum.iterations) - 1
    oc = np.array([it.cost for it in osum.iterations])
    assert np.max(np.abs(np.array(summ["costs"]) - oc) / oc) <= 1e-6
    assert summ["successful"][1:] == [int(it.step_is_successful) for it in osum.iterations[1:]]
    assert np.max(np.abs(t - osum.state.t)) < 1e-6 and np.max(np.abs(X - osum.state.X)) < 1e-6


def test_cpu_port_exact_solve_and_pcg_agree_on_a_ring():
    """The envelope Cholesky (default: what SPARSE_SCHUR does) and the block-Jacobi PCG land on the same trajectory;
    60 images on a closed ring exercise the reverse Cuthill-McKee ordering (the band wraps around)."""
    prob = make_scene(60, 1500, 5, seed=4)
    lay = engine.build_layout(prob)
    a, qa, ta, Xa = cpu_port.solve(lay)
    b, qb, tb, Xb = cpu_port.solve(lay, linear_solver="pcg", pcg_tolerance=1e-12)
    assert a["num_iterations"] == b["num_iterations"] and a["total_pcg_iterations"] == 0 < b["total_pcg_iterations"]
    assert np.max(np.abs(np.array(a["costs"]) - np.array(b["costs"])) / np.array(a["costs"])) <= 1e-9
    assert np.max(np.abs(ta - tb)) < 1e-7 and np.max(np.abs(Xa - Xb)) < 1e-7
    osum = lm.solve(prob)
    assert a["num_iterations"] == len(osum.iterations) - 1
    assert abs(a["final_cost"] - osum.final_cost) <= 1e-6 * osum.final_cost


def test_cpu_port_structure_only_and_threads():
    prob = make_scene(10, 400, 5, seed=2, depth_loss=LossType.TRIVIAL)
    prob.img_free_mask[:] = 0  # refine_3d_points: every pose constant
    osum = lm.solve(prob)
    lay = engine.build_layout(prob)
    a, _, _, Xa = cpu_port.solve(lay, num_threads=1)
    b, _, _, Xb = cpu_port.solve(lay, num_threads=4)
    assert a["num_iterations"] == b["num_iterations"] == len(osum.iterations) - 1
    assert abs(a["final_cost"] - osum.final_cost) <= 1e-6 * osum.final_cost
    assert np.max(np.abs(Xa - Xb)) <= 1e-9
    assert cpu_port.max_threads() >= 1
