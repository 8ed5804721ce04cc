"""Normal-prior rows (EXTENSION, "parity n/a": the reference has no such residual -- SURVEY.md section 8 row a-9) on the CUDA
path, through the C ABI, against their definition in oracle/residuals.py at the gates of the reference rows: cost
1e-12, normal-equation blocks 1e-11, LM trajectory / final state 1e-6.  And: a problem WITHOUT the rows is untouched
by them (same kernels as before, same numbers)."""

import dataclasses

import numpy as np
import pytest

from mpsfm_b200 import engine, normal_prior
from mpsfm_b200.problem import CameraModel, LossType
from mpsfm_b200.synthetic import make_scene
from oracle import lm
from oracle.residuals import State, evaluate, free_columns

pytestmark = pytest.mark.gpu

SCENES = {
    "local_14_simple_radial": dict(num_images=14, num_points=900, model=CameraModel.SIMPLE_RADIAL, depth_frac=0.7),
    "local_14_opencv_huber": dict(num_images=14, num_points=900, model=CameraModel.OPENCV, depth_loss=LossType.HUBER,
                                  const_point_frac=0.1),
    # beyond the dense direct solve: Schur + cluster PCG, host LM loop vs device loop both covered below
    "global_60_fisheye": dict(num_images=60, num_points=4000, model=CameraModel.OPENCV_FISHEYE,
                              reproj_loss=LossType.CAUCHY),
}


def relerr(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def _angle(qa, qb):
    return 2.0 * np.arccos(np.abs(np.sum(qa * qb, axis=1)).clip(max=1.0))


@pytest.fixture(scope="module", params=sorted(SCENES))
def scene(request):
    kw = dict(SCENES[request.param])
    base = make_scene(kw.pop("num_images"), kw.pop("num_points"), 5, seed=11, **kw)
    prob = normal_prior.synthetic_normal_priors(base, frac=0.6, seed=3)
    return base, prob, engine.BundleAdjuster(prob, device=0)


def test_cost_and_normal_equation_blocks(scene):
    base, prob, ba = scene
    ba.reset_state()
    st = State.of(prob)
    ref = evaluate(prob, st, jacobians=True)
    rr, rd, cost, fixed = ba.evaluate()
    assert abs(cost - ref.cost) <= 1e-12 * ref.cost and abs(fixed - ref.fixed_cost) <= 1e-12 * max(ref.fixed_cost, 1.0)
    assert cost > evaluate(base, st, jacobians=False).cost * (1 + 1e-6)  # the rows are in the cost
    assert np.max(np.abs(rr - ref.r_reproj) / (np.abs(ref.r_reproj) + 1e-6)) < 1e-9  # the other rows are as before
    Hcc, gc, Hpp, gp, cost2, _ = ba.linearize()
    ne = lm.NormalEquations(prob, ref, free_columns(prob).astype(float),
                            np.repeat((prob.pt_const == 0)[:, None], 3, axis=1).astype(float))
    assert relerr(Hcc, ne.Hcc) < 1e-11 and relerr(gc, ne.gc) < 1e-11
    assert relerr(Hpp, ne.Hpp) < 1e-11 and relerr(gp, ne.gp) < 1e-11
    assert abs(cost2 - ref.cost) <= 1e-12 * ref.cost


def test_full_solve_follows_the_definition(scene):
    base, prob, ba = scene
    osum = lm.solve(prob)
    for loop in (1, 2):  # host controller, device controller (CUDA-graph WHILE)
        ba.reset_state()
        summ = ba.solve(engine.default_options(lm_loop=loop))
        assert summ.termination == osum.termination
        assert summ.num_iterations == len(osum.iterations) - 1
        assert abs(summ.initial_cost - osum.initial_cost) <= 1e-12 * osum.initial_cost
        for a, b in zip(summ.iterations, osum.iterations):
            assert abs(a["cost"] - b.cost) <= 1e-6 * b.cost
            assert bool(a["step_is_successful"]) == bool(b.step_is_successful)
        q, t, X = ba.state()
        assert np.max(_angle(q, osum.state.q)) < 1e-6
        assert np.max(np.abs(t - osum.state.t)) < 1e-6 and np.max(np.abs(X - osum.state.X)) < 1e-6
    # ... and it is another optimum than without the rows
    obase = lm.solve(base)
    assert np.max(np.abs(osum.state.X - obase.state.X)) > 1e-4


def test_zero_records_switch_the_rows_off(scene):
    """A problem without the arrays runs the kernels that never read them (the <false> instantiations: the code of
    before the extension).  All-zero records run the <true> kernels with every row off: the same problem through
    other code (other FMA contraction), so the same answer to rounding."""
    base, prob, _ = scene
    s_base = engine.BundleAdjuster(base, device=0)
    r0 = s_base.solve()
    q0, t0, X0 = s_base.state()
    zero = dataclasses.replace(prob, plane_m=np.zeros_like(prob.plane_m), plane_c=np.zeros_like(prob.plane_c))
    s_zero = engine.BundleAdjuster(zero, device=0)
    r1 = s_zero.solve()
    q1, t1, X1 = s_zero.state()
    assert r1.num_iterations == r0.num_iterations and abs(r1.final_cost - r0.final_cost) <= 1e-10 * r0.final_cost
    assert r1.initial_cost == pytest.approx(r0.initial_cost, rel=1e-13)
    assert np.max(np.abs(q1 - q0)) < 1e-8 and np.max(np.abs(t1 - t0)) < 1e-7 and np.max(np.abs(X1 - X0)) < 1e-7


def test_sharded_path_on_one_gpu_carries_the_rows():
    prob = normal_prior.synthetic_normal_priors(
        make_scene(40, 3000, 5, seed=5, model=CameraModel.SIMPLE_RADIAL), frac=0.5, seed=1)
    ba = engine.BundleAdjuster(prob, device=0)
    s = ba.solve(engine.default_options(lm_loop=1))
    sh = engine.BundleAdjuster(prob, device=0, shard=(0, 1))
    s1 = sh.solve(engine.default_options(lm_loop=1))
    assert s1.num_iterations == s.num_iterations and abs(s1.final_cost - s.final_cost) <= 1e-9 * s.final_cost
    for a, b in zip(ba.state(), sh.state()):
        assert np.max(np.abs(a - b)) < 1e-7


def test_bad_plane_arguments_are_rejected():
    prob = normal_prior.synthetic_normal_priors(make_scene(8, 300, 4, seed=2), frac=0.5, seed=1)
    ba = engine.BundleAdjuster(prob, device=0)
    ba.prob.plane_loss_scale = 0.0
    with pytest.raises(engine.EngineError, match="plane_loss_scale"):
        ba.solve()
    ba.prob.plane_loss_scale = 2.0
    keep = ba.prob.cm_plane
    ba.prob.cm_plane = None
    with pytest.raises(engine.EngineError, match="both"):
        ba.solve()
    ba.prob.cm_plane = keep
    assert ba.solve().termination == "CONVERGENCE"


def test_optimizer_ba_with_the_option_follows_the_oracle():
    import copy
    import sys

    sys.path.insert(0, __file__.rsplit("/", 1)[0])
    from mpsfm_b200.optimizer import Optimizer
    from scene_factory import attach_normals, build_scene

    rec = attach_normals(build_scene(seed=11, num_images=8, num_points=260))
    conf = {"gpu": {"normal_prior": {"enabled": True, "rel_sigma": 0.05}}}
    bundle = {"optim_ids": [2, 1, 3, 4, 5, 6, 7, 8], "pts3D": set(rec.points3D)}
    opt = Optimizer(conf, rec, None)
    opt.update_truncation_multiplier(rec.reg_image_ids())
    opt_o = Optimizer(conf, copy.deepcopy(rec), None)
    opt_o.truncation_multiplier = opt.truncation_multiplier
    asm_o = opt_o.assemble(bundle, fix_pose=False, mode="global", allow_scale_filter=True)
    assert asm_o.problem.num_plane > 0
    osum = lm.solve(asm_o.problem)
    out, ok = opt.ba(bundle, mode="global", allow_scale_filter=True)
    assert ok is True
    summ = opt.last_summary
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    for k, pid in enumerate(opt.last_assembly.point_ids):
        assert np.max(np.abs(rec.points3D[int(pid)].xyz - osum.state.X[k])) < 1e-6
