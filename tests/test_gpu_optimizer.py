"""The drop-in ``Optimizer`` end to end on the GPU: same entry points as the reference's class
(/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:244-293), results against the CPU oracle run on the very
same assembled problem, and the in-place update contract on the pycolmap-shaped scene."""

import copy

import numpy as np
import pytest

from scene_factory import build_scene

from mpsfm_b200.optimizer import Optimizer
from oracle import lm
from oracle.residuals import State, evaluate

pytestmark = pytest.mark.gpu


def _angle(qa, qb):
    return 2.0 * np.arccos(np.abs(np.sum(qa * qb, axis=-1)).clip(max=1.0))


def _snapshot(rec):
    return ({i: (im.cam_from_world.rotation.quat.copy(), im.cam_from_world.translation.copy()) for i, im in rec.images.items()},
            {p: pt.xyz.copy() for p, pt in rec.points3D.items()})


def _buffers(rec):
    return ({i: (id(im.cam_from_world.rotation.quat), id(im.cam_from_world.translation)) for i, im in rec.images.items()},
            {p: id(pt.xyz) for p, pt in rec.points3D.items()})


@pytest.mark.parametrize("mode", ["global", "local"])
def test_ba_updates_the_reconstruction_in_place_like_ceres(mode):
    rec = build_scene(seed=11, num_images=8, num_points=260)
    opt = Optimizer({}, rec, None)
    opt.update_truncation_multiplier(rec.reg_image_ids())
    if mode == "global":
        optim_ids = [2, 1, 3, 4, 5, 6, 7, 8]
        bundle = {"optim_ids": optim_ids, "pts3D": set(rec.points3D)}
    else:
        optim_ids = [5, 4, 6, 3]
        ref = rec.images[5]
        bundle = {"optim_ids": optim_ids, "pts3D": set(ref.point3D_ids(ref.get_observation_point2D_idxs())), "ref_id": 5}
    # what the oracle does with the identical problem (assembled from an identical copy of the scene)
    rec_o = copy.deepcopy(rec)
    opt_o = Optimizer({}, rec_o, None)
    opt_o.truncation_multiplier = opt.truncation_multiplier
    asm_o = opt_o.assemble(bundle, fix_pose=False, mode=mode, allow_scale_filter=True)
    osum = lm.solve(asm_o.problem)

    before_pose, before_pts = _snapshot(rec)
    buf_pose, buf_pts = _buffers(rec)
    out, ok = opt.ba(bundle, mode=mode, allow_scale_filter=True)
    assert ok is True
    summ, asm = opt.last_summary, opt.last_assembly
    assert summ.num_iterations == len(osum.iterations) - 1
    assert abs(summ.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    # same buffers, new values (Ceres parameter blocks alias pycolmap memory in the reference)
    assert _buffers(rec) == (buf_pose, buf_pts)
    pr = asm.problem
    for k, imid in enumerate(asm.image_ids):
        q, t = rec.images[imid].cam_from_world.rotation.quat, rec.images[imid].cam_from_world.translation
        if k < asm.num_optim_images:
            assert abs(np.linalg.norm(q) - 1) < 1e-12  # COLMAP normalises the images of the config
            assert _angle(q, osum.state.q[k]) < 1e-6 and np.max(np.abs(t - osum.state.t[k])) < 1e-6
        q0, t0 = before_pose[imid]
        if k == 0 or k >= asm.num_optim_images:  # gauge anchor / images outside the bundle: untouched
            assert np.allclose(q, q0 / np.linalg.norm(q0) if k == 0 else q0, rtol=0, atol=1e-15) and np.array_equal(t, t0)
        elif k == 1:
            assert t[0] == t0[0] and not np.array_equal(t[1:], t0[1:])  # SubsetManifold(3, [0])
        else:
            assert not np.array_equal(t, t0)
    moved = 0
    for j, pid in enumerate(asm.point_ids.tolist()):
        x = rec.points3D[pid].xyz
        if pr.pt_const[j]:
            assert np.array_equal(x, before_pts[pid])
        else:
            assert np.max(np.abs(x - osum.state.X[j])) < 1e-6
            moved += 1
    assert moved > 0
    untouched = set(rec.points3D) - set(asm.point_ids.tolist())
    assert all(np.array_equal(rec.points3D[p].xyz, before_pts[p]) for p in untouched)


def test_refine_3d_points_keeps_poses():
    rec = build_scene(seed=12, num_images=6, num_points=180)
    opt = Optimizer({}, rec, None)
    bundle = {"optim_ids": [1, 2, 3, 4], "pts3D": set(rec.points3D)}
    rec_o = copy.deepcopy(rec)
    asm_o = Optimizer({}, rec_o, None).assemble(bundle, fix_pose=True, depth_loss_name="trivial", depth_type="prior")
    osum = lm.solve(asm_o.problem)
    before_pose, _ = _snapshot(rec)
    out, ok = opt.refine_3d_points(bundle, depth_type="prior")
    assert ok is True and opt.last_assembly.problem.depth_loss == 0  # ref3d_loss_name = "trivial"
    assert np.all(opt.last_assembly.problem.img_free_mask == 0)
    for imid, (q0, t0) in before_pose.items():
        im = rec.images[imid]
        assert np.array_equal(im.cam_from_world.translation, t0)
        assert np.allclose(im.cam_from_world.rotation.quat, q0 / np.linalg.norm(q0) if imid in bundle["optim_ids"] else q0, atol=1e-15)
    asm = opt.last_assembly
    assert opt.last_summary.num_iterations == len(osum.iterations) - 1
    assert abs(opt.last_summary.final_cost - osum.final_cost) <= 1e-6 * osum.final_cost
    for j in np.flatnonzero(asm.problem.pt_const == 0):
        assert np.max(np.abs(rec.points3D[int(asm.point_ids[j])].xyz - osum.state.X[j])) < 1e-6


def test_calculate_point_covs():
    rec = build_scene(seed=13, num_images=6, num_points=150)
    opt = Optimizer({}, rec, None)
    bundle = {"optim_ids": [1, 2, 3, 4, 5, 6], "pts3D": set(rec.points3D)}
    opt.calculate_point_covs(bundle)
    asm = Optimizer({}, copy.deepcopy(rec), None).assemble(bundle, fix_pose=False, all_points_variable=True,
                                                           reproj_loss="TRIVIAL", with_depth=False)
    lin = evaluate(asm.problem, State.of(asm.problem), jacobians=True)
    Hpp = lm.segment_sum(asm.problem.obs_pt, np.einsum("nki,nkj->nij", lin.Jp, lin.Jp), asm.problem.num_points)
    assert set(rec.point_covs.data) == set(bundle["pts3D"])
    for j, pid in enumerate(asm.point_ids.tolist()):
        cov = rec.point_covs.data[pid]
        ref = np.linalg.inv(Hpp[j])
        assert cov.shape == (3, 3) and np.max(np.abs(cov - ref)) <= 1e-9 * np.max(np.abs(ref))
