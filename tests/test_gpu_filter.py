"""Post-BA track filter on the GPU (mpba_filter_tracks through the C ABI) against the numpy restatement of COLMAP's
FilterPoints3D (oracle/filtering.py): flags and the filtered count exact, errors to 1e-9 relative."""

import copy

import numpy as np
import pytest

from scene_factory import build_scene
from test_filtering import filter_scene

from mpsfm_b200 import engine
from mpsfm_b200.optimizer import Optimizer
from mpsfm_b200.problem import CameraModel
from mpsfm_b200.synthetic import make_scene
from oracle import filtering

pytestmark = pytest.mark.gpu


def _check(got, want):
    assert np.array_equal(got[0], want[0])
    assert np.array_equal(got[1], want[1])
    assert got[3] == want[3]
    assert np.allclose(got[2], want[2], rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("model", [CameraModel.PINHOLE, CameraModel.SIMPLE_RADIAL, CameraModel.OPENCV])
@pytest.mark.parametrize("thresholds", [(4.0, 1.5), (1.0, 0.5), (0.0, 1.5)])
def test_filter_matches_oracle(model, thresholds):
    prob = filter_scene(seed=int(model), model=model)
    want = filtering.filter_points3D(prob, prob.img_q, prob.img_t, prob.pt_xyz, *thresholds)
    ba = engine.BundleAdjuster(prob, device=0)
    _check(ba.filter_tracks(*thresholds), want)
    if thresholds == (4.0, 1.5):
        assert 0 < want[1].sum() < prob.num_points and 0 < want[0].sum() < prob.num_obs


def test_selection_and_state_after_a_solve():
    prob = filter_scene(seed=2)
    ba = engine.BundleAdjuster(prob, device=0)
    sel = np.arange(prob.num_points) % 3 != 1
    want = filtering.filter_points3D(prob, prob.img_q, prob.img_t, prob.pt_xyz, 4.0, 1.5, select=sel)
    _check(ba.filter_tracks(4.0, 1.5, select=sel), want)
    ba.close()
    # the filter reads the handle's CURRENT state: after a solve that is the refined one
    prob = make_scene(8, 300, 4, seed=1, outlier_frac=0.1)
    ba = engine.BundleAdjuster(prob, device=0)
    before = ba.filter_tracks(4.0, 1.5)
    ba.solve()
    q, t, X = ba.state()
    want = filtering.filter_points3D(prob, q, t, X, 4.0, 1.5)
    after = ba.filter_tracks(4.0, 1.5)
    _check(after, want)
    assert after[1].sum() > before[1].sum()  # the noisy start state loses more points than the refined one


def test_filter_is_idempotent_at_scale():
    """1 M observations (no oracle at this size): the count matches the flags, unobserved / short tracks go, and a
    second pass over the surviving observations removes nothing."""
    import dataclasses

    prob = make_scene(200, 200_000, 5, model=CameraModel.SIMPLE_RADIAL, seed=9, outlier_frac=0.05, noise_px=0.7,
                      pose_rot_noise=0.0, pose_pos_noise=0.0, point_noise=0.002, depth_prior=False)
    ba = engine.BundleAdjuster(prob, device=0)
    ok, pk, pe, nf = ba.filter_tracks(4.0, 1.5)
    assert nf == (~ok).sum() and 0 < nf < prob.num_obs // 4
    assert np.array_equal(np.bincount(prob.obs_pt[ok], minlength=prob.num_points) >= 2, pk)
    assert (pe[pk] >= 0).all() and (pe[pk] <= 4.0).all() and (pe[~pk] == -1).all()
    ba.close()
    new_pt = np.cumsum(pk) - 1
    sub = dataclasses.replace(prob, obs_img=prob.obs_img[ok], obs_pt=new_pt[prob.obs_pt[ok]].astype(np.int32),
                              obs_xy=prob.obs_xy[ok], pt_xyz=prob.pt_xyz[pk], pt_const=prob.pt_const[pk])
    ba2 = engine.BundleAdjuster(sub, device=0)
    ok2, pk2, pe2, nf2 = ba2.filter_tracks(4.0, 1.5)
    assert nf2 == 0 and ok2.all() and pk2.all()
    assert np.allclose(pe2, pe[pk], rtol=1e-12, atol=1e-12)


def test_optimizer_filter_points3D_edits_the_reconstruction_like_pycolmap():
    rec = build_scene(seed=21, num_images=8, num_points=260, pose_noise=(0.0, 0.0), point_noise=0.0)
    rng = np.random.default_rng(0)
    for pid in list(rec.points3D)[::9]:  # gross errors on a few points, far-away points on others
        rec.points3D[pid].xyz += rng.normal(0, 0.4, 3)
    for pid in list(rec.points3D)[4::13]:
        rec.points3D[pid].xyz *= 300.0
    opt = Optimizer({}, rec, None)
    pids = set(list(rec.points3D)[::2]) | {10**9}  # an id that does not exist is skipped, as in COLMAP
    before = copy.deepcopy(rec)
    # expectation: the oracle on the arrays of the selected tracks, applied by hand
    exp_tracks, exp_nf = {}, 0
    for pid in sorted(p for p in pids if p in before.points3D):
        els = [(e.image_id, e.point2D_idx) for e in before.points3D[pid].track.elements]
        ids = sorted({i for i, _ in els})
        row = {i: k for k, i in enumerate(ids)}
        q = np.array([before.images[i].cam_from_world.rotation.quat for i in ids])
        t = np.array([before.images[i].cam_from_world.translation for i in ids])
        cam = before.rec.cameras[before.images[ids[0]].camera_id]
        prm = np.zeros((1, 12))
        prm[0, : len(cam.params)] = cam.params
        from mpsfm_b200.optimizer import _model_id
        from types import SimpleNamespace
        pr = SimpleNamespace(obs_img=np.array([row[i] for i, _ in els]), obs_pt=np.zeros(len(els), int),
                             obs_xy=np.array([before.images[i].points2D[j].xy for i, j in els]).reshape(-1, 2),
                             img_camera=np.zeros(len(ids), int), cam_model=np.array([_model_id(cam.model)]),
                             cam_params=prm, num_obs=len(els), num_points=1)
        ok, pk, pe, nf = filtering.filter_points3D(pr, q, t, before.points3D[pid].xyz[None], 3.0, 1.5)
        exp_nf += nf
        exp_tracks[pid] = [el for el, k in zip(els, ok) if k] if pk[0] else None
    nf = opt.filter_points3D(3.0, 1.5, pids)
    assert nf == exp_nf and nf > 0
    assert any(v is None for v in exp_tracks.values()) and any(v for v in exp_tracks.values())
    for pid, want in exp_tracks.items():
        if want is None:
            assert pid not in rec.points3D
        else:
            assert [(e.image_id, e.point2D_idx) for e in rec.points3D[pid].track.elements] == want
            assert rec.points3D[pid].error >= 0
    for pid in before.points3D:  # unselected points untouched
        if pid not in pids:
            assert len(rec.points3D[pid].track.elements) == len(before.points3D[pid].track.elements)
    # image side of the edits: a deleted observation no longer points at a 3D point
    for im in rec.images.values():
        for j, p2 in enumerate(im.points2D):
            if p2.has_point3D():
                assert any((e.image_id, e.point2D_idx) == (im.image_id, j) for e in rec.points3D[p2.point3D_id].track.elements)
