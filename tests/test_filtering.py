"""Post-BA track filter: the product's per-track arithmetic (ba_math.cuh: filter_track, built with g++) against the
numpy restatement of COLMAP's FilterPoints3D (oracle/filtering.py), plus hand-checked cases of the rule itself."""

import ctypes as C
import dataclasses

import numpy as np
import pytest

from mpsfm_b200.problem import CameraModel
from mpsfm_b200.synthetic import make_scene
from oracle import filtering


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def filter_scene(seed=0, model=CameraModel.SIMPLE_RADIAL, num_images=10, num_points=400, track_len=5):
    """A scene with every case of the rule: reprojection outliers, tracks of length 0 / 1 / 2, far-away points (small
    triangulation angle), a point behind one of its cameras, a point on a projection centre."""
    prob = make_scene(num_images, num_points, track_len, model=model, seed=seed, outlier_frac=0.08, noise_px=0.7,
                      pose_rot_noise=0.0, pose_pos_noise=0.0, point_noise=0.002)
    rng = np.random.default_rng(seed + 100)
    keep = np.ones(prob.num_obs, dtype=bool)
    for p, n_left in ((3, 0), (7, 1), (11, 2), (12, 2), (13, 3)):
        ks = np.flatnonzero(prob.obs_pt == p)
        keep[ks[n_left:]] = False
    keep &= ~((prob.obs_pt % 9 == 4) & (rng.random(prob.num_obs) < 0.5))  # ragged tracks
    new_index = np.cumsum(keep) - 1
    dkeep = keep[prob.depth_obs]
    prob = dataclasses.replace(
        prob, obs_img=prob.obs_img[keep], obs_pt=prob.obs_pt[keep], obs_xy=prob.obs_xy[keep],
        depth_obs=new_index[prob.depth_obs[dkeep]].astype(np.int32), depth_d=prob.depth_d[dkeep],
        depth_w=prob.depth_w[dkeep], depth_a=prob.depth_a[dkeep])
    X = prob.pt_xyz.copy()
    q = prob.img_q / np.linalg.norm(prob.img_q, axis=1, keepdims=True)
    from oracle import geometry
    centers = -np.einsum("nji,nj->ni", geometry.rotmat(q), prob.img_t)
    # far points along their viewing direction from image 0's centre: tiny angles, still small reprojection error
    for p in range(20, 60):
        X[p] = centers.mean(axis=0) + (X[p] - centers.mean(axis=0)) * 10.0 ** rng.uniform(1.0, 4.0)
    # a point behind the first camera of its track, one exactly on a projection centre
    k = np.flatnonzero(prob.obs_pt == 70)[0]
    X[70] = centers[prob.obs_img[k]] - 0.5 * (X[70] - centers[prob.obs_img[k]])
    k = np.flatnonzero(prob.obs_pt == 71)[0]
    X[71] = centers[prob.obs_img[k]]
    return dataclasses.replace(prob, pt_xyz=X)


def track_order(prob):
    order = np.argsort(prob.obs_pt, kind="stable")
    counts = np.bincount(prob.obs_pt, minlength=prob.num_points)
    ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    return order, ptr


def host_filter(hc, prob, max_err, min_deg, select=None):
    order, ptr = track_order(prob)
    q = prob.img_q / np.linalg.norm(prob.img_q, axis=1, keepdims=True)
    pose8 = np.zeros((prob.num_images, 8))
    pose8[:, :4], pose8[:, 4:7], pose8[:, 7] = q, prob.img_t, 1.0
    pm_img = np.ascontiguousarray(prob.obs_img[order].astype(np.int32))
    pm_xy = np.ascontiguousarray(prob.obs_xy[order])
    X = np.ascontiguousarray(prob.pt_xyz)
    N, P = prob.num_obs, prob.num_points
    ok = np.zeros(max(N, 1), np.uint8)
    pk = np.zeros(max(P, 1), np.uint8)
    pe = np.zeros(max(P, 1))
    sel = None if select is None else np.ascontiguousarray(select.astype(np.uint8))
    hc.hc_filter_tracks.restype = C.c_longlong
    nf = hc.hc_filter_tracks(C.c_int(P), _p(ptr), _p(pose8), _p(np.ascontiguousarray(prob.img_camera.astype(np.int32))),
                             _p(np.ascontiguousarray(prob.cam_model.astype(np.int32))),
                             _p(np.ascontiguousarray(prob.cam_params)), _p(pm_img), _p(pm_xy), _p(X),
                             C.c_double(max_err), C.c_double(min_deg), None if sel is None else _p(sel), _p(ok), _p(pk),
                             _p(pe))
    obs_keep = np.empty(N, dtype=bool)
    obs_keep[order] = ok[:N] != 0
    return obs_keep, pk[:P] != 0, pe[:P], int(nf)


@pytest.mark.parametrize("model", [CameraModel.PINHOLE, CameraModel.SIMPLE_RADIAL, CameraModel.RADIAL,
                                   CameraModel.OPENCV])
@pytest.mark.parametrize("thresholds", [(4.0, 1.5), (1.0, 0.5), (50.0, 0.0), (0.0, 1.5)])
def test_host_build_of_the_kernel_arithmetic_matches_oracle(hostcheck, model, thresholds):
    prob = filter_scene(seed=int(model), model=model)
    want = filtering.filter_points3D(prob, prob.img_q, prob.img_t, prob.pt_xyz, *thresholds)
    got = host_filter(hostcheck, prob, *thresholds)
    assert np.array_equal(got[0], want[0])
    assert np.array_equal(got[1], want[1])
    assert got[3] == want[3]
    assert np.allclose(got[2], want[2], rtol=1e-9, atol=1e-9)
    # every case of the rule is present in the scene
    if thresholds == (4.0, 1.5):
        assert (~want[0]).sum() > 0 and (~want[1]).sum() > 5 and want[1].sum() > 100
        assert not want[1][3] and not want[1][7] and not want[1][70] and not want[1][71]
        assert (~want[1][20:60]).sum() > 5


def test_selection_leaves_other_points_untouched(hostcheck):
    prob = filter_scene(seed=5)
    sel = np.arange(prob.num_points) % 3 == 0
    want = filtering.filter_points3D(prob, prob.img_q, prob.img_t, prob.pt_xyz, 4.0, 1.5, select=sel)
    got = host_filter(hostcheck, prob, 4.0, 1.5, select=sel)
    for a, b in zip(got[:2], want[:2]):
        assert np.array_equal(a, b)
    assert got[3] == want[3]
    assert want[1][~sel].all() and (want[2][~sel] == -1).all()
    assert want[0][~sel[prob.obs_pt]].all()


def test_rule_by_hand():
    """Two cameras 1 apart looking down +z, one point at depth 10: angle = 2 atan(0.05) = 5.72 deg."""
    from mpsfm_b200.problem import BundleProblem, LossType
    q = np.tile([0.0, 0.0, 0.0, 1.0], (3, 1))
    t = np.array([[0.5, 0, 0], [-0.5, 0, 0], [-0.5, 0, 0]])  # centres (-0.5,0,0), (0.5,0,0), (0.5,0,0)
    X = np.array([[0.0, 0.0, 10.0], [0.0, 0.0, 10.0]])
    f = 100.0
    xy = np.array([[5.0, 0.0], [-5.0, 0.0], [-5.0, 3.0],  # point 0: third observation 3 px off
                   [5.0, 0.0], [-5.0, 0.0]])
    z = np.zeros(0)
    prob = BundleProblem(img_q=q, img_t=t, img_camera=np.zeros(3, np.int32), img_free_mask=np.full(3, 63, np.uint8),
                         cam_model=np.array([int(CameraModel.SIMPLE_PINHOLE)], np.int32),
                         cam_params=np.array([[f, 0.0, 0.0] + [0.0] * 9]), pt_xyz=X, pt_const=np.zeros(2, np.uint8),
                         obs_img=np.array([0, 1, 2, 1, 2], np.int32), obs_pt=np.array([0, 0, 0, 1, 1], np.int32),
                         obs_xy=xy, depth_obs=z.astype(np.int32), depth_d=z, depth_w=z, depth_a=z,
                         shift_scale=np.zeros((3, 2)), reproj_loss=LossType.TRIVIAL, depth_loss=LossType.TRIVIAL)
    ok, pk, pe, nf = filtering.filter_points3D(prob, q, t, X, 2.0, 1.5)
    # point 0: the 3 px observation goes, two remain with 5.72 deg; point 1: both cameras share a centre -> angle 0
    assert ok.tolist() == [True, True, False, False, False]
    assert pk.tolist() == [True, False] and nf == 1 + 2
    assert pe[0] == 0.0 and pe[1] == -1.0
    ok, pk, pe, nf = filtering.filter_points3D(prob, q, t, X, 4.0, 6.0)  # 5.72 deg < 6 deg: everything goes
    assert not pk.any() and not ok.any() and nf == 5
    ok, pk, pe, nf = filtering.filter_points3D(prob, q, t, X, 4.0, 5.0)
    assert pk.tolist() == [True, False] and abs(pe[0] - 1.0) < 1e-12 and nf == 2


def test_optimizer_filter_points3D_host_logic(hostcheck, monkeypatch):
    """Host side of Optimizer.filter_points3D (track gathering, the ObservationManager edits) with the engine call
    replaced by the g++ build of the same arithmetic: the CUDA path itself is covered by tests/test_gpu_filter.py."""
    from scene_factory import build_scene

    from mpsfm_b200 import engine
    from mpsfm_b200.optimizer import Optimizer

    class HostAdjuster:
        def __init__(self, problem, device=0):
            self.problem = problem

        def filter_tracks(self, max_err, min_deg, select=None):
            return host_filter(hostcheck, self.problem, max_err, min_deg, select)

        def close(self):
            pass

    monkeypatch.setattr(engine, "BundleAdjuster", HostAdjuster)
    rec = build_scene(seed=21, num_images=8, num_points=260, pose_noise=(0.0, 0.0), point_noise=0.0)
    rng = np.random.default_rng(0)
    for pid in list(rec.points3D)[::9]:
        rec.points3D[pid].xyz += rng.normal(0, 0.4, 3)
    for pid in list(rec.points3D)[4::13]:
        rec.points3D[pid].xyz *= 300.0
    n_obs = sum(p.track.length() for p in rec.points3D.values())
    n_pts = len(rec.points3D)
    pids = set(list(rec.points3D)[::2]) | {10**9}
    untouched = {p: rec.points3D[p].track.length() for p in rec.points3D if p not in pids}
    nf = Optimizer({}, rec, None).filter_points3D(3.0, 1.5, pids)
    assert 0 < nf == n_obs - sum(p.track.length() for p in rec.points3D.values())
    assert len(rec.points3D) < n_pts
    assert all(rec.points3D[p].track.length() == n for p, n in untouched.items())
    for im in rec.images.values():
        for j, p2 in enumerate(im.points2D):
            if p2.has_point3D():
                assert any((e.image_id, e.point2D_idx) == (im.image_id, j) for e in rec.points3D[p2.point3D_id].track.elements)
    for pid in pids & set(rec.points3D):
        assert rec.points3D[pid].track.length() >= 2 and 0 <= rec.points3D[pid].error <= 3.0
    assert Optimizer({}, rec, None).filter_points3D(3.0, 1.5, pids) == 0  # idempotent


def test_reprojection_error_of_all_camera_models_and_angle_edge_cases(hostcheck):
    """reproj_error2 (ba_math.cuh) for the 11 COLMAP models against the oracle's projection; points behind the camera;
    triangulation_angle on exact and degenerate configurations."""
    from oracle import camera_models, geometry

    hostcheck.hc_reproj_error2.restype = C.c_double
    hostcheck.hc_triangulation_angle.restype = C.c_double
    params = {
        0: [900.0, 500.0, 400.0], 1: [900.0, 950.0, 500.0, 400.0], 2: [900.0, 500.0, 400.0, 0.05],
        3: [900.0, 500.0, 400.0, 0.05, -0.01], 4: [900.0, 950.0, 500.0, 400.0, 0.05, -0.01, 1e-3, -2e-3],
        5: [900.0, 950.0, 500.0, 400.0, 0.02, -0.01, 3e-3, -1e-3],
        6: [900.0, 950.0, 500.0, 400.0, 0.05, -0.01, 1e-3, -2e-3, 3e-3, 0.02, -4e-3, 1e-3],
        7: [900.0, 950.0, 500.0, 400.0, 0.9], 8: [900.0, 500.0, 400.0, 0.03], 9: [900.0, 500.0, 400.0, 0.03, -0.004],
        10: [900.0, 950.0, 500.0, 400.0, 0.02, -0.01, 1e-3, -2e-3, 3e-3, -1e-3, 2e-3, -3e-3],
    }
    rng = np.random.default_rng(1)
    for model, p in params.items():
        prm = np.zeros(12)
        prm[: len(p)] = p
        for _ in range(20):
            q = rng.standard_normal(4)
            q /= np.linalg.norm(q)
            t = rng.standard_normal(3)
            pc = np.array([rng.uniform(-0.4, 0.4), rng.uniform(-0.3, 0.3), 1.0]) * rng.uniform(2.0, 30.0)
            X = geometry.rotmat(q[None])[0].T @ (pc - t)
            obs = rng.uniform(0, 1000, 2)
            pose = np.concatenate([q, t, [1.0]])
            got = hostcheck.hc_reproj_error2(_p(pose), C.c_int(model), _p(prm), _p(np.ascontiguousarray(X)),
                                             C.c_double(obs[0]), C.c_double(obs[1]))
            xy = camera_models.img_from_cam_values(np.array([model]), prm[None], pc[None])[0]
            want = float(np.sum((xy - obs) ** 2))
            assert abs(got - want) <= 1e-9 * want
        # behind the camera (and exactly on the principal plane): COLMAP's "infinite" error
        for z in (-1.0, 0.0, 1e-17):
            pose = np.array([0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0])
            got = hostcheck.hc_reproj_error2(_p(pose), C.c_int(model), _p(prm), _p(np.array([0.1, 0.2, z])),
                                             C.c_double(0.0), C.c_double(0.0))
            assert got == np.finfo(np.float64).max
    ang = lambda a, b, x: hostcheck.hc_triangulation_angle(_p(np.array(a, float)), _p(np.array(b, float)),  # noqa: E731
                                                           _p(np.array(x, float)))
    assert abs(ang([1, 0, 0], [0, 1, 0], [0, 0, 0]) - np.pi / 2) < 1e-15          # right angle
    assert abs(ang([1, 0, 0], [-1, 0, 0], [0, 1, 0]) - np.pi / 2) < 1e-15
    assert abs(ang([1, 0, 0], [-1, 0, 0], [0, 0.1, 0]) - 2 * np.arctan(0.1)) < 1e-14  # obtuse folds to pi - angle
    assert ang([0, 0, 0], [0, 0, 0], [0, 0, 0]) == 0.0                            # zero denominator
    assert ang([1, 2, 3], [4, 5, 6], [1, 2, 3]) == 0.0                            # point on a projection centre
    a = ang([0, 0, 0], [1e-9, 0, 0], [0, 0, 1e3])                                  # nearly parallel rays
    assert np.isnan(a) or 0.0 <= a < 1e-8
    for _ in range(50):
        c1, c2, x = rng.standard_normal((3, 3))
        want = filtering.triangulation_angle(c1, c2, x)
        assert abs(ang(c1, c2, x) - want) <= 1e-14


def test_filter_local_points3D_follows_the_mapper(hostcheck, monkeypatch):
    """MpsfmMapper.filter_local_points3D (reference base.py:769-797) on the Optimizer: risky points (no valid depth in
    any bundle image) at 1.5 degrees, then the bundle's points at the configured angle."""
    import copy

    from scene_factory import build_scene

    from mpsfm_b200 import engine
    from mpsfm_b200.optimizer import Optimizer

    class HostAdjuster:
        def __init__(self, problem, device=0):
            self.problem = problem

        def filter_tracks(self, max_err, min_deg, select=None):
            return host_filter(hostcheck, self.problem, max_err, min_deg, select)

        def close(self):
            pass

    monkeypatch.setattr(engine, "BundleAdjuster", HostAdjuster)
    rec = build_scene(seed=4, num_images=8, num_points=260, pose_noise=(0.0, 0.0), point_noise=0.0)
    rng = np.random.default_rng(2)
    for pid in list(rec.points3D)[::7]:
        rec.points3D[pid].xyz += rng.normal(0, 0.4, 3)
    optim_ids = [5, 4, 6]
    ref = rec.images[5]
    pts = set(ref.point3D_ids(ref.get_observation_point2D_idxs()))
    allp = set()
    for i in optim_ids:
        allp |= set(rec.images[i].point3D_ids(rec.images[i].get_observation_point2D_idxs()))
    bundle = {"optim_ids": optim_ids, "pts3D": pts, "constpoints": allp - pts, "ref_id": 5}
    opt = Optimizer({}, rec, None)
    risky = opt.find_invalid_depth_points(optim_ids)
    assert len(risky) == 3 and all(isinstance(r, set) for r in risky)
    for imid, r in zip(optim_ids, risky):  # restated independently from the image's own accessors
        im = rec.images[imid]
        idx = im.get_observation_point2D_idxs()
        valid = np.asarray(im.depth.valid_at_kps(im.keypoint_coordinates(idx))).reshape(-1)
        assert r == {p for p, v in zip(im.point3D_ids(idx), valid) if not v}
    # expectation: the two pycolmap calls of the reference, through the single-call method on a copy
    rec2 = copy.deepcopy(rec)
    opt2 = Optimizer({}, rec2, None)
    want = opt2.filter_points3D(3.0, 1.5, set.intersection(*risky)) + opt2.filter_points3D(3.0, 2.5, allp)
    got = opt.filter_local_points3D(bundle, 3.0, 2.5)
    assert got == want and got > 0
    assert set(rec.points3D) == set(rec2.points3D)
    for pid in rec.points3D:
        a = [(e.image_id, e.point2D_idx) for e in rec.points3D[pid].track.elements]
        b = [(e.image_id, e.point2D_idx) for e in rec2.points3D[pid].track.elements]
        assert a == b
    assert all(rec.points3D[p].track.length() >= 2 for p in rec.points3D)
    n_before = len(rec.points3D)
    assert opt.filter_all_points3D(3.0, 2.5) >= 0 and len(rec.points3D) <= n_before
