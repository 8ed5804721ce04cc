"""COLMAP binary model files (cameras.bin / images.bin / points3D.bin) <-> the engine's arrays: a hand-packed model
(bytes laid out here with ``struct``, independently of the writer) read and re-written byte for byte, round trips of
synthetic scenes, and the track-filter edits applied to a model."""

import struct

import numpy as np

from mpsfm_b200 import colmap_io
from mpsfm_b200.problem import CameraModel
from mpsfm_b200.synthetic import make_scene
from oracle import filtering

INVALID = 2**64 - 1


def hand_packed_model():
    cams = struct.pack("<Q", 2)
    cams += struct.pack("<IiQQ", 1, 1, 640, 480) + struct.pack("<4d", 500.0, 510.0, 320.0, 240.0)  # PINHOLE
    cams += struct.pack("<IiQQ", 7, 2, 800, 600) + struct.pack("<4d", 700.0, 400.0, 300.0, 0.01)  # SIMPLE_RADIAL
    imgs = struct.pack("<Q", 2)
    imgs += struct.pack("<I7dI", 3, 1.0, 0.0, 0.0, 0.0, 0.1, 0.2, 0.3, 1) + b"a/first.jpg\0"
    imgs += struct.pack("<Q", 3) + struct.pack("<ddQ", 10.5, 20.25, 42) + struct.pack("<ddQ", 1.0, 2.0, INVALID)
    imgs += struct.pack("<ddQ", 300.0, 200.0, 43)
    imgs += struct.pack("<I7dI", 5, 0.5, 0.5, -0.5, 0.5, -1.0, 0.0, 2.0, 7) + b"second.png\0"
    imgs += struct.pack("<Q", 2) + struct.pack("<ddQ", 11.0, 21.0, 42) + struct.pack("<ddQ", 310.0, 190.0, 43)
    pts = struct.pack("<Q", 2)
    pts += struct.pack("<Q3d3BdQ", 42, 1.0, 2.0, 3.0, 255, 128, 0, 0.75, 2) + struct.pack("<IIII", 3, 0, 5, 0)
    pts += struct.pack("<Q3d3BdQ", 43, -1.0, 0.5, 9.0, 1, 2, 3, 1.25, 2) + struct.pack("<IIII", 3, 2, 5, 1)
    return {"cameras.bin": cams, "images.bin": imgs, "points3D.bin": pts}


def test_hand_packed_files_read_and_rewrite_byte_for_byte(tmp_path):
    files = hand_packed_model()
    for name, data in files.items():
        (tmp_path / name).write_bytes(data)
    m = colmap_io.read_model(str(tmp_path))
    assert m.cam_id.tolist() == [1, 7] and m.cam_model.tolist() == [1, 2]
    assert m.cam_width.tolist() == [640, 800] and m.cam_height.tolist() == [480, 600]
    assert m.cam_params[1, :4].tolist() == [700.0, 400.0, 300.0, 0.01] and not m.cam_params[:, 4:].any()
    assert m.img_id.tolist() == [3, 5] and m.img_name == ["a/first.jpg", "second.png"]
    assert m.img_q[1].tolist() == [0.5, -0.5, 0.5, 0.5]  # stored w x y z, held x y z w
    assert m.img_t[0].tolist() == [0.1, 0.2, 0.3] and m.img_camera_id.tolist() == [1, 7]
    assert m.kp_ptr.tolist() == [0, 3, 5] and m.kp_xy[2].tolist() == [300.0, 200.0]
    assert m.kp_point3D_id.tolist() == [42, INVALID, 43, 42, 43]
    assert m.pt_id.tolist() == [42, 43] and m.pt_rgb[0].tolist() == [255, 128, 0] and m.pt_error.tolist() == [0.75, 1.25]
    assert m.track_ptr.tolist() == [0, 2, 4] and m.track_image_id.tolist() == [3, 5, 3, 5]
    assert m.track_point2D_idx.tolist() == [0, 0, 2, 1]
    out = tmp_path / "out"
    colmap_io.write_model(str(out), m)
    for name, data in files.items():
        assert (out / name).read_bytes() == data
    prob, _ = colmap_io.to_bundle_problem(m)
    assert prob.obs_img.tolist() == [0, 1, 0, 1] and prob.obs_pt.tolist() == [0, 0, 1, 1]
    assert prob.obs_xy.tolist() == [[10.5, 20.25], [11.0, 21.0], [300.0, 200.0], [310.0, 190.0]]
    assert prob.img_camera.tolist() == [0, 1]


def model_of(prob, extra_kps=3, seed=0):
    """A COLMAP model holding the problem's images / points, with a few unmatched keypoints per image."""
    rng = np.random.default_rng(seed)
    C, P, K = prob.num_images, prob.num_points, prob.cam_model.shape[0]
    order = np.lexsort((prob.obs_pt, prob.obs_img))  # keypoints of an image in point order, then the unmatched ones
    counts = np.bincount(prob.obs_img, minlength=C)
    kp_ptr = np.concatenate([[0], np.cumsum(counts + extra_kps)]).astype(np.int64)
    M = int(kp_ptr[-1])
    kp_xy, kp_pid = rng.uniform(0, 500, (M, 2)), np.full(M, colmap_io.INVALID_POINT3D, dtype=np.uint64)
    p2d_of_obs = np.empty(prob.num_obs, dtype=np.int64)
    start = np.concatenate([[0], np.cumsum(counts)])
    for rank, k in enumerate(order):
        p2d_of_obs[k] = rank - start[prob.obs_img[k]]
    slot = kp_ptr[prob.obs_img] + p2d_of_obs
    pt_id = (np.arange(P, dtype=np.uint64) * 3 + 11)
    kp_xy[slot], kp_pid[slot] = prob.obs_xy, pt_id[prob.obs_pt]
    torder = np.argsort(prob.obs_pt, kind="stable")
    return colmap_io.ColmapModel(
        cam_id=np.arange(1, K + 1, dtype=np.uint32), cam_model=prob.cam_model.astype(np.int32),
        cam_width=np.full(K, 1600, np.uint64), cam_height=np.full(K, 1200, np.uint64), cam_params=prob.cam_params.copy(),
        img_id=(np.arange(C, dtype=np.uint32) * 2 + 1), img_q=prob.img_q.copy(), img_t=prob.img_t.copy(),
        img_camera_id=(prob.img_camera + 1).astype(np.uint32), img_name=[f"im{k:04d}.jpg" for k in range(C)],
        kp_ptr=kp_ptr, kp_xy=kp_xy, kp_point3D_id=kp_pid, pt_id=pt_id, pt_xyz=prob.pt_xyz.copy(),
        pt_rgb=rng.integers(0, 256, (P, 3)).astype(np.uint8), pt_error=rng.uniform(0, 2, P),
        track_ptr=np.concatenate([[0], np.cumsum(np.bincount(prob.obs_pt, minlength=P))]).astype(np.int64),
        track_image_id=(prob.obs_img[torder].astype(np.uint32) * 2 + 1),
        track_point2D_idx=p2d_of_obs[torder].astype(np.uint32))


def assert_models_equal(a, b):
    for f in ("cam_id", "cam_model", "cam_width", "cam_height", "cam_params", "img_id", "img_q", "img_t", "img_camera_id",
              "kp_ptr", "kp_xy", "kp_point3D_id", "pt_id", "pt_xyz", "pt_rgb", "pt_error", "track_ptr", "track_image_id",
              "track_point2D_idx"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f
    assert a.img_name == b.img_name


def test_round_trip_and_problem_view(tmp_path):
    for model in (CameraModel.PINHOLE, CameraModel.OPENCV):
        prob = make_scene(9, 400, 4, model=model, seed=3, depth_prior=False).normalized()
        m = model_of(prob)
        colmap_io.write_model(str(tmp_path / model.name), m)
        back = colmap_io.read_model(str(tmp_path / model.name))
        assert_models_equal(m, back)  # doubles survive bit for bit
        view, _ = colmap_io.to_bundle_problem(back)
        torder = np.argsort(prob.obs_pt, kind="stable")
        assert np.array_equal(view.obs_img, prob.obs_img[torder]) and np.array_equal(view.obs_pt, prob.obs_pt[torder])
        assert np.array_equal(view.obs_xy, prob.obs_xy[torder]) and np.array_equal(view.pt_xyz, prob.pt_xyz)
        assert np.array_equal(view.img_q, prob.img_q) and np.array_equal(view.cam_params, prob.cam_params)
    empty = colmap_io.ColmapModel(*(np.zeros(0, t) for t in (np.uint32, np.int32, np.uint64, np.uint64)),
                                  np.zeros((0, 12)), np.zeros(0, np.uint32), np.zeros((0, 4)), np.zeros((0, 3)),
                                  np.zeros(0, np.uint32), [], np.zeros(1, np.int64), np.zeros((0, 2)), np.zeros(0, np.uint64),
                                  np.zeros(0, np.uint64), np.zeros((0, 3)), np.zeros((0, 3), np.uint8), np.zeros(0),
                                  np.zeros(1, np.int64), np.zeros(0, np.uint32), np.zeros(0, np.uint32))
    colmap_io.write_model(str(tmp_path / "empty"), empty)
    assert (tmp_path / "empty" / "images.bin").read_bytes() == struct.pack("<Q", 0)
    assert_models_equal(empty, colmap_io.read_model(str(tmp_path / "empty")))


def test_filter_edits_on_a_model(tmp_path):
    prob = make_scene(10, 300, 5, model=CameraModel.SIMPLE_RADIAL, seed=5, depth_prior=False, outlier_frac=0.1,
                      pose_rot_noise=0.0, pose_pos_noise=0.0, point_noise=0.002).normalized()
    m = model_of(prob)
    view, _ = colmap_io.to_bundle_problem(m)
    obs_keep, pt_keep, pt_err, nf = filtering.filter_points3D(view, view.img_q, view.img_t, view.pt_xyz, 4.0, 1.5)
    assert 0 < nf and 0 < (~pt_keep).sum() < prob.num_points
    f = colmap_io.apply_filter(m, obs_keep, pt_keep, pt_err)
    assert len(f.pt_id) == pt_keep.sum() and int(f.track_ptr[-1]) == m.track_ptr[-1] - nf
    assert (np.diff(f.track_ptr) >= 2).all() and np.array_equal(f.pt_error, pt_err[pt_keep])
    assert (f.kp_point3D_id != colmap_io.INVALID_POINT3D).sum() == f.track_ptr[-1]
    colmap_io.write_model(str(tmp_path), f)
    back = colmap_io.read_model(str(tmp_path))
    assert_models_equal(f, back)
    view2, _ = colmap_io.to_bundle_problem(back)  # also checks that tracks and keypoint ids still agree
    again = filtering.filter_points3D(view2, view2.img_q, view2.img_t, view2.pt_xyz, 4.0, 1.5)
    assert again[3] == 0 and again[0].all() and again[1].all()
    q2 = view2.img_q.copy()
    q2[:, :3] *= -1.0
    colmap_io.update_state(back, q2, view2.img_t + 1.0, view2.pt_xyz * 2.0)
    assert np.array_equal(back.pt_xyz, view2.pt_xyz * 2.0) and np.array_equal(back.img_t, view2.img_t + 1.0)


def test_depth_cache_twin_round_trip(tmp_path):
    """The reference's depths.h5 (reconstruction/base.py:122-133) or, without h5py, its .npz twin with the same keys."""
    from types import SimpleNamespace as NS

    rng = np.random.default_rng(0)
    images = {
        4: NS(depth=NS(data=rng.random((6, 8)), continuity_mask=rng.random((6, 8)) > 0.5, scale=1.25,
                       data_prior=rng.random((6, 8)))),
        9: NS(depth=NS(data=rng.random((5, 7)), continuity_mask=None, scale=None, data_prior=None)),
    }
    written = colmap_io.cache_depths(images, tmp_path / "depths.h5")
    back = colmap_io.read_cached_depths(written)
    assert set(back["depths"]) == {4, 9} and set(back["continuity"]) == {4} and set(back["pdepth"]) == {4}
    assert np.array_equal(back["depths"][9], images[9].depth.data)
    assert np.array_equal(back["continuity"][4], images[4].depth.continuity_mask)
    assert float(back["scales"][4]) == 1.25 and 9 not in back["scales"]
