"""The native host layout builder (mpba_host_layout_create / mpba_host_gather) against the numpy statement of
the same layout: every index array bit for bit, on ragged and degenerate problems.  Needs no GPU."""

import dataclasses

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.synthetic import make_scene

FIELDS = [f.name for f in dataclasses.fields(engine.HostLayout)]


def _same(a, b):
    for name in FIELDS:
        x, y = getattr(a, name), getattr(b, name)
        if isinstance(x, np.ndarray):
            assert x.shape == y.shape, name
            assert np.array_equal(x, y), name
            assert x.dtype == y.dtype, (name, x.dtype, y.dtype)
        else:
            assert x == y, name


@pytest.mark.parametrize("C,P,L,seed", [(6, 200, 3, 0), (15, 3000, 5, 1), (40, 9000, 6, 2), (3, 10, 2, 3)])
def test_native_layout_equals_numpy(C, P, L, seed):
    prob = make_scene(C, P, L, seed=seed, depth_prior=True)
    _same(engine.build_layout(prob), engine.build_layout_numpy(prob))


def test_shuffled_observations_constant_points_and_gaps():
    prob = make_scene(12, 5000, 5, seed=5, depth_prior=True)
    rng = np.random.default_rng(0)
    perm = rng.permutation(prob.num_obs)
    inv = np.empty_like(perm)
    inv[perm] = np.arange(perm.size)
    keep = (prob.obs_pt % 11 != 0) & (prob.obs_img != 4)  # unobserved points, an image without observations
    sel = perm[keep[perm]]
    new_index = np.full(prob.num_obs, -1)
    new_index[sel] = np.arange(sel.size)
    dkeep = new_index[prob.depth_obs] >= 0
    prob = dataclasses.replace(
        prob, obs_img=prob.obs_img[sel], obs_pt=prob.obs_pt[sel], obs_xy=prob.obs_xy[sel],
        depth_obs=new_index[prob.depth_obs[dkeep]].astype(np.int32), depth_d=prob.depth_d[dkeep],
        depth_w=prob.depth_w[dkeep], depth_a=prob.depth_a[dkeep])
    prob.pt_const[::7] = 1
    _same(engine.build_layout(prob), engine.build_layout_numpy(prob))


def test_structure_only_and_empty():
    prob = make_scene(5, 100, 3, seed=7)
    prob.img_free_mask[:] = 0
    a, b = engine.build_layout(prob), engine.build_layout_numpy(prob)
    _same(a, b)
    assert a.item_a.size == 0 and a.num_blocks == 5
    empty = dataclasses.replace(prob, obs_img=prob.obs_img[:0], obs_pt=prob.obs_pt[:0], obs_xy=prob.obs_xy[:0],
                                depth_obs=prob.depth_obs[:0], depth_d=prob.depth_d[:0], depth_w=prob.depth_w[:0],
                                depth_a=prob.depth_a[:0])
    _same(engine.build_layout(empty), engine.build_layout_numpy(empty))


def test_long_tracks_and_threaded_placement_is_deterministic(monkeypatch):
    prob = make_scene(30, 6000, 12, seed=9)
    ref = engine.build_layout_numpy(prob)
    for threads in ("1", "3", "16"):
        monkeypatch.setenv("MPBA_HOST_THREADS", threads)
        _same(engine.build_layout(prob), ref)


def test_out_of_range_index_is_rejected():
    prob = make_scene(4, 50, 3, seed=1)
    lib = engine.load_library()
    import ctypes as C

    img = np.ascontiguousarray(prob.obs_img, dtype=np.int32).copy()
    img[3] = 99
    pt = np.ascontiguousarray(prob.obs_pt, dtype=np.int32)
    view, handle = engine.MpbaHostLayoutView(), C.c_void_p()
    rc = lib.mpba_host_layout_create(4, 50, img.size, img.ctypes.data, pt.ctypes.data,
                                     np.zeros(50, np.uint8).ctypes.data, 0, C.byref(handle), C.byref(view))
    assert rc == -1


def test_index_arrays_outlive_the_layout_object():
    """build_layout returns views of the native builder's buffers: the memory must stay alive as long as any array
    (or a slice of one, or a torch tensor made from one) does, and pickles must be plain copies."""
    import gc
    import pickle

    from mpsfm_b200.synthetic import make_scene

    prob = make_scene(12, 900, 4, seed=5)
    want = engine.build_layout_numpy(prob)
    lay = engine.build_layout(prob)
    item_a, tail = lay.item_a, lay.pm_src[100:200]
    blob = pickle.dumps(lay.s_col)
    del lay
    gc.collect()
    for _ in range(3):  # churn the native allocator
        engine.build_layout(make_scene(12, 900, 4, seed=6))
    gc.collect()
    assert np.array_equal(item_a, want.item_a) and np.array_equal(tail, want.pm_src[100:200])
    assert np.array_equal(pickle.loads(blob), want.s_col)
