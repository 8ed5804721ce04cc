"""Host-side layout (numpy) and library loading: no GPU needed."""

import ctypes as C
import os

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.synthetic import make_scene


def test_library_exports_every_symbol_of_the_header():
    lib = engine.load_library()
    hdr = open(os.path.join(os.path.dirname(engine.__file__), "..", "include", "mpba.h")).read()
    import re

    declared = set(re.findall(r"\b(mpba_[a-z_]+)\s*\(", hdr))
    assert declared == set(engine.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.mpba_version() == 100
    opt = engine.default_options()
    assert opt.max_num_iterations == 50 and opt.initial_trust_region_radius == 1e4
    assert opt.function_tolerance == 1e-6 and opt.min_relative_decrease == 1e-3


def test_struct_sizes_match_the_c_abi():
    # natural alignment, same as the C compiler's
    assert C.sizeof(engine.MpbaIteration) == 64
    assert C.sizeof(engine.MpbaOptions) == 16 + 10 * 8 + 32
    assert C.sizeof(engine.MpbaProblem) % 8 == 0
    from mpsfm_b200 import integration

    assert C.sizeof(integration.MpbaIntegrationConf) == 8 + 10 * 8
    assert C.sizeof(integration.MpbaIntegrationJob) == 16 + 32 + 11 * 8 + 8 + 8
    assert C.sizeof(integration.MpbaIntegrationResult) == 16 + 8 + 33 * 8 + 32 * 4
    conf = integration.default_conf()
    assert (conf.max_iter, conf.cg_max_iter, conf.tol, conf.cg_tol, conf.scale_filter_factor) == (10, 5000, 5e-2, 1e-3, 1.5)


def test_create_without_gpu_fails_loudly():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    prob = make_scene(4, 20, 3, seed=0)
    with pytest.raises(engine.EngineUnavailable):
        engine.BundleAdjuster(prob)
    lib = engine.load_library()
    h = C.c_void_p()
    assert lib.mpba_create(C.byref(h), 0, None) == -2  # MPBA_E_CUDA


@pytest.mark.parametrize("seed", [0, 1])
def test_layout_orders_and_structure(seed):
    prob = make_scene(9, 200, 4, seed=seed, depth_frac=0.6, const_point_frac=0.2)
    lay = engine.build_layout(prob)
    N, Cn, P = prob.num_obs, prob.num_images, prob.num_points
    # both orders are permutations of the natural order
    assert np.array_equal(np.sort(lay.pm_src), np.arange(N))
    assert np.array_equal(np.sort(lay.cm_src), np.arange(N))
    assert np.array_equal(lay.pm_img, prob.obs_img[lay.pm_src])
    assert np.array_equal(lay.pm_xy, prob.obs_xy[lay.pm_src])
    assert np.array_equal(lay.cm_pt, prob.obs_pt[lay.cm_src])
    # tracks contiguous, images ascending inside a track
    for p in range(P):
        sl = slice(lay.pt_ptr[p], lay.pt_ptr[p + 1])
        assert np.all(prob.obs_pt[lay.pm_src[sl]] == p)
        assert np.all(np.diff(lay.pm_img[sl]) >= 0)
    for i in range(Cn):
        sl = slice(lay.img_ptr[i], lay.img_ptr[i + 1])
        assert np.all(prob.obs_img[lay.cm_src[sl]] == i)
    # chunks tile the image order
    assert lay.chunk_begin[0] == 0 and lay.chunk_end[-1] == N
    assert np.all(lay.chunk_end - lay.chunk_begin <= engine.chunk_size(N))
    lib = engine.load_library()
    lib.mpba_chunk_size.restype, lib.mpba_chunk_size.argtypes = C.c_int32, [C.c_int64]
    assert all(lib.mpba_chunk_size(n) == engine.chunk_size(n) for n in (0, 1, 75_775, 75_776, 12_000, 303_103, 303_104, 10**7))
    assert engine.chunk_size(12_000) == 128 and engine.chunk_size(300_000) == 256 and engine.chunk_size(5_000_000) == 1024
    assert np.all(lay.chunk_end[:-1] == lay.chunk_begin[1:])
    # depth rows: w > 0 exactly on the prior observations
    has = np.zeros(N, bool)
    has[prob.depth_obs] = True
    assert np.array_equal(lay.pm_depth[1] > 0, has[lay.pm_src])
    # reduced-system structure: symmetric, diagonal present, pairs map to the right block
    rows = np.repeat(np.arange(Cn), np.diff(lay.s_row_ptr))
    keys = set(zip(rows.tolist(), lay.s_col.tolist()))
    assert all((j, i) in keys for i, j in keys)
    assert all(lay.s_col[lay.s_diag[i]] == i and rows[lay.s_diag[i]] == i for i in range(Cn))
    assert np.array_equal(rows[lay.s_transpose], lay.s_col) and np.array_equal(lay.s_col[lay.s_transpose], rows)
    assert np.array_equal(lay.s_row, rows)
    # Schur work lists: one unit per upper block; every (track, a <= b) pair of a variable point appears
    # exactly once, in the unit of block (img_a, img_b), sorted by track inside the unit
    assert np.array_equal(lay.u_block, np.flatnonzero(rows <= lay.s_col))
    expect = set()
    for p in range(P):
        if prob.pt_const[p]:
            continue
        for a in range(lay.pt_ptr[p], lay.pt_ptr[p + 1]):
            for b in range(a, lay.pt_ptr[p + 1]):
                expect.add((a, b))
    got = set(zip(lay.item_a.tolist(), lay.item_b.tolist()))
    assert got == expect and len(lay.item_a) == len(expect)
    for u, blk in enumerate(lay.u_block):
        sl = slice(lay.u_item_ptr[u], lay.u_item_ptr[u + 1])
        assert np.all(lay.pm_img[lay.item_a[sl]] == rows[blk]) and np.all(lay.pm_img[lay.item_b[sl]] == lay.s_col[blk])
        assert np.all(np.diff(lay.item_a[sl]) >= 0)


def test_layout_edge_cases():
    # an image without observations, a point without observations, no depth priors at all
    prob = make_scene(6, 40, 3, seed=4, depth_prior=False)
    keep = prob.obs_img != 3
    prob.obs_img, prob.obs_pt, prob.obs_xy = prob.obs_img[keep], prob.obs_pt[keep], prob.obs_xy[keep]
    keep = prob.obs_pt != 7
    prob.obs_img, prob.obs_pt, prob.obs_xy = prob.obs_img[keep], prob.obs_pt[keep], prob.obs_xy[keep]
    lay = engine.build_layout(prob)
    assert lay.img_ptr[4] == lay.img_ptr[3]
    assert lay.pt_ptr[8] == lay.pt_ptr[7]
    assert np.all(lay.pm_depth[1] == 0)
    assert lay.flags == 0
    prob.img_free_mask[:] = 0
    lay = engine.build_layout(prob)
    assert lay.flags == engine.FLAG_STRUCTURE_ONLY and lay.num_blocks == prob.num_images
