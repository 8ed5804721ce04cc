"""Host side of the sharded solve (mpba_shard): shard ranges, per-rank layouts, and the world_size-2
plumbing over gloo.  No GPU needed; the device side is covered by tests/test_gpu_sharded.py."""

import ctypes as C
import os
import socket

import numpy as np
import pytest

from mpsfm_b200 import engine
from mpsfm_b200.synthetic import make_scene


def _scene():
    return make_scene(11, 260, 4, seed=3, depth_frac=0.7, const_point_frac=0.15)


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shards_partition_both_observation_orders(world):
    prob = _scene()
    lay = engine.build_layout(prob)
    specs = engine.shard_ranges(lay, world)
    assert specs[0].img_begin == 0 and specs[-1].img_end == lay.C
    assert all(a.img_end == b.img_begin for a, b in zip(specs, specs[1:]))
    assert all(s.pt_begin == min(s.rank * s.pt_stride, lay.P) for s in specs)
    assert specs[-1].pt_end == lay.P and world * specs[0].pt_stride >= lay.P
    pm, cm, items = [], [], 0
    for r in range(world):
        sh, spec = engine.shard_layout(lay, r, world)
        s0, s1 = int(lay.pt_ptr[spec.pt_begin]), int(lay.pt_ptr[spec.pt_end])
        c0, c1 = int(lay.img_ptr[spec.img_begin]), int(lay.img_ptr[spec.img_end])
        assert sh.N >= max(s1 - s0, c1 - c0) and sh.C == lay.C and sh.P == lay.P
        # tracks: empty outside the owned range, identical (rebased) inside
        n = np.diff(sh.pt_ptr)
        assert np.all(n[: spec.pt_begin] == 0) and np.all(n[spec.pt_end:] == 0)
        assert np.array_equal(n[spec.pt_begin:spec.pt_end], np.diff(lay.pt_ptr)[spec.pt_begin:spec.pt_end])
        assert np.array_equal(sh.pm_img[: s1 - s0], lay.pm_img[s0:s1])
        assert np.array_equal(sh.pm_depth[:, : s1 - s0], lay.pm_depth[:, s0:s1])
        pm.append(sh.pm_src[: s1 - s0])
        # image order: chunks of the owned images only, inside the local slots
        ni = np.diff(sh.img_ptr)
        assert np.all(ni[: spec.img_begin] == 0) and np.all(ni[spec.img_end:] == 0)
        assert np.array_equal(sh.cm_pt[: c1 - c0], lay.cm_pt[c0:c1])
        assert np.all((sh.chunk_img >= spec.img_begin) & (sh.chunk_img < spec.img_end))
        assert np.all(sh.chunk_begin >= 0) and np.all(sh.chunk_end <= c1 - c0)
        assert int(np.sum(sh.chunk_end - sh.chunk_begin)) == c1 - c0
        assert sh.img_chunk_ptr[-1] == sh.chunk_img.shape[0]
        cm.append(sh.cm_src[: c1 - c0])
        # Schur items: pairs of local slots of one owned track, grouped by the (global) unit list
        assert sh.u_item_ptr.shape == lay.u_item_ptr.shape and sh.u_item_ptr[-1] == sh.item_a.shape[0]
        assert np.all((sh.item_a >= 0) & (sh.item_b < s1 - s0) & (sh.item_a <= sh.item_b))
        track_of = np.searchsorted(sh.pt_ptr, sh.item_a, side="right")
        assert np.array_equal(track_of, np.searchsorted(sh.pt_ptr, sh.item_b, side="right"))
        items += sh.item_a.shape[0]
        # the reduced-system structure is replicated
        assert sh.s_col is lay.s_col and sh.u_block is lay.u_block
    assert np.array_equal(np.concatenate(pm), lay.pm_src)
    assert np.array_equal(np.concatenate(cm), lay.cm_src)
    assert items == lay.item_a.shape[0]


def test_partial_schur_item_sums_add_up():
    """Summing a per-item quantity unit by unit over the shards reproduces the unsharded per-unit sums."""
    prob = _scene()
    lay = engine.build_layout(prob)
    rng = np.random.default_rng(0)
    val = rng.standard_normal(lay.N)  # one number per track-order slot
    U = lay.u_block.shape[0]

    def unit_sums(la, v):
        unit = np.repeat(np.arange(U), np.diff(la.u_item_ptr))
        return np.bincount(unit, weights=v[la.item_a] * v[la.item_b], minlength=U)

    full = unit_sums(lay, val)
    acc = np.zeros(U)
    for r in range(3):
        sh, spec = engine.shard_layout(lay, r, 3)
        s0, s1 = int(lay.pt_ptr[spec.pt_begin]), int(lay.pt_ptr[spec.pt_end])
        loc = np.zeros(sh.N)
        loc[: s1 - s0] = val[s0:s1]
        acc += unit_sums(sh, loc)
    assert np.allclose(acc, full, rtol=1e-12, atol=1e-12)


def test_shard_struct_matches_the_c_abi():
    assert C.sizeof(engine.MpbaShard) == 32


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, out):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        prob = _scene()
        lay = engine.build_layout(prob)
        sh, spec = engine.shard_layout(lay, rank, world)
        # the 128-byte communicator id travels from rank 0 exactly as BundleAdjuster._init_comm sends it
        ident = bytes(range(128)) if rank == 0 else bytes(128)
        got = engine.broadcast_bytes(ident, src=0)
        s0, s1 = int(lay.pt_ptr[spec.pt_begin]), int(lay.pt_ptr[spec.pt_end])
        c0, c1 = int(lay.img_ptr[spec.img_begin]), int(lay.img_ptr[spec.img_end])
        pieces = [None] * world
        dist.all_gather_object(pieces, (sh.pm_src[: s1 - s0], sh.cm_src[: c1 - c0], int(sh.item_a.shape[0])))
        ok = (got == bytes(range(128))
              and np.array_equal(np.concatenate([p[0] for p in pieces]), lay.pm_src)
              and np.array_equal(np.concatenate([p[1] for p in pieces]), lay.cm_src)
              and sum(p[2] for p in pieces) == lay.item_a.shape[0])
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_world_size_2_over_gloo():
    import torch.multiprocessing as mp

    world, port = 2, _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_gloo_worker, args=(world, port, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}


def test_sharded_entry_points_reject_missing_arguments():
    lib = engine.load_library()
    n = C.c_size_t()
    assert lib.mpba_workspace_bytes_sharded(None, None, C.byref(n)) == -1
    assert lib.mpba_solve_sharded(None, None, None, None, 0, None, None) == -1
    assert lib.mpba_comm_init(None, 2, 0, None, None) == -1
