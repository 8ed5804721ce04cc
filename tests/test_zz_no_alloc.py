"""mpba_segmented_median_static -- the allocation-free variant of mpba_segmented_median (scratch in a file-scope
``static __device__`` array of libmpba.so; no counterpart in the reference)
-- through the C ABI against the existing entry point: BIT-EXACT (the tolerance of tests/test_gpu_stats.py:
both are exact selects), on two calls in a row that share the static scratch, the second one smaller and with a NaN
segment so that stale contents of the first would show.

Collected last on purpose (``test_zz_``)."""

import ctypes as C

import numpy as np
import pytest

from mpsfm_b200 import engine

pytestmark = pytest.mark.gpu

MAX_VALUES, MAX_SEGMENTS = 262144, 256  # MPBA_MEDIAN_STATIC_MAX_VALUES / _SEGMENTS of include/mpba.h


def _both(lib, torch, values, seg_ptr):
    """(median, mad) of the existing entry point (caller-owned workspace) and of the static one, same device inputs."""
    dev = torch.device("cuda", 0)
    n, S = int(values.size), int(len(seg_ptr) - 1)
    t_val = torch.from_numpy(np.ascontiguousarray(values, dtype=np.float64)).to(dev)
    t_seg = torch.from_numpy(np.ascontiguousarray(seg_ptr, dtype=np.int32)).to(dev)
    outs = [torch.full((S,), -7.0, dtype=torch.float64, device=dev) for _ in range(4)]
    nbytes = C.c_size_t()
    assert lib.mpba_segmented_median_workspace_bytes(n, S, C.byref(nbytes)) == 0
    ws = torch.empty(max(int(nbytes.value), 256), dtype=torch.uint8, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = t_val.data_ptr() if n else None
    torch.cuda.synchronize()
    assert lib.mpba_segmented_median(None, n, vp, S, t_seg.data_ptr(), ws.data_ptr(), ws.numel(), outs[0].data_ptr(),
                                     outs[1].data_ptr(), stream) == 0
    before = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
    rc = lib.mpba_segmented_median_static(None, n, vp, S, t_seg.data_ptr(), outs[2].data_ptr(), outs[3].data_ptr(), stream)
    assert rc == 0
    assert (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)) == before  # nothing went through torch
    return [o.cpu().numpy() for o in outs]


def _same(a, b):
    return np.array_equal(a, b, equal_nan=True)


def test_static_scratch_variant_equals_the_workspace_variant_on_two_calls_in_a_row():
    import torch

    lib = engine.load_library()
    rng = np.random.default_rng(7)
    # call 1: many ragged segments (an empty one, single elements, ties), close to both limits
    lens = rng.integers(0, 2000, size=200)
    lens[3], lens[10], lens[11] = 0, 1, 2
    values = rng.normal(0, 3, size=int(lens.sum()))
    values[::17] = np.round(values[::17])  # ties
    seg_ptr = np.concatenate([[0], np.cumsum(lens)])
    assert values.size <= MAX_VALUES and len(lens) <= MAX_SEGMENTS
    med_w, mad_w, med_s, mad_s = _both(lib, torch, values, seg_ptr)
    assert _same(med_s, med_w) and _same(mad_s, mad_w)
    want = np.array([np.median(values[a:b]) if b > a else np.nan for a, b in zip(seg_ptr[:-1], seg_ptr[1:])])
    assert _same(med_s, want)

    # call 2, straight after: fewer and shorter segments (the scratch still holds call 1's states, histograms and
    # deviations beyond what this call uses), one segment with a NaN, negative values and infinities
    lens2 = np.array([5, 0, 64, 1, 999, 4096, 7])
    values2 = rng.standard_cauchy(size=int(lens2.sum()))
    seg2 = np.concatenate([[0], np.cumsum(lens2)])
    values2[seg2[2] + 5] = np.nan
    values2[seg2[4] + 1], values2[seg2[4] + 2] = np.inf, -np.inf
    med_w2, mad_w2, med_s2, mad_s2 = _both(lib, torch, values2, seg2)
    assert _same(med_s2, med_w2) and _same(mad_s2, mad_w2)
    with np.errstate(invalid="ignore"):
        want2 = np.array([np.median(values2[a:b]) if b > a else np.nan for a, b in zip(seg2[:-1], seg2[1:])])
    assert _same(med_s2, want2) and np.isnan(med_s2[1]) and np.isnan(med_s2[2])

    # and call 1 again: same answer as the first time
    _, _, med_s3, mad_s3 = _both(lib, torch, values, seg_ptr)
    assert _same(med_s3, med_s) and _same(mad_s3, mad_s)


def test_sizes_beyond_the_static_scratch_are_rejected_with_a_message():
    import torch

    lib = engine.load_library()
    dev = torch.device("cuda", 0)
    handle = C.c_void_p()
    assert lib.mpba_create(C.byref(handle), 0, None) == 0
    try:
        t_val = torch.zeros(8, dtype=torch.float64, device=dev)
        t_seg = torch.zeros(MAX_SEGMENTS + 2, dtype=torch.int32, device=dev)
        out = torch.zeros(MAX_SEGMENTS + 1, dtype=torch.float64, device=dev)
        for n, S in ((MAX_VALUES + 1, 1), (8, MAX_SEGMENTS + 1)):
            rc = lib.mpba_segmented_median_static(handle, n, t_val.data_ptr(), S, t_seg.data_ptr(), out.data_ptr(), None, None)
            assert rc == -3  # MPBA_E_WORKSPACE
            msg = lib.mpba_last_error(handle).decode()
            assert "exceed the static scratch" in msg and str(MAX_VALUES) in msg and str(MAX_SEGMENTS) in msg
        # the limits themselves are accepted
        big = torch.arange(MAX_VALUES, dtype=torch.float64, device=dev)
        seg = torch.tensor([0, MAX_VALUES], dtype=torch.int32, device=dev)
        o = torch.zeros(2, dtype=torch.float64, device=dev)
        assert lib.mpba_segmented_median_static(handle, MAX_VALUES, big.data_ptr(), 1, seg.data_ptr(), o.data_ptr(),
                                                o[1:].data_ptr(), None) == 0
        assert o[0].item() == (MAX_VALUES - 1) / 2 and o[1].item() == MAX_VALUES / 4
    finally:
        lib.mpba_destroy(handle)
