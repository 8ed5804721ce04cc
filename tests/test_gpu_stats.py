"""mpba_segmented_median / mpba_depth_prior_terms against numpy (the reference computes both with numpy:
bundle_adjustment.py:10-15, 136-161, 239-241, 311-330)."""

import numpy as np
import pytest

from mpsfm_b200 import engine

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.asarray(a, dtype=np.float64).view(np.uint64)


def _np_median(x):
    with np.errstate(all="ignore"):
        import warnings

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return np.median(x) if x.size else np.nan


def _check(values, seg_ptr):
    med, mad = engine.segmented_median(values, seg_ptr, mad=True)
    for s in range(len(seg_ptr) - 1):
        x = values[seg_ptr[s]:seg_ptr[s + 1]]
        m = _np_median(x)
        if np.isnan(m):
            assert np.isnan(med[s]) and np.isnan(mad[s]), s
            continue
        assert med[s] == m, (s, med[s], m)
        assert _bits(med[s]) == _bits(m) or m == 0, (s, med[s], m)
        d = _np_median(np.abs(x - m))
        assert _bits(mad[s]) == _bits(d) or d == 0, (s, mad[s], d)


def test_segment_lengths_around_the_chunk_size():
    rng = np.random.default_rng(0)
    lens = [0, 1, 2, 3, 0, 4095, 4096, 4097, 8192, 10001, 0, 7, 100000, 1]
    seg_ptr = np.concatenate([[0], np.cumsum(lens)])
    values = rng.standard_normal(seg_ptr[-1]) * 3.0
    _check(values, seg_ptr)


def test_ties_signs_and_infinities():
    rng = np.random.default_rng(1)
    lens = [1000, 1001, 6, 5000, 64]
    seg_ptr = np.concatenate([[0], np.cumsum(lens)])
    values = np.round(rng.standard_normal(seg_ptr[-1]) * 2.0)  # many ties, +-0
    values[1000:1004] = [np.inf, -np.inf, np.inf, -np.inf]
    values[2001:2007] = [-1.0, -1.0, 2.0, 2.0, -3.0, 5.0]
    values[seg_ptr[4]:] = 1e-310 * rng.integers(-5, 5, 64)  # denormals
    _check(values, seg_ptr)


def test_nan_segment_is_nan_like_numpy():
    rng = np.random.default_rng(2)
    seg_ptr = np.array([0, 100, 300, 301])
    values = rng.standard_normal(301)
    values[150] = np.nan
    med, mad = engine.segmented_median(values, seg_ptr, mad=True)
    assert np.isnan(med[1]) and np.isnan(mad[1])
    assert med[0] == np.median(values[:100]) and med[2] == values[300]
    assert mad[2] == 0.0


def test_many_small_segments_and_one_large():
    rng = np.random.default_rng(3)
    lens = rng.integers(0, 9000, 700)
    seg_ptr = np.concatenate([[0], np.cumsum(lens)])
    values = rng.lognormal(size=seg_ptr[-1])
    med = engine.segmented_median(values, seg_ptr)
    ref = np.array([_np_median(values[a:b]) for a, b in zip(seg_ptr[:-1], seg_ptr[1:])])
    assert np.array_equal(_bits(med[~np.isnan(ref)]), _bits(ref[~np.isnan(ref)]))
    assert np.all(np.isnan(med[np.isnan(ref)]))
    big = rng.standard_normal(3_000_001)
    m, d = engine.segmented_median(big, [0, big.size], mad=True)
    assert m[0] == np.median(big) and d[0] == np.median(np.abs(big - np.median(big)))


def _terms_numpy(img_of, row3, xyz, d, v, f, gross, m):
    z = np.einsum("nk,nk->n", row3[img_of, :3], xyz) + row3[img_of, 3]
    with np.errstate(all="ignore"):
        mask = d > 0
        if f > 0:
            div = d / z
            mask = mask & (div < f) & (div > 1 / f)
        if gross:
            mask = mask & (np.abs(np.log(d).clip(1e-6, None) - np.log(z).clip(1e-6, None)) / v**0.5 < 3)
        w = d**2 * (1 / v.clip(1e-6, None))
        a = m * v**0.5 / d
        logratio = np.log((z / d).clip(1e-6, None))
        whitened = (np.log(d) - np.log(z)) / np.clip(v**0.5 / d, 1e-6, None)
    return dict(z=z, mask=mask, w=w, a=a, logratio=logratio, whitened=whitened)


@pytest.mark.parametrize("f,gross", [(0.0, False), (1.5, False), (1.5, True)])
def test_depth_prior_terms_match_numpy(f, gross):
    rng = np.random.default_rng(5)
    n, C = 200_003, 37
    img_of = rng.integers(0, C, n).astype(np.int32)
    row3 = np.concatenate([rng.standard_normal((C, 3)) * 0.3 + [0, 0, 1], rng.standard_normal((C, 1))], axis=1)
    xyz = rng.standard_normal((n, 3)) * 2 + [0, 0, 8]
    z = np.einsum("nk,nk->n", row3[img_of, :3], xyz) + row3[img_of, 3]
    d = np.abs(z) * np.exp(rng.standard_normal(n) * 0.2)
    d[::97] = 0.0  # invalid prior depth
    d[5::101] *= 3.0  # outside the scale filter
    v = (0.05 * np.maximum(d, 1e-3)) ** 2
    v[::53] = 1e-9  # below the variance clip
    got = engine.depth_prior_terms(img_of, row3, xyz, d, v, scale_filter_factor=f, gross_outliers=gross, loss_multiplier=2.5)
    ref = _terms_numpy(img_of, row3, xyz, d, v, f, gross, 2.5)
    ok = d > 0
    assert np.array_equal(_bits(got["w"]), _bits(ref["w"]))
    assert np.array_equal(_bits(got["a"][ok]), _bits(ref["a"][ok]))
    assert np.max(np.abs(got["z"] - ref["z"])) < 1e-13
    good = ok & (z > 1e-3)
    # thresholded quantities: identical except where a value sits within rounding of its threshold
    assert np.mean(got["mask"][good] != ref["mask"][good]) < 1e-4
    assert np.array_equal(got["mask"][~ok], ref["mask"][~ok])
    assert np.max(np.abs(got["logratio"][good] - ref["logratio"][good])) < 1e-12
    assert np.max(np.abs(got["whitened"][good] - ref["whitened"][good]) / (1 + np.abs(ref["whitened"][good]))) < 1e-11
