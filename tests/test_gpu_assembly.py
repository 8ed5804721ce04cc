"""GPU problem assembly (SURVEY.md section 8 row f, next-2): batched keypoint sampling of the prior maps.

Pinned against the REFERENCE ITSELF: tests/golden/sampling.npz holds the outputs of the reference's own
``PriorUtils._data_at_kps`` (mpsfm/sfm/scene/image/mixins/priorutils.py:49-62, torch grid_sample), written by
tests/golden/make_golden.py."""

import os

import numpy as np
import pytest

from mpsfm_b200 import engine
from oracle import sampling

pytestmark = pytest.mark.gpu
S = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sampling.npz"))


def test_sampling_matches_the_reference_golden_vectors():
    sx, sy = float(S["sx"]), float(S["sy"])
    data, valid = engine.sample_maps([S["kps"], S["kps"]], [S["data"], S["valid"].astype(np.float64)], [(sx, sy)] * 2)
    assert np.max(np.abs(data - S["s_data"])) <= 1e-12 * np.max(np.abs(S["s_data"]))
    assert np.array_equal(valid == 1, S["s_valid"])  # valid_at_kps: all four neighbours valid, bit-exact
    outside = (S["kps"][:, 0] < -40) | (S["kps"][:, 1] < -40)
    assert np.all(data[outside] == 0)  # zero padding


def test_batched_sampling_is_bit_identical_to_the_oracle_restatement():
    rng = np.random.default_rng(3)
    kps, maps, scales = [], [], []
    for m in range(7):
        H, W = int(rng.integers(20, 90)), int(rng.integers(20, 120))
        sx, sy = W / 1600.0, H / 1200.0
        n = int(rng.integers(0, 4000)) if m != 3 else 0  # one image without keypoints (ragged batch)
        k = rng.uniform(-60, 1700, size=(n, 2))
        k[: n // 10] = np.round(k[: n // 10] * sx) / sx  # samples on exact pixel centres
        kps.append(k)
        maps.append(rng.uniform(0.5, 30.0, size=(H, W)) if m % 2 == 0 else (rng.uniform(size=(H, W)) > 0.2).astype(float))
        scales.append((sx, sy))
    got = engine.sample_maps(kps, maps, scales)
    for k, mp, sc, g in zip(kps, maps, scales, got):
        ref = sampling.data_at_kps(k, mp, *sc) if k.shape[0] else np.zeros(0)
        assert g.shape == ref.shape and np.array_equal(g, ref)


def test_optimizer_assembly_with_device_sampling_equals_the_reference_objects_path():
    """Optimizer.assemble with gpu.sample_on_device: the same depth-prior rows as through the scene objects' own
    valid_at_kps / data_at_kps (the reference's three grid_sample calls per image)."""
    from scene_factory import build_scene

    from mpsfm_b200.optimizer import Optimizer

    rec = build_scene(seed=4, num_images=7, num_points=240)
    bundle = {"optim_ids": [2, 1, 3, 4, 5, 6, 7], "pts3D": set(rec.points3D)}
    a = Optimizer({}, rec, None).assemble(bundle, fix_pose=False, mode="global", allow_scale_filter=True)
    b = Optimizer({"gpu": {"sample_on_device": True}}, rec, None).assemble(bundle, fix_pose=False, mode="global",
                                                                             allow_scale_filter=True)
    pa, pb = a.problem, b.problem
    assert pa.num_depth > 0
    assert np.array_equal(pa.depth_obs, pb.depth_obs)  # the same observations carry a prior (validity bit-exact)
    assert np.array_equal(pa.depth_d, pb.depth_d) and np.array_equal(pa.depth_w, pb.depth_w)
    assert np.array_equal(pa.depth_a, pb.depth_a)


def test_optimizer_stats_on_device_equal_the_numpy_path():
    """gpu.stats_on_device: depth-prior masks / weights (mpba_depth_prior_terms) and the medians of
    optimize_prior_shiftscale / update_truncation_multiplier (mpba_segmented_median) against the reference's
    per-image numpy on the same scene objects."""
    from scene_factory import build_scene

    from mpsfm_b200.optimizer import Optimizer

    rec = build_scene(seed=6, num_images=8, num_points=300)
    ids = [3, 1, 2, 4, 5, 6, 7, 8]
    bundle = {"optim_ids": ids, "pts3D": set(rec.points3D)}
    host, dev = Optimizer({}, rec, None), Optimizer({"gpu": {"stats_on_device": True, "sample_on_device": True}}, rec, None)
    for kw in ({}, {"allow_scale_filter": True}):
        pa = host.assemble(bundle, fix_pose=False, mode="global", **kw).problem
        pb = dev.assemble(bundle, fix_pose=False, mode="global", **kw).problem
        assert pa.num_depth > 0 and np.array_equal(pa.depth_obs, pb.depth_obs)
        assert np.array_equal(pa.depth_d, pb.depth_d)
        assert np.array_equal(pa.depth_w, pb.depth_w) and np.array_equal(pa.depth_a, pb.depth_a)  # IEEE order kept

    host.update_truncation_multiplier(ids)
    dev.update_truncation_multiplier(ids)
    assert host.truncation_multiplier > 0
    assert abs(host.truncation_multiplier - dev.truncation_multiplier) <= 1e-12 * host.truncation_multiplier

    for kw in ({}, {"allow_scale_filter": True}):
        for ref_id in ids[:3]:
            b = dict(bundle, ref_id=ref_id)
            sa, oka = host.optimize_prior_shiftscale(b, **kw)
            sb, okb = dev.optimize_prior_shiftscale(b, **kw)
            assert oka and okb and sa.keys() == sb.keys() and len(sa) > 0
            for imid in sa:
                assert sa[imid][0] == sb[imid][0] == 0.0
                assert abs(sa[imid][1] - sb[imid][1]) <= 1e-13 * abs(sa[imid][1])
