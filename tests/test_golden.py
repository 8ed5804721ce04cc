"""Golden vectors recorded from the REFERENCE's own Python code (tests/golden/make_golden.py) against
mpsfm_b200's problem assembly and the oracle's sampling restatement.  No GPU, no /root/reference needed."""

import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from make_golden import CONF_OVERRIDES, SCENARIOS, make_bundle  # noqa: E402  (scenario table only; no stubs installed)
from scene_factory import build_scene  # noqa: E402

from mpsfm_b200.optimizer import Optimizer  # noqa: E402
from mpsfm_b200.problem import FREE_ALL, FREE_FIX_TX, FREE_NONE, LossType  # noqa: E402
from oracle import sampling  # noqa: E402

G = np.load(os.path.join(HERE, "golden", "assembly.npz"))
S = np.load(os.path.join(HERE, "golden", "sampling.npz"))


def test_sampling_restatement_matches_reference_grid_sample():
    got = sampling.data_at_kps(S["kps"], S["data"], float(S["sx"]), float(S["sy"]))
    assert np.max(np.abs(got - S["s_data"])) <= 1e-12 * np.max(np.abs(S["s_data"]))
    assert np.array_equal(sampling.valid_at_kps(S["kps"], S["valid"], float(S["sx"]), float(S["sy"])), S["s_valid"])
    # out-of-image keypoints sample zeros (padding_mode="zeros") and are invalid
    outside = (S["kps"][:, 0] < -40) | (S["kps"][:, 1] < -40)
    assert np.all(got[outside] == 0)


def test_projection_restatement_matches_reference():
    pts, depth = sampling.project3D(S["X"], S["H"], S["K"])
    assert np.allclose(pts, S["pts"], rtol=1e-14, atol=0) and np.allclose(depth, S["depth"], rtol=1e-14, atol=0)


@pytest.mark.parametrize("scenario", SCENARIOS, ids=[s[0] for s in SCENARIOS])
def test_assembly_matches_what_the_reference_hands_to_its_solver(scenario):
    name, scene_kw, optim_ids, call, kw = scenario
    rec = build_scene(**scene_kw)
    opt = Optimizer(CONF_OVERRIDES.get(name, {}), rec, None)
    if name.startswith("global"):
        opt.update_truncation_multiplier(rec.reg_image_ids())  # row a-12
        assert abs(opt.truncation_multiplier - float(G[f"{name}/truncation_multiplier"])) <= 1e-12
    bundle = make_bundle(rec, name, optim_ids)
    kwargs = dict(kw)
    if call == "refine_3d_points":
        asm = opt.assemble(bundle, fix_pose=True, depth_loss_name=opt.conf.ref3d_loss_name, **kwargs)
    else:
        asm = opt.assemble(bundle, fix_pose=False, **kwargs)
    pr = asm.problem

    # --- images in config order, explicit variable points, reprojection loss options (bit exact)
    assert asm.image_ids[: asm.num_optim_images] == G[f"{name}/config_images"].tolist()
    assert sorted(asm.variable_points) == G[f"{name}/variable_points"].tolist()
    assert pr.reproj_loss_weight == float(G[f"{name}/loss_magnitude"])
    assert pr.reproj_loss_scale == float(G[f"{name}/loss_scale"])
    assert LossType(pr.reproj_loss).name == str(G[f"{name}/loss_type"])

    # --- gauge: constant blocks and manifolds -> free masks
    for k, (cq, ct, tfix, qman) in enumerate(G[f"{name}/gauge"].tolist()):
        if cq and ct:
            expect = FREE_NONE
        elif tfix == 1 and qman:
            expect = FREE_FIX_TX
        else:
            assert qman and not cq and not ct and tfix == 0
            expect = FREE_ALL
        assert pr.img_free_mask[k] == expect
    assert np.all(pr.img_free_mask[asm.num_optim_images :] == FREE_NONE)  # images outside the config: constant pose
    assert np.all(pr.shift_scale == 0)

    # --- depth-prior rows: same images in the same order, identical ids, identical numbers
    assert list(asm.depth_rows) == G[f"{name}/depth_images"].tolist()
    for imid, row in asm.depth_rows.items():
        assert np.array_equal(row["p3Ds"], G[f"{name}/depth/{imid}/p3Ds"].astype(np.int64))  # indexing: bit exact
        for key in ("depths", "magnitudes", "params"):
            ref = G[f"{name}/depth/{imid}/{key}"]
            assert np.max(np.abs(row[key] - ref) / np.abs(ref)) <= 1e-12, (imid, key)
        assert LossType(pr.depth_loss).name.replace("_", "") == str(G[f"{name}/depth/{imid}/loss"]).replace("_", "")
    # the rows landed on the right observations of the flat problem
    n_rows = sum(len(r["depths"]) for r in asm.depth_rows.values())
    assert pr.num_depth == n_rows
    assert np.array_equal(asm.obs_point3D_id[pr.depth_obs], np.concatenate([r["p3Ds"] for r in asm.depth_rows.values()]))
    assert np.array_equal(asm.obs_point2D_idx[pr.depth_obs], np.concatenate([r["p2Ds"] for r in asm.depth_rows.values()]))
    assert np.array_equal(pr.depth_d, np.concatenate([r["depths"] for r in asm.depth_rows.values()]))

    # --- closed-form scale (row a-13)
    for tag, skw in (("plain", {}), ("scale", dict(allow_scale_filter=True)),
                     ("metric", dict(allow_scale_filter=True, allow_metric_scale_filter=True))):
        b2 = dict(bundle)
        if tag == "metric":
            b2["ref_id"] = optim_ids[0]
            for k, imid in enumerate(optim_ids):
                rec.images[imid].depth.scale = 1.0 + 0.02 * k
        ss, ok = opt.optimize_prior_shiftscale(b2, **skw)
        assert ok and sorted(ss) == G[f"{name}/shiftscale/{tag}/ids"].tolist()
        got = np.array([ss[i][1] for i in sorted(ss)])
        assert np.max(np.abs(got - G[f"{name}/shiftscale/{tag}/scale"])) <= 1e-12
        assert all(ss[i][0] == 0.0 for i in ss)
        for imid in optim_ids:
            rec.images[imid].depth.scale = 1


def test_colmap_point_rules():
    """Upstream COLMAP semantics restated in assemble(): a point whose track leaves the problem is constant;
    explicit variable points pull in constant-pose observations from images outside the config."""
    rec = build_scene(seed=3, num_images=9, num_points=260)
    opt = Optimizer({}, rec, None)
    optim_ids = [5, 4, 6, 3]
    ref = rec.images[5]
    pts = sorted(set(ref.point3D_ids(ref.get_observation_point2D_idxs())))
    asm = opt.assemble({"optim_ids": optim_ids, "pts3D": pts}, fix_pose=False, mode="local")
    pr = asm.problem
    inside = set(optim_ids)
    var = set(asm.variable_points)
    for j, pid in enumerate(asm.point_ids.tolist()):
        track = rec.points3D[pid].track
        n_obs = int(np.sum(pr.obs_pt == j))
        if pid in var:
            assert n_obs == track.length() and pr.pt_const[j] == 0
        else:
            n_in = sum(el.image_id in inside for el in track.elements)
            assert n_obs == n_in and pr.pt_const[j] == (1 if n_in < track.length() else 0)
    extra = asm.image_ids[asm.num_optim_images :]
    assert extra and not (set(extra) & inside)
    # global mode: nothing outside the config
    asm = opt.assemble({"optim_ids": optim_ids, "pts3D": pts}, fix_pose=False, mode="global")
    assert asm.image_ids == optim_ids and not asm.variable_points


def test_conf_struct_semantics():
    rec = build_scene(seed=0, num_images=4, num_points=30)
    with pytest.raises(Exception):
        Optimizer({"no_such_key": 1}, rec, None)
    opt = Optimizer({"rob_std": 3, "gpu": {"device": 0, "extra": 1}}, rec, None)
    assert opt.conf.rob_std == 3 and opt.conf.depth_loss_name == "cauchy" and opt.truncation_multiplier == 1
