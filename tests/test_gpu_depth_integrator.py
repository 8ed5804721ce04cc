"""``DepthIntegrator`` -- the mapper-facing half of the depth-integration path (``MpsfmMapper.integrate_bundle``,
mpsfm/sfm/mapper/base.py:619-631, over ``Integration.integrate`` / ``_prepare_integration_variables``,
mpsfm/sfm/scene/image/integration.py:92-137) -- on stub objects offering exactly what the reference's classes read:
the bundle's images go through ONE launch and end as the oracle's per-image solve does; the cached state makes the
second call a no-op; the propagated variances land in ``depth.uncertainty_update`` at the keypoints."""

import os
import sys
from types import SimpleNamespace as NS

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden_integration import make_inputs  # noqa: E402

from mpsfm_b200.integration import DepthIntegrator  # noqa: E402
from oracle import integration as oi  # noqa: E402

pytestmark = pytest.mark.gpu


class _Depth:
    def __init__(self, inp, n_kps):
        self.data = inp["depth_init"].copy()
        self.data_prior = inp["depth_prior"]
        self.uncertainty = inp["uncertainty"]
        self.valid = inp["valid"]
        self.activated = True
        self.uncertainty_update = np.zeros(n_kps)

    def data_prior_at_kps(self, kps):
        xy = np.round(np.asarray(kps)).astype(int)
        return self.data_prior[xy[:, 1], xy[:, 0]]


def _scene(seeds, H=40, W=56, n_sparse=60):
    images, proj, kps_all = {}, {}, {}
    for imid, seed in enumerate(seeds, start=1):
        inp = make_inputs(seed, H, W, n_sparse)
        K = inp["K"]  # [K11 sy, K00 sx, K12 sy, K02 sx] with sx = sy = 1
        cam = NS(sx=1.0, sy=1.0, calibration_matrix=lambda K=K: np.array([[K[1], 0, K[3]], [0, K[0], K[2]], [0, 0, 1.0]]))
        kps = np.asarray(inp["kps"], dtype=np.float64)  # integer pixels: (kps + 0.5).astype(int) gives them back
        images[imid] = NS(depth=_Depth(inp, len(kps)), normals=NS(data=inp["normals"], uncertainty=inp["normals_uncertainty"]),
                          camera=cam, conf=NS(robust_triangles=None, ignore_depths=True), _inp=inp)
        proj[imid] = (None, np.arange(len(kps)), kps, inp["depth3d"], True)
        kps_all[imid] = kps
    rec = NS(images=images,
             project_image_3d_points=lambda imid: proj[imid],
             find_points3D_with_small_triangulation_angle=lambda min_angle, point3D_ids: np.zeros(len(point3D_ids), bool),
             point_covs=NS(points_zvars=lambda image, ids: (None, image._inp["zvars3d"][np.asarray(ids)])),
             keypoints=lambda imid: kps_all[imid])
    return rec


def test_bundle_of_images_matches_the_per_image_oracle_and_caches_state():
    rec = _scene([21, 22, 23])
    integ = DepthIntegrator(rec, device=0)
    before = {i: im.depth.data.copy() for i, im in rec.images.items()}
    changed = integ.integrate_bundle([1, 2, 3], int_covs=True)
    assert changed == {1: True, 2: True, 3: True}
    conf = oi.IntegrationConf()
    for imid, image in rec.images.items():
        inp = image._inp
        pr = oi.prepare(inp["depth_prior"], inp["uncertainty"], inp["valid"], inp["normals"], inp["normals_uncertainty"],
                        before[imid], inp["kps"], inp["depth3d"], inp["zvars3d"], inp["K"], conf)
        ok, depth, _ = oi.integrate(pr, conf)
        assert ok
        assert np.max(np.abs(np.log(image.depth.data) - np.log(depth))) <= 1e-9 * np.max(np.abs(np.log(depth)))
        # variances at the keypoints: H^-1 1 of the oracle's Hessian at the integrated depth, times prior depth squared
        prh = oi.prepare(inp["depth_prior"], inp["uncertainty"], inp["valid"], inp["normals"], inp["normals_uncertainty"],
                         image.depth.data, inp["kps"], inp["depth3d"], inp["zvars3d"], inp["K"],
                         oi.IntegrationConf(scale_filter=False))
        diag, e_r, e_d = oi.hessian(prh, conf, ignore_depths=True)
        H, W = inp["depth_prior"].shape
        y = oi.variances_at(H, W, diag, e_r, e_d, np.asarray(inp["kps"], dtype=np.float64))
        want = y * image.depth.data_prior_at_kps(inp["kps"]) ** 2
        got = image.depth.uncertainty_update
        assert np.max(np.abs(got - want) / np.abs(want)) < 1e-5
    # second call: energies did not move -> nothing changes, no variance pass (first_refinement is over)
    snapshot = {i: im.depth.data.copy() for i, im in rec.images.items()}
    again = integ.integrate_bundle([1, 2, 3], int_covs=True)
    assert again == {1: False, 2: False, 3: False}
    assert all(np.array_equal(snapshot[i], rec.images[i].depth.data) for i in snapshot)
