#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE's own Python code.  Runs only where /root/reference exists.

The reference's bundle-adjustment class (/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py) is imported
unmodified.  Its three native dependencies are not installable here, so they are replaced by RECORDING stubs
that capture every call the reference makes into them:

  pycolmap   BundleAdjustmentConfig / BundleAdjustmentOptions / create_default_bundle_adjuster /
             create_depth_bundle_adjuster / LossFunctionType
  pyceres    Problem.set_parameter_block_constant / set_manifold, SubsetManifold, EigenQuaternionManifold,
             SolverOptions, solve
  omegaconf  OmegaConf.create / merge / set_struct (mpsfm_b200.conf.Conf has the same semantics)

and the scene is the fake pycolmap-shaped reconstruction of tests/fake_scene.py with the reference's REAL
``PriorUtils._data_at_kps`` (torch grid_sample) and ``project3D_colmap`` plugged in.  What is recorded is
therefore exactly what the reference hands to its solver stack for a given scene: image order, gauge
(constant blocks / manifolds), variable points, loss options and, per image, the arrays given to the
depth-prior adjuster (p3Ds, depths, magnitudes, params) -- rows a-2, a-10, a-11, a-12, a-13 of SURVEY.md
section 8.  tests/test_golden.py rebuilds the same scenes on any box and compares mpsfm_b200's assembly.

    python tests/golden/make_golden.py        # rewrites tests/golden/assembly.npz and sampling.npz
"""

from __future__ import annotations

import enum
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from mpsfm_b200.conf import Conf, _to_conf  # noqa: E402


# ------------------------------------------------------------------------------------------------
# stubs
# ------------------------------------------------------------------------------------------------
def install_stubs(log):
    om = types.ModuleType("omegaconf")

    class DictConfig(Conf):
        pass

    def _dc(d):
        out = DictConfig()
        for k, v in dict(d).items():
            out[k] = _dc(v) if isinstance(v, dict) else v
        return out

    class OmegaConf:
        @staticmethod
        def create(d=None):
            return _dc(d or {})

        @staticmethod
        def set_struct(conf, flag):
            conf._frozen = flag

        @staticmethod
        def merge(a, b):
            out = _dc(a)
            for k, v in dict(b).items():
                if getattr(a, "_frozen", False) and k not in a:
                    raise AttributeError(k)
                if isinstance(out.get(k), dict) and isinstance(v, dict):
                    out[k].update(_dc(v))
                else:
                    out[k] = v
            return out

        @staticmethod
        def load(path):
            raise NotImplementedError

    om.OmegaConf, om.DictConfig = OmegaConf, DictConfig
    sys.modules["omegaconf"] = om

    pc = types.ModuleType("pycolmap")

    class LossFunctionType(enum.Enum):
        TRIVIAL = "TRIVIAL"
        SOFT_L1 = "SOFT_L1"
        CAUCHY = "CAUCHY"

    class BundleAdjustmentConfig:
        def __init__(self):
            self.images, self.variable_points, self.constant_cams = [], [], []

        def add_image(self, imid):
            self.images.append(imid)

        def add_variable_point(self, pid):
            self.variable_points.append(pid)

        def set_constant_cam_intrinsics(self, camid):
            self.constant_cams.append(camid)

    class BundleAdjustmentOptions:
        def __init__(self, **kw):
            self.kw = kw

    class Problem:
        def set_parameter_block_constant(self, arr):
            log["constant"].append(id(arr))

        def set_manifold(self, arr, manifold):
            log["manifold"].append((id(arr), manifold))

    class Bundler:
        def __init__(self):
            self.problem = Problem()

    def create_default_bundle_adjuster(options, config, rec):
        log["default"].append((options, config))
        # upstream COLMAP (BundleAdjuster::AddImageToProblem) normalises the quaternion of every image of the
        # config in place before adding its residuals; the stub keeps that side effect
        for imid in config.images:
            q = log["scene"].images[imid].cam_from_world.rotation.quat
            q /= np.linalg.norm(q)
        return Bundler()

    def create_depth_bundle_adjuster(problem, imid, p3Ds, depths, magnitudes, params, loss_type, shift_scale, rec,
                                     fix_shift, fix_scale, logloss):
        log["depth"].append(dict(imid=imid, p3Ds=np.array(p3Ds), depths=np.array(depths), magnitudes=np.array(magnitudes),
                                 params=np.array(params), loss=loss_type.name, shift_scale=np.array(shift_scale),
                                 fix_shift=fix_shift, fix_scale=fix_scale, logloss=logloss))

    pc.LossFunctionType, pc.BundleAdjustmentConfig, pc.BundleAdjustmentOptions = LossFunctionType, BundleAdjustmentConfig, BundleAdjustmentOptions
    pc.create_default_bundle_adjuster, pc.create_depth_bundle_adjuster = create_default_bundle_adjuster, create_depth_bundle_adjuster
    sys.modules["pycolmap"] = pc

    ce = types.ModuleType("pyceres")

    class SubsetManifold:
        def __init__(self, size, constant):
            self.size, self.constant = size, list(constant)

    class EigenQuaternionManifold:
        pass

    class SolverOptions:
        pass

    class LinearSolverType(enum.Enum):
        SPARSE_SCHUR = "SPARSE_SCHUR"

    class SolverSummary:
        def BriefReport(self):
            return "stub"

    def solve(options, problem, summary):
        log["solve"].append(dict(linear_solver_type=options.linear_solver_type.name, num_threads=options.num_threads))

    ce.Problem, ce.SubsetManifold, ce.EigenQuaternionManifold = Problem, SubsetManifold, EigenQuaternionManifold
    ce.SolverOptions, ce.LinearSolverType, ce.SolverSummary, ce.solve = SolverOptions, LinearSolverType, SolverSummary, solve
    sys.modules["pyceres"] = ce
    return SubsetManifold, EigenQuaternionManifold


def load_reference_modules():
    sys.path.insert(0, REF)

    def by_path(name, rel):
        spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod

    # mpsfm/sfm/mapper/__init__.py imports the whole mapper: load the single files instead
    import mpsfm.baseclass  # noqa: F401  (real file, with the omegaconf stub)

    ba = by_path("ref_bundle_adjustment", "mpsfm/sfm/mapper/bundle_adjustment.py")
    pu = by_path("ref_priorutils", "mpsfm/sfm/scene/image/mixins/priorutils.py")
    geo = by_path("ref_geometry", "mpsfm/utils/geometry.py")
    return ba, pu, geo


# the scenarios tests/test_golden.py replays: (name, scene kwargs, optim image ids, call, kwargs)
SCENARIOS = [
    ("global", dict(seed=1, num_images=7, num_points=220), [3, 1, 2, 4, 5, 6, 7], "ba",
     dict(mode="global", allow_scale_filter=True)),
    ("global_prior_final", dict(seed=2, num_images=6, num_points=150), [2, 4, 1, 3, 5, 6], "ba",
     dict(mode="global", allow_scale_filter=False, param_multiplier=0.125, depth_type="prior", final=True)),
    ("local", dict(seed=3, num_images=9, num_points=260), [5, 4, 6, 3], "ba",
     dict(mode="local", allow_scale_filter=True, refimid=5)),
    ("refine3d", dict(seed=4, num_images=6, num_points=160), [1, 2, 3], "refine_3d_points", dict(depth_type="prior")),
    ("global_gross", dict(seed=5, num_images=6, num_points=180), [1, 2, 3, 4, 5, 6], "ba",
     dict(mode="global", allow_scale_filter=True)),
]
CONF_OVERRIDES = {"global_gross": {"gross_outliers": True, "rob_std": 3, "reproj_loss_name": "CAUCHY", "depth_loss_name": "softl1"}}


def make_bundle(rec, name, optim_ids):
    bundle = {"optim_ids": list(optim_ids)}
    if name == "local":
        ref = rec.images[optim_ids[0]]
        bundle["pts3D"] = sorted(set(ref.point3D_ids(ref.get_observation_point2D_idxs())))
        bundle["ref_id"] = optim_ids[0]
    else:
        bundle["pts3D"] = sorted(rec.points3D.keys())
    return bundle


def main():
    log = {"constant": [], "manifold": [], "default": [], "depth": [], "solve": [], "scene": None}
    SubsetManifold, EigenQuaternionManifold = install_stubs(log)
    ba_mod, pu_mod, geo_mod = load_reference_modules()
    import fake_scene as fs
    from scene_factory import build_scene

    # plug the reference's own sampling / projection into the fake scene
    fs.Depth.sampler = staticmethod(lambda depth, kps, data: pu_mod.PriorUtils._data_at_kps(depth, kps, data))
    fs.Reconstruction.projector = staticmethod(lambda image, camera, pts: geo_mod.project3D_colmap(image, camera, pts))

    out = {}
    for name, scene_kw, optim_ids, call, kw in SCENARIOS:
        rec = build_scene(**scene_kw)
        for k in log:
            if k != "scene":
                log[k].clear()
        log["scene"] = rec
        opt = ba_mod.Optimizer(CONF_OVERRIDES.get(name, {}), rec, None)
        if name.startswith("global"):
            opt.update_truncation_multiplier(rec.reg_image_ids())
            out[f"{name}/truncation_multiplier"] = np.float64(opt.truncation_multiplier)
        ids = {}
        for imid, im in rec.images.items():
            ids[id(im.cam_from_world.rotation.quat)] = (imid, 0)
            ids[id(im.cam_from_world.translation)] = (imid, 1)
        bundle = make_bundle(rec, name, optim_ids)
        getattr(opt, call)(bundle, **kw)
        options, config = log["default"][0]
        out[f"{name}/config_images"] = np.array(config.images, dtype=np.int64)
        out[f"{name}/variable_points"] = np.array(sorted(config.variable_points), dtype=np.int64)
        out[f"{name}/loss_magnitude"] = np.float64(options.kw["loss_function_magnitude"])
        out[f"{name}/loss_scale"] = np.float64(options.kw["loss_function_scale"])
        out[f"{name}/loss_type"] = np.array(options.kw["loss_function_type"].name)
        # gauge: per image (constant q, constant t, fixed translation coordinates, quaternion manifold)
        gauge = np.zeros((len(optim_ids), 4), dtype=np.int64)
        row = {imid: k for k, imid in enumerate(optim_ids)}
        for a in log["constant"]:
            if a in ids:
                imid, kind = ids[a]
                gauge[row[imid], kind] = 1
        n_ss = 0
        for a, man in log["manifold"]:
            if a in ids:
                imid, kind = ids[a]
                if isinstance(man, SubsetManifold):
                    assert kind == 1 and man.size == 3
                    gauge[row[imid], 2] = sum(1 << c for c in man.constant)
                else:
                    assert kind == 0 and isinstance(man, EigenQuaternionManifold)
                    gauge[row[imid], 3] = 1
            else:  # shift_scale block: always SubsetManifold(2, [0, 1]) in the reference's call sites
                assert isinstance(man, SubsetManifold) and man.size == 2 and man.constant == [0, 1]
                n_ss += 1
        out[f"{name}/gauge"] = gauge
        out[f"{name}/num_shift_scale_manifolds"] = np.int64(n_ss)
        out[f"{name}/depth_images"] = np.array([d["imid"] for d in log["depth"]], dtype=np.int64)
        for d in log["depth"]:
            for key in ("p3Ds", "depths", "magnitudes", "params"):
                out[f"{name}/depth/{d['imid']}/{key}"] = d[key]
            out[f"{name}/depth/{d['imid']}/loss"] = np.array(d["loss"])
            assert d["fix_shift"] and d["fix_scale"] and d["logloss"] and np.all(d["shift_scale"] == 0)
        assert log["solve"] == [dict(linear_solver_type="SPARSE_SCHUR", num_threads=-1)]
        # closed-form scale (a-13) on the same scene
        for tag, skw in (("plain", {}), ("scale", dict(allow_scale_filter=True)),
                         ("metric", dict(allow_scale_filter=True, allow_metric_scale_filter=True))):
            b2 = dict(bundle)
            if tag == "metric":
                b2["ref_id"] = optim_ids[0]
                for k, imid in enumerate(optim_ids):
                    rec.images[imid].depth.scale = 1.0 + 0.02 * k
            ss, ok = opt.optimize_prior_shiftscale(b2, **skw)
            out[f"{name}/shiftscale/{tag}/ids"] = np.array(sorted(ss), dtype=np.int64)
            out[f"{name}/shiftscale/{tag}/scale"] = np.array([ss[i][1] for i in sorted(ss)], dtype=np.float64)
            for imid in optim_ids:
                rec.images[imid].depth.scale = 1
    np.savez_compressed(os.path.join(HERE, "assembly.npz"), **out)

    # sampling / projection known answers straight from the reference functions
    rng = np.random.Generator(np.random.PCG64(123))
    H, W = 37, 53
    data = rng.uniform(0.5, 9.0, (H, W))
    valid = rng.random((H, W)) > 0.2
    kps = np.concatenate([rng.uniform(-30, 1630, (400, 2)) * np.array([1.0, 0.75]),
                          np.array([[0.0, 0.0], [1599.0, 1199.0], [800.0, 600.0], [1599.0, 0.0]])])
    kps = kps.astype(np.float16).astype(np.float64)
    cam = types.SimpleNamespace(sx=(W - 1) / 1599.0, sy=(H - 1) / 1199.0)
    holder = types.SimpleNamespace(camera=cam)
    s_data = pu_mod.PriorUtils._data_at_kps(holder, kps.copy(), data)
    s_valid = pu_mod.PriorUtils._data_at_kps(holder, kps.copy(), valid) == 1
    X = rng.normal(0, 3, (50, 3)) + np.array([0, 0, 20.0])
    Hm = np.eye(4)
    Hm[:3, :3] = np.linalg.qr(rng.normal(size=(3, 3)))[0]
    Hm[:3, 3] = rng.normal(size=3)
    K = np.array([[1200.0, 0, 800.0], [0, 1190.0, 600.0], [0, 0, 1]])
    pts, depth = geo_mod.project3D(X, Hm, K)
    np.savez_compressed(os.path.join(HERE, "sampling.npz"), data=data, valid=valid, kps=kps, sx=cam.sx, sy=cam.sy,
                        s_data=s_data, s_valid=s_valid, X=X, H=Hm, K=K, pts=pts, depth=depth)
    print("wrote", os.path.join(HERE, "assembly.npz"), "and sampling.npz;", len(out), "arrays")


if __name__ == "__main__":
    main()
