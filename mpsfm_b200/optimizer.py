"""Drop-in for the reference's pyceres-backed bundle-adjustment ``Optimizer``.

Same constructor, ``default_conf`` keys, methods, return shapes and in-place update semantics as
``mpsfm.sfm.mapper.bundle_adjustment.Optimizer`` (/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:18-333);
the Ceres problem the reference assembles through pycolmap / pyceres is assembled here as arrays
(:class:`mpsfm_b200.problem.BundleProblem`) and solved by the sm_100a kernels behind ``include/mpba.h``.

What replaces what:

=====================================================  ==========================================================
reference (bundle_adjustment.py)                        here
=====================================================  ==========================================================
``pycolmap.BundleAdjustmentConfig`` + ``create_default_bundle_adjuster`` (:85-104)   ``assemble()`` rows "reprojection"
per-image depth selection + ``create_depth_bundle_adjuster`` (:124-176)             ``assemble()`` rows "depth prior"
``set_parameter_block_constant`` / manifolds (:111-122, :178-182)                   ``img_free_mask`` / ``pt_const``
``pyceres.solve`` SPARSE_SCHUR (:285-293)                                           ``engine.BundleAdjuster.solve``
in-place update through aliased parameter blocks                                    ``_write_back()``
``estimate_ba_covariance`` (:244-261)                                               ``engine.BundleAdjuster.point_covariances``
=====================================================  ==========================================================

There is no CPU path for the solve: without ``libmpba.so`` and a CUDA device ``ba`` / ``refine_3d_points`` /
``calculate_point_covs`` raise :class:`mpsfm_b200.engine.EngineUnavailable`.  ``assemble()`` itself is host
code (as it is in the reference) and runs anywhere, which is what the golden-vector tests use.
"""

from __future__ import annotations

import dataclasses

import numpy as np

from .conf import BaseClass
from .problem import FREE_ALL, FREE_FIX_TX, FREE_NONE, MAX_CAMERA_PARAMS, BundleProblem, CameraModel, LossType


def fit_robust_gaussian_mad(data):
    """bundle_adjustment.py:10-15"""
    mu = np.median(data)
    mad = np.median(np.abs(data - mu))
    return mu, 1.4826 * mad


def _model_id(model) -> int:
    """pycolmap.CameraModelId / its name / its integer value -> COLMAP model id."""
    if hasattr(model, "name"):
        model = model.name
    if isinstance(model, str):
        return int(CameraModel[model.split(".")[-1]])
    return int(model)


@dataclasses.dataclass
class Assembly:
    """The array form of one Ceres problem plus the maps back to the reconstruction."""

    problem: BundleProblem
    image_ids: list  # problem image index -> image id; the bundle's optim_ids first (same order), then images
    #                  outside the bundle that only contribute constant-pose residuals of variable points
    num_optim_images: int
    point_ids: np.ndarray  # problem point index -> point3D id
    obs_image_id: np.ndarray  # [N] image id of every reprojection residual block, in the order they are added
    obs_point2D_idx: np.ndarray  # [N]
    obs_point3D_id: np.ndarray  # [N]
    depth_rows: dict  # image id -> {"p2Ds","p3Ds","depths","magnitudes","params"}: the arguments the reference
    #                   hands to pycolmap.create_depth_bundle_adjuster (:163-176)
    kp_std: float
    variable_points: list  # explicit variable points (local mode, :88-91)


class Optimizer(BaseClass):
    """Optimizer class for Bundle Adjustment (B200 engine)."""

    default_conf = {
        # --- the reference's keys and defaults (bundle_adjustment.py:21-36)
        "depth_loss_name": "cauchy",
        "ref3d_loss_name": "trivial",
        "reproj_loss_name": "SOFT_L1",
        "reproj_loss_scale": 1.5,
        "scale_filter": True,
        "scale_filter_factor": 1.5,
        "metric_scale_filter": True,
        "rob_std": 2,
        "truncation_mode": "mad",  # [quantile, mad]
        "gross_outliers": False,
        "single_rescale": True,
        "min_truncation_mult": None,
        "verbose": 0,
        # --- new, engine-only knobs (nested dict: stays open under the struct freeze, existing YAMLs still load)
        "gpu": {"device": 0, "pcg_tolerance": 1e-10, "pcg_max_iterations": 2000, "max_num_iterations": 50,
                # sample the depth / validity maps of the whole bundle in ONE kernel launch (mpba_sample_maps) instead
                # of the reference's three torch.grid_sample calls per image (priorutils.py:49-62)
                "sample_on_device": False,
                # depth-prior masks / weights of the whole bundle in one launch (mpba_depth_prior_terms) and the
                # medians of optimize_prior_shiftscale / update_truncation_multiplier on the GPU
                # (mpba_segmented_median) instead of per-image numpy
                "stats_on_device": False,
                # under torchrun: shard every solve over the ranks' GPUs (mpba_solve_sharded)
                "sharded": False,
                # EXTENSION, off by default (nothing upstream does this): one
                # point-to-plane row per depth-prior observation of an image that has monocular normals
                # (mpsfm_b200/normal_prior.py).  rel_sigma: plane-distance std as a fraction of the prior depth;
                # loss_scale: scale of the depth loss on the whitened row
                "normal_prior": {"enabled": False, "rel_sigma": 0.02, "loss_scale": 2.0}},
    }

    def _sample_bundle_on_device(self, optim_ids, img_p2ds, depth_type):
        """{imid: (valid [n] bool, depth [n])} for every activated image of the bundle from one mpba_sample_maps
        launch, or None when the option is off.  Same values as image.depth.valid_at_kps / data_at_kps /
        data_prior_at_kps (bundle_adjustment.py:130-135): the kernel keeps torch.grid_sample's operation order."""
        g = self.conf["gpu"] if "gpu" in self.conf else {}
        if not bool(g.get("sample_on_device", False)):
            return None
        rec = self.mpsfm_rec
        ids, kps, vmaps, dmaps, scales = [], [], [], [], []
        for imid in optim_ids:
            image = rec.images[imid]
            if image.depth is None or not image.depth.activated or img_p2ds[imid].size == 0:
                continue
            ids.append(imid)
            kps.append(np.asarray(image.keypoint_coordinates(img_p2ds[imid]), dtype=np.float64).reshape(-1, 2))
            vmaps.append(np.asarray(image.depth.valid, dtype=np.float64))
            dmaps.append(np.asarray(image.depth.data if depth_type == "update" else image.depth.data_prior, dtype=np.float64))
            cam = image.depth.camera
            scales.append((float(cam.sx), float(cam.sy)))
        if not ids:
            return {}
        from . import engine

        out = engine.sample_maps(kps + kps, vmaps + dmaps, scales + scales, device=int(g.get("device", 0)))
        n = len(ids)
        return {imid: (out[i] == 1, out[n + i]) for i, imid in enumerate(ids)}

    def _stats_on_device(self) -> bool:
        g = self.conf["gpu"] if "gpu" in self.conf else {}
        return bool(g.get("stats_on_device", False))

    def _row3_xyz(self, rows_imid_p3ds):
        """(img_of, row3, xyz, counts) of the samples of several images: third row of cam_from_world.matrix() per
        image (what project3D reads for the depth, utils/geometry.py:13-19) and the world points."""
        rec = self.mpsfm_rec
        row3 = np.array([np.asarray(rec.images[imid].cam_from_world.matrix(), dtype=np.float64)[2] for imid, _ in rows_imid_p3ds]
                        ).reshape(-1, 4)
        counts = [int(len(p3Ds)) for _, p3Ds in rows_imid_p3ds]
        img_of = np.repeat(np.arange(len(counts), dtype=np.int32), counts)
        xyz = [np.asarray(rec.point3D_coordinates([int(p) for p in p3Ds]), dtype=np.float64).reshape(-1, 3)
               for _, p3Ds in rows_imid_p3ds if len(p3Ds)]
        xyz = np.concatenate(xyz) if xyz else np.zeros((0, 3))
        return img_of, row3, xyz, counts

    def _prior_terms(self, rows, scale_f, gross, m):
        """[(mask, w, a)] per image of ``rows`` = [(imid, p2Ds, p3Ds, depths, variances)] (:136-161)."""
        rec = self.mpsfm_rec
        if not rows:
            return []
        if self._stats_on_device():
            from . import engine

            img_of, row3, xyz, counts = self._row3_xyz([(r[0], r[2]) for r in rows])
            out = engine.depth_prior_terms(img_of, row3, xyz, np.concatenate([r[3] for r in rows]),
                                           np.concatenate([r[4] for r in rows]), scale_filter_factor=scale_f,
                                           gross_outliers=gross, loss_multiplier=m, want=("mask", "w", "a"),
                                           device=int(self.conf.gpu.get("device", 0)))
            cuts = np.cumsum(counts)[:-1]
            return list(zip(np.split(out["mask"], cuts), np.split(out["w"], cuts), np.split(out["a"], cuts)))
        terms = []
        for imid, p2Ds, p3Ds, depths, variances in rows:
            with np.errstate(divide="ignore", invalid="ignore"):
                mask = depths > 0
                if scale_f > 0 or gross:
                    depth3d = (np.asarray(rec.project_image_3d_points(imid, p3Ds)[3], dtype=np.float64)
                               if p2Ds.size else np.zeros(0))
                if scale_f > 0:
                    div = depths / depth3d
                    mask = mask & (div < scale_f) & (div > (1 / scale_f))
                if gross:
                    whitened = np.abs(np.log(depths).clip(1e-6, None) - np.log(depth3d).clip(1e-6, None)) / variances**0.5
                    mask = mask & (whitened < 3)
                inv_uncert = 1 / variances.clip(1e-6, None)
                params = m * variances**0.5 / depths  # :160 loss scale a
                magnitudes = depths**2 * inv_uncert  # :161 ScaledLoss weight w
            terms.append((mask, magnitudes, params))
        return terms

    def _init(self, mpsfm_rec, correspondences=None):
        self.mpsfm_rec = mpsfm_rec
        self.correspondences = correspondences
        self.truncation_multiplier = 1
        self.get_loss = {"trivial": LossType.TRIVIAL, "cauchy": LossType.CAUCHY, "softl1": LossType.SOFT_L1,
                         "huber": LossType.HUBER}
        self.last_summary = None
        self.last_assembly = None

    # ------------------------------------------------------------------------------------------
    # problem assembly (host, like the reference)
    # ------------------------------------------------------------------------------------------
    def assemble(self, bundle, fix_pose, fix_scale=True, mode=None, depth_loss_name=None, allow_scale_filter=False,
                 param_multiplier=1, depth_type="update", all_points_variable=False, reproj_loss=None,
                 with_depth=True, **kw) -> Assembly:
        """Array form of ``Optimizer.__build_problem`` (:67-185) up to, not including, the solve."""
        conf, rec = self.conf, self.mpsfm_rec
        optim_ids = list(bundle["optim_ids"])
        depth_loss_name = depth_loss_name or conf.depth_loss_name
        depth_loss = self.get_loss[depth_loss_name]
        in_config = set(optim_ids)

        # --- BundleAdjustmentConfig (:85-94)
        if all_points_variable:  # calculate_point_covs (:249-250)
            variable_points = list(bundle["pts3D"])
        elif mode == "local":
            variable_points = [p for p in bundle["pts3D"] if rec.points3D[p].track.length() < 15]
        else:
            variable_points = []
        kp_std = float(np.median([rec.images[imid].kp_std for imid in optim_ids]))

        # --- reprojection residual blocks of the images in the config (COLMAP AddImageToProblem)
        image_ids = list(optim_ids)
        image_index = {imid: k for k, imid in enumerate(image_ids)}
        o_img, o_p2d, o_p3d, o_xy, o_row = [], [], [], [], []
        img_obs_offset, img_p2ds = {}, {}
        n = 0
        for imid in optim_ids:
            image = rec.images[imid]
            q = image.cam_from_world.rotation.quat
            q /= np.linalg.norm(q)  # COLMAP normalises the quaternion in place before adding the image
            p2ds = np.asarray(image.get_observation_point2D_idxs(), dtype=np.int64)
            img_obs_offset[imid], img_p2ds[imid] = n, p2ds
            if p2ds.size == 0:
                continue
            o_img.append(np.full(p2ds.size, imid, dtype=np.int64))
            o_row.append(np.full(p2ds.size, image_index[imid], dtype=np.int32))
            o_p2d.append(p2ds)
            o_p3d.append(np.asarray(image.point3D_ids(p2ds), dtype=np.uint64).astype(np.int64))
            o_xy.append(np.asarray(image.keypoint_coordinates(p2ds), dtype=np.float64).reshape(-1, 2))
            n += p2ds.size
        cat = lambda parts, dt, shape: (np.concatenate(parts) if parts else np.zeros(shape, dtype=dt))  # noqa: E731
        obs_img_id = cat(o_img, np.int64, (0,))
        obs_row = cat(o_row, np.int32, (0,))  # row of the observation's image in the problem, built per image
        obs_p2d = cat(o_p2d, np.int64, (0,))
        obs_p3d = cat(o_p3d, np.int64, (0,))
        obs_xy = cat(o_xy, np.float64, (0, 2))

        # --- explicit variable points: constant-pose residuals from images outside the config
        #     (COLMAP AddPointToProblem), then ParameterizePoints: a point whose track is not completely
        #     inside the problem is constant
        point_ids, inverse, counts = np.unique(obs_p3d, return_inverse=True, return_counts=True)
        n_in_problem = dict(zip(point_ids.tolist(), counts.tolist()))
        x_img, x_p2d, x_p3d, x_xy = [], [], [], []
        for pid in variable_points:
            track = rec.points3D[pid].track
            if n_in_problem.get(pid, 0) == track.length():
                continue
            for el in track.elements:
                if el.image_id in in_config:
                    continue
                n_in_problem[pid] = n_in_problem.get(pid, 0) + 1
                if el.image_id not in image_index:
                    image_index[el.image_id] = len(image_ids)
                    image_ids.append(el.image_id)
                x_img.append(el.image_id)
                x_p2d.append(el.point2D_idx)
                x_p3d.append(pid)
                x_xy.append(np.asarray(rec.images[el.image_id].keypoint_coordinates([el.point2D_idx])).reshape(2))
        if x_img:
            obs_img_id = np.concatenate([obs_img_id, np.asarray(x_img, dtype=np.int64)])
            obs_row = np.concatenate([obs_row, np.asarray([image_index[i] for i in x_img], dtype=np.int32)])
            obs_p2d = np.concatenate([obs_p2d, np.asarray(x_p2d, dtype=np.int64)])
            obs_p3d = np.concatenate([obs_p3d, np.asarray(x_p3d, dtype=np.int64)])
            obs_xy = np.concatenate([obs_xy, np.asarray(x_xy, dtype=np.float64).reshape(-1, 2)])
            point_ids, inverse = np.unique(obs_p3d, return_inverse=True)
        pid_list = point_ids.tolist()
        points3D = rec.points3D
        track_len = np.fromiter((points3D[pid].track.length() for pid in pid_list), dtype=np.int64, count=len(pid_list))
        in_problem = np.fromiter((n_in_problem[pid] for pid in pid_list), dtype=np.int64, count=len(pid_list))
        pt_const = (track_len > in_problem).astype(np.uint8)

        # --- images: poses, gauge (:111-122), cameras (intrinsics constant, :92-94)
        C = len(image_ids)
        img_q, img_t = np.empty((C, 4)), np.empty((C, 3))
        free = np.zeros(C, dtype=np.uint8)
        cam_ids = []
        for k, imid in enumerate(image_ids):
            image = rec.images[imid]
            img_q[k] = image.cam_from_world.rotation.quat
            img_t[k] = image.cam_from_world.translation
            cam_ids.append(image.camera_id)
            if k >= len(optim_ids) or fix_pose or k == 0:
                free[k] = FREE_NONE
            elif k == 1:
                free[k] = FREE_FIX_TX
            else:
                free[k] = FREE_ALL
        uniq_cams = list(dict.fromkeys(cam_ids))
        cam_row = {c: r for r, c in enumerate(uniq_cams)}
        cam_model = np.array([_model_id(rec.rec.cameras[c].model) for c in uniq_cams], dtype=np.int32)
        cam_params = np.zeros((len(uniq_cams), MAX_CAMERA_PARAMS))
        for r, c in enumerate(uniq_cams):
            prm = np.asarray(rec.rec.cameras[c].params, dtype=np.float64)
            cam_params[r, : prm.size] = prm

        # --- depth-prior residual blocks (:124-176)
        depth_rows = {}
        d_obs, d_d, d_w, d_a = [], [], [], []
        np_conf = conf.gpu.get("normal_prior", None) or {}
        want_normals = bool(np_conf.get("enabled", False)) and with_depth
        n_obs, n_vec, n_depth = [], [], []
        pm = param_multiplier * self.truncation_multiplier  # :109
        sampled = self._sample_bundle_on_device(optim_ids, img_p2ds, depth_type) if with_depth else None
        if with_depth:
            # phase 1 (per image, host): which observations carry a valid prior sample (:126-135, :143)
            rows = []
            for imid in optim_ids:
                image = rec.images[imid]
                if image.depth is None or not image.depth.activated:
                    continue
                p2Ds = img_p2ds[imid]
                if p2Ds.size == 0:
                    continue
                if sampled is not None:
                    valid, depths = sampled[imid]
                    depths = depths[valid]
                else:
                    kps = np.asarray(image.keypoint_coordinates(p2Ds), dtype=np.float64).reshape(-1, 2)
                    valid = np.asarray(image.depth.valid_at_kps(kps), dtype=bool)
                    kps = kps[valid]
                    depths = np.asarray(image.depth.data_at_kps(kps) if depth_type == "update"
                                        else image.depth.data_prior_at_kps(kps), dtype=np.float64)
                p2Ds = p2Ds[valid]
                p3Ds = np.asarray(image.point3D_ids(p2Ds), dtype=np.uint64).astype(np.int64)
                uu = image.depth.uncertainty_update
                variances = (np.asarray(uu, dtype=np.float64)[p2Ds] if isinstance(uu, np.ndarray)
                             else np.array([uu[i] for i in p2Ds], dtype=np.float64))
                rows.append((imid, p2Ds, p3Ds, depths, variances))
            # phase 2: masks (:139-147), loss scale a and weight w (:156-161) -- one launch for the bundle
            #          (mpba_depth_prior_terms) or the reference's per-image numpy
            scale_f = float(conf.scale_filter_factor) if (allow_scale_filter and conf.scale_filter) else 0.0
            m = pm * conf.rob_std
            terms = self._prior_terms(rows, scale_f, bool(conf.gross_outliers), m)
            # phase 3 (per image): keep the selected observations
            for (imid, p2Ds, p3Ds, depths, variances), (mask, magnitudes, params) in zip(rows, terms):
                if np.sum(mask) == 0:
                    self.log("No valid points for depth regularizing", level=1)
                    continue
                depths, p2Ds, p3Ds = depths[mask], p2Ds[mask], p3Ds[mask]
                magnitudes, params = magnitudes[mask], params[mask]
                depth_rows[imid] = {"p2Ds": p2Ds, "p3Ds": p3Ds, "depths": depths, "magnitudes": magnitudes, "params": params}
                pos = np.searchsorted(img_p2ds[imid], p2Ds)
                if not np.array_equal(img_p2ds[imid][np.minimum(pos, img_p2ds[imid].size - 1)], p2Ds):
                    # get_observation_point2D_idxs() of this pycolmap build is not ascending: look the rows up
                    where = {int(v): k for k, v in enumerate(img_p2ds[imid])}
                    pos = np.array([where[int(v)] for v in p2Ds], dtype=np.int64)
                d_obs.append(img_obs_offset[imid] + pos)
                d_d.append(depths)
                d_w.append(magnitudes)
                d_a.append(params)
                normals = getattr(rec.images[imid], "normals", None) if want_normals else None
                if normals is not None and getattr(normals, "data", None) is not None:
                    # nearest pixel of the normal map at the keypoint (the reference's bilinear sampler,
                    # priorutils.py:49-62, is single-channel); rows with a degenerate normal are left out
                    nmap = np.asarray(normals.data, dtype=np.float64)
                    cam = rec.images[imid].depth.camera
                    kps = np.asarray(rec.images[imid].keypoint_coordinates(p2Ds), dtype=np.float64).reshape(-1, 2)
                    col = np.clip(np.rint(kps[:, 0] * cam.sx).astype(np.int64), 0, nmap.shape[1] - 1)
                    row = np.clip(np.rint(kps[:, 1] * cam.sy).astype(np.int64), 0, nmap.shape[0] - 1)
                    vec = nmap[row, col]
                    ok = np.isfinite(vec).all(axis=1) & (np.linalg.norm(vec, axis=1) > 1e-6)
                    n_obs.append((img_obs_offset[imid] + pos)[ok])
                    n_vec.append(vec[ok])
                    n_depth.append(depths[ok])

        problem = BundleProblem(
            img_q=img_q, img_t=img_t,
            img_camera=np.array([cam_row[c] for c in cam_ids], dtype=np.int32),
            img_free_mask=free, cam_model=cam_model, cam_params=cam_params,
            pt_xyz=rec.point3D_coordinates(pid_list) if point_ids.size else np.zeros((0, 3)),
            pt_const=pt_const,
            obs_img=obs_row,
            obs_pt=inverse.astype(np.int32), obs_xy=obs_xy,
            depth_obs=cat(d_obs, np.int64, (0,)).astype(np.int32), depth_d=cat(d_d, np.float64, (0,)),
            depth_w=cat(d_w, np.float64, (0,)), depth_a=cat(d_a, np.float64, (0,)),
            shift_scale=np.zeros((C, 2)),  # :83; both coordinates constant at every call site (:176-182)
            reproj_loss=LossType.parse(reproj_loss if reproj_loss is not None else conf.reproj_loss_name),
            reproj_loss_scale=float(conf.reproj_loss_scale * kp_std),  # :101
            reproj_loss_weight=float(1 / kp_std**2),  # :99
            depth_loss=depth_loss,
        )
        if not fix_scale:
            raise NotImplementedError("fix_scale=False is never used by the reference's call sites (:263-283)")
        if n_obs and sum(a.size for a in n_obs):
            from . import normal_prior

            sigma = float(np_conf.get("rel_sigma", 0.02)) * np.concatenate(n_depth)
            problem = normal_prior.with_normal_priors(problem, np.concatenate(n_obs), np.concatenate(n_vec),
                                                      1.0 / sigma**2, float(np_conf.get("loss_scale", 2.0)))
        return Assembly(problem=problem, image_ids=image_ids, num_optim_images=len(optim_ids), point_ids=point_ids,
                        obs_image_id=obs_img_id, obs_point2D_idx=obs_p2d, obs_point3D_id=obs_p3d,
                        depth_rows=depth_rows, kp_std=kp_std, variable_points=variable_points)

    # ------------------------------------------------------------------------------------------
    # solve + in-place update
    # ------------------------------------------------------------------------------------------
    def solve(self, assembly: Assembly):
        """``Optimizer.solve`` (:285-293): SPARSE_SCHUR with Ceres defaults -> LM + Schur + PCG on the GPU."""
        from . import engine

        g = self.conf.gpu
        shard = None
        if bool(g.get("sharded", False)):
            # one process per GPU (torchrun), every rank running the same mapper: the solve is sharded over the
            # ranks (mpba_solve_sharded) and every rank receives the full, bit-identical solution
            import torch.distributed as dist

            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                shard = (dist.get_rank(), dist.get_world_size())
        ba = engine.BundleAdjuster(assembly.problem, device=int(g["device"]), shard=shard)
        try:
            opt = engine.default_options(pcg_tolerance=float(g["pcg_tolerance"]),
                                         pcg_max_iterations=int(g["pcg_max_iterations"]),
                                         max_num_iterations=int(g["max_num_iterations"]),
                                         verbose=1 if self.conf.verbose > 3 else 0)
            summary = ba.solve(opt)
            q, t, X = ba.state()
        finally:
            ba.close()
        if summary.termination == "FAILURE" and summary.num_iterations == 0:
            # Ceres' FAILURE at the initial evaluation (e.g. a depth prior on a point with z <= 0): pyceres.solve
            # does not raise and the reference ignores the summary (:285-293, :266, :283) -- parameters stay put
            self.log(f"bundle adjustment skipped: {summary.message}", level=1)
        else:
            self._write_back(assembly, q, t, X)
        self.last_summary = summary
        self.log(summary.brief_report(), level=2)
        return summary

    def _write_back(self, assembly: Assembly, q, t, X):
        """Ceres parameter blocks alias pycolmap memory in the reference; here the non-constant blocks are
        copied back into the very same numpy buffers."""
        rec, pr = self.mpsfm_rec, assembly.problem
        for k in range(assembly.num_optim_images):
            mask = int(pr.img_free_mask[k])
            if mask == 0:
                continue
            pose = rec.images[assembly.image_ids[k]].cam_from_world
            if mask & 0b000111:
                pose.rotation.quat[:] = q[k]
            if mask & 0b111000:
                pose.translation[:] = t[k]
        # one pass over the variable points; the ids and rows as plain Python ints (numpy scalars cost ~3x in this loop)
        rows = np.flatnonzero(pr.pt_const == 0)
        points3D = rec.points3D
        for pid, xyz in zip(assembly.point_ids[rows].tolist(), X[rows]):
            points3D[pid].xyz[:] = xyz

    def __build_problem(self, bundle, fix_pose, fix_scale, **kwargs):
        assembly = self.assemble(bundle, fix_pose, fix_scale, **kwargs)
        self.last_assembly = assembly
        summary = self.solve(assembly)
        shift_scale = {imid: np.array([0.0, 0.0]) for imid in bundle["optim_ids"]}
        return summary, shift_scale

    # ------------------------------------------------------------------------------------------
    # the entry points the mapper calls (base.py:456-648)
    # ------------------------------------------------------------------------------------------
    def ba(self, bundle, mode, **kwargs):
        """Optimizes per frame data and 3d points in entire reconstruction (:263-266)."""
        problem, _ = self.__build_problem(bundle, fix_pose=False, fix_scale=True, mode=mode, **kwargs)
        return problem, True

    def refine_3d_points(self, bundle, **kwargs):
        """Refines triangulated 3d points with depth maps keeping poses fixed (:276-283)."""
        problem, _ = self.__build_problem(bundle, fix_pose=True, fix_scale=True,
                                          depth_loss_name=self.conf.ref3d_loss_name, **kwargs)
        return problem, True

    def calculate_point_covs(self, bundle):
        """3x3 covariance of every point of the bundle (:244-261): TRIVIAL loss with magnitude 1/kp_std^2,
        all points of the bundle variable, intrinsics constant; COLMAP's POINTS covariance is conditioned on
        the poses, i.e. the inverse of the point's own normal-equation block."""
        from . import engine

        assembly = self.assemble(bundle, fix_pose=False, all_points_variable=True, reproj_loss=LossType.TRIVIAL,
                                 with_depth=False)
        ba = engine.BundleAdjuster(assembly.problem, device=int(self.conf.gpu["device"]))
        try:
            covs = ba.point_covariances()
        finally:
            ba.close()
        index = {int(pid): j for j, pid in enumerate(assembly.point_ids)}
        for p3Did in bundle["pts3D"]:
            j = index.get(int(p3Did))
            if j is not None:
                self.mpsfm_rec.point_covs.data[p3Did] = covs[j]

    def filter_points3D(self, max_reproj_error, min_tri_angle, point3D_ids):
        """``mpsfm_rec.obs.filter_points3D(max_reproj_error, min_tri_angle, point3D_ids)`` as the mapper calls it after
        every bundle (mpsfm/sfm/mapper/base.py:700-711 ``filter_bundle``, :772-797 ``filter_local_points3D``), on the
        GPU: one ``mpba_filter_tracks`` pass over the tracks of ``point3D_ids``, then the same in-place edits pycolmap's
        ObservationManager makes (``delete_observation``, ``delete_point3D``, ``point3D.error``).  Returns the number
        of filtered observations."""
        from . import engine

        rec = self.mpsfm_rec
        pids = [int(p) for p in point3D_ids if int(p) in rec.points3D]
        if not pids:
            return 0
        image_index, image_ids = {}, []
        o_img, o_pt, o_el, o_xy = [], [], [], []
        for j, pid in enumerate(pids):
            for el in rec.points3D[pid].track.elements:
                if el.image_id not in image_index:
                    image_index[el.image_id] = len(image_ids)
                    image_ids.append(el.image_id)
                o_img.append(image_index[el.image_id])
                o_pt.append(j)
                o_el.append((el.image_id, el.point2D_idx))
                o_xy.append(np.asarray(rec.images[el.image_id].keypoint_coordinates([el.point2D_idx])).reshape(2))
        C = len(image_ids)
        img_q, img_t = np.empty((C, 4)), np.empty((C, 3))
        cam_ids = []
        for k, imid in enumerate(image_ids):
            image = rec.images[imid]
            img_q[k] = image.cam_from_world.rotation.quat
            img_t[k] = image.cam_from_world.translation
            cam_ids.append(image.camera_id)
        uniq_cams = list(dict.fromkeys(cam_ids))
        cam_row = {c: r for r, c in enumerate(uniq_cams)}
        cam_params = np.zeros((max(len(uniq_cams), 1), MAX_CAMERA_PARAMS))
        for r, c in enumerate(uniq_cams):
            prm = np.asarray(rec.rec.cameras[c].params, dtype=np.float64)
            cam_params[r, : prm.size] = prm
        none = np.zeros(0)
        problem = BundleProblem(
            img_q=img_q, img_t=img_t, img_camera=np.array([cam_row[c] for c in cam_ids], dtype=np.int32),
            img_free_mask=np.full(C, FREE_NONE, dtype=np.uint8),
            cam_model=np.array([_model_id(rec.rec.cameras[c].model) for c in uniq_cams] or [0], dtype=np.int32),
            cam_params=cam_params, pt_xyz=rec.point3D_coordinates(pids), pt_const=np.zeros(len(pids), dtype=np.uint8),
            obs_img=np.asarray(o_img, dtype=np.int32), obs_pt=np.asarray(o_pt, dtype=np.int32),
            obs_xy=np.asarray(o_xy, dtype=np.float64).reshape(-1, 2), depth_obs=none.astype(np.int32), depth_d=none,
            depth_w=none, depth_a=none, shift_scale=np.zeros((C, 2)), reproj_loss=LossType.TRIVIAL,
            depth_loss=LossType.TRIVIAL).normalized()
        ba = engine.BundleAdjuster(problem, device=int(self.conf.gpu["device"]))
        try:
            obs_keep, pt_keep, pt_error, num_filtered = ba.filter_tracks(max_reproj_error, min_tri_angle)
        finally:
            ba.close()
        obs = rec.obs
        for j, pid in enumerate(pids):
            if not pt_keep[j]:
                obs.delete_point3D(pid)
            else:
                rec.points3D[pid].error = float(pt_error[j])
        for k in np.flatnonzero(~obs_keep):
            if pt_keep[o_pt[k]]:
                obs.delete_observation(*o_el[k])
        return num_filtered

    def filter_all_points3D(self, max_reproj_error, min_tri_angle):
        """``mpsfm_rec.obs.filter_all_points3D`` (mpsfm/sfm/mapper/base.py:748-755): the same rule over every point."""
        return self.filter_points3D(max_reproj_error, min_tri_angle, set(self.mpsfm_rec.points3D))

    def find_invalid_depth_points(self, imids):
        """Per image, the 3D points whose keypoint falls on an invalid depth pixel (base.py:757-767)."""
        collect_risky_p3d = []
        for imid in imids:
            image = self.mpsfm_rec.images[imid]
            pt2D_ids = image.get_observation_point2D_idxs()
            kps_with3D = image.keypoint_coordinates(pt2D_ids)
            p3d_ids = np.array(image.point3D_ids(pt2D_ids))
            valid = np.asarray(image.depth.valid_at_kps(kps_with3D), dtype=bool).reshape(-1)
            collect_risky_p3d.append(set(int(p) for p in p3d_ids[~valid]))
        return collect_risky_p3d

    def filter_local_points3D(self, local_bundle, filter_max_reproj_error, filter_min_tri_angle):
        """``MpsfmMapper.filter_local_points3D`` (base.py:769-797) on the GPU filter: points whose keypoints have no
        valid depth in every image of the bundle are held to COLMAP's default 1.5 degrees, the bundle's points (and
        its constant points) to ``filter_min_tri_angle``.  Returns the number of filtered observations."""
        collect_risky_p3d = self.find_invalid_depth_points(local_bundle["optim_ids"])
        collect_risky_p3d = set.intersection(*collect_risky_p3d)
        pts3d = local_bundle["pts3D"]
        if "constpoints" in local_bundle:
            pts3d = pts3d.union(local_bundle["constpoints"])
        num_changed_observations = self.filter_points3D(filter_max_reproj_error, 1.5, collect_risky_p3d)  # default colmap
        num_changed_observations += self.filter_points3D(filter_max_reproj_error, filter_min_tri_angle, pts3d)
        return num_changed_observations

    def optimize_prior_shiftscale(self, bundle, **kwargs):
        """Closed-form per-image log-scale of the depth prior (:268-274, :187-242)."""
        shift_scale, success = self.__build_shiftscale_problem(bundle, **kwargs)
        if not success:
            return None, False
        shift_scale = {imid: (shift, np.exp(scale)) for imid, (shift, scale) in shift_scale.items()}
        return shift_scale, True

    def __yield_problem_parameters(self, optim_ids, proj_depths=False):
        for imid in optim_ids:  # :50-65
            image = self.mpsfm_rec.images[imid]
            pt2D_ids = image.get_observation_point2D_idxs()
            kps_with3D = image.keypoint_coordinates(pt2D_ids)
            kwargs = {"imid": imid, "image": image, "pt3D_ids": image.point3D_ids(pt2D_ids), "kps": kps_with3D}
            kwargs["obsdepths"] = image.depth.data_prior_at_kps(kps_with3D)
            kwargs["valid"] = image.depth.valid_at_kps(kps_with3D)
            if proj_depths:
                kwargs["projdepths"] = self.mpsfm_rec.project_image_3d_points(imid, kwargs["pt3D_ids"])[3]
            yield kwargs

    def __build_shiftscale_problem(self, bundle, allow_scale_filter=False, allow_metric_scale_filter=False):
        conf, rec = self.conf, self.mpsfm_rec
        shift_scale = {}
        pending = []  # gpu.stats_on_device: (imid, p3D ids, prior depths) of the images still to be reduced
        for kw in self.__yield_problem_parameters(bundle["optim_ids"], proj_depths=conf.scale_filter or conf.metric_scale_filter):
            pose = kw["image"].cam_from_world
            imid, p3dids = kw["imid"], kw["pt3D_ids"]
            if (conf.scale_filter_factor or conf.metric_scale_filter) and (
                "ref_id" in bundle and imid != bundle["ref_id"] and conf.single_rescale
            ):
                continue  # :204-207
            if allow_metric_scale_filter and conf.metric_scale_filter and (
                (imid == bundle["ref_id"]) or (not conf.single_rescale)
            ):  # :208-232
                scale = kw["projdepths"] / (kw["obsdepths"].clip(1e-6, None))
                proposed_scale = scale * rec.images[imid].depth.scale
                map_scale = np.mean([rec.images[i].depth.scale for i in bundle["optim_ids"] if i != imid])
                div = map_scale / proposed_scale
                valid = (div < 1.5) * (div > (1 / 1.5))
                presum = kw["valid"].sum()
                kw["valid"] = kw["valid"] * valid
                if kw["valid"].sum() == 0:
                    print("WARNING: Settin all points as outliers for metric scale optim and using map scale!!")
                    shift_scale[imid] = np.array([0.0, np.log(map_scale / rec.images[imid].depth.scale)])
                    break  # the reference returns here (:224-226); images reduced so far are kept
                self.log(f"Setting {presum - kw['valid'].sum()}points as outliers for metric scale optim, out of {presum}", level=3)
            if allow_scale_filter and conf.scale_filter and not allow_metric_scale_filter:  # :233-238
                div = kw["obsdepths"] / kw["projdepths"]
                kw["valid"] = kw["valid"] * ((div < conf.scale_filter_factor) * (div > (1 / conf.scale_filter_factor)))
            if self._stats_on_device():
                keep = np.asarray(kw["valid"], dtype=bool)
                pending.append((imid, np.asarray(p3dids)[keep], np.asarray(kw["obsdepths"], dtype=np.float64)[keep]))
                continue
            z = (pose * rec.point3D_coordinates(p3dids))[:, -1]
            z = z[kw["valid"]]
            odepth = kw["obsdepths"][kw["valid"]]
            shift_scale[imid] = np.array([0.0, np.median(np.log((z / odepth).clip(1e-6, None)))])  # :239-241
        if pending:
            # every image's log(z / d) in one launch, then one segmented median (one segment per image)
            from . import engine

            device = int(self.conf.gpu.get("device", 0))
            img_of, row3, xyz, counts = self._row3_xyz([(imid, p3Ds) for imid, p3Ds, _ in pending])
            odepth = np.concatenate([d for _, _, d in pending])
            terms = engine.depth_prior_terms(img_of, row3, xyz, odepth, np.ones_like(odepth), want=("logratio",),
                                             device=device, as_torch=True)
            med = engine.segmented_median(terms["logratio"], np.concatenate([[0], np.cumsum(counts)]), device=device)
            for (imid, _, _), value in zip(pending, med):
                shift_scale[imid] = np.array([0.0, float(value)])
        return shift_scale, True

    def update_truncation_multiplier(self, imids):
        """Robust spread (1.4826 MAD) of the whitened log-depth discrepancy over the given images (:295-333)."""
        D, D3d, dstds, P3, V = [], [], [], [], []
        on_device = self._stats_on_device()
        for imid in imids:
            image = self.mpsfm_rec.images[imid]
            p2Ds = np.asarray(image.get_observation_point2D_idxs(), dtype=np.int64)
            kps = np.asarray(image.keypoint_coordinates(p2Ds), dtype=np.float64).reshape(-1, 2)
            valid = np.asarray(image.depth.valid_at_kps(kps), dtype=bool)
            depths = np.asarray(image.depth.data_at_kps(kps[valid]), dtype=np.float64)
            p2Ds = p2Ds[valid]
            p3Ds = np.asarray(image.point3D_ids(p2Ds), dtype=np.uint64).astype(np.int64)
            mask = depths > 0
            P3.append(p3Ds[mask])
            depth3d = np.zeros(0) if on_device else self.mpsfm_rec.project_image_3d_points(imid, [int(p) for p in p3Ds[mask]])[3]
            uu = image.depth.uncertainty_update
            variances = (np.asarray(uu, dtype=np.float64)[p2Ds[mask]] if isinstance(uu, np.ndarray)
                         else np.array([uu[i] for i in p2Ds[mask]], dtype=np.float64))
            D.append(depths[mask])
            D3d.append(np.asarray(depth3d, dtype=np.float64))
            dstds.append(variances**0.5)
            V.append(variances)
        if on_device:
            # whitened log discrepancy of every sample in one launch, then median and MAD as one radix select each
            from . import engine

            device = int(self.conf.gpu.get("device", 0))
            img_of, row3, xyz, _ = self._row3_xyz(list(zip(imids, P3)))
            depths, dvars = np.concatenate(D), np.concatenate(V)
            terms = engine.depth_prior_terms(img_of, row3, xyz, depths, dvars, want=("whitened",), device=device,
                                             as_torch=True)
            _, mad = engine.segmented_median(terms["whitened"], [0, depths.size], mad=True, device=device)
            sigma = 1.4826 * float(mad[0])
        else:
            depths, depth3ds, dstds = np.concatenate(D), np.concatenate(D3d), np.concatenate(dstds)
            log_stds = np.clip(dstds / depths, 1e-6, None)
            _, sigma = fit_robust_gaussian_mad((np.log(depths) - np.log(depth3ds)) / log_stds)
        self.truncation_multiplier = sigma
        if self.conf.min_truncation_mult is not None:
            self.truncation_multiplier = max(self.truncation_multiplier, self.conf.min_truncation_mult)
