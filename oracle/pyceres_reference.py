"""Opportunistic cross-check against the REAL solver stack of the reference.  TEST / BENCH INFRASTRUCTURE ONLY.

The reference's bundle adjustment is `pycolmap.create_default_bundle_adjuster` + `pyceres.solve`
(/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:85-122 problem, :285-293 solve).  Neither library is
installable in the build container and there is no network, so everything under oracle/ is "parity unpinned"
against Ceres itself.  This module is the one exit SURVEY.md section 8(c) names: WHEN a
box happens to carry `pycolmap` and `pyceres`, it builds the same problem out of real pycolmap objects with exactly
the reference's calls (file:line beside each), solves it with the reference's solver options, and hands back the
result for comparison.  `available()` is False everywhere else and callers skip.

It cannot run the reference's own file (`/root/reference` does not travel to the GPU box), so the call sequence is
restated here, reprojection blocks only unless the fork-only `create_depth_bundle_adjuster` is present.
Written against the pycolmap >= 3.11 / pyceres 2.x Python API from its documentation; UNTESTED in this container
(ModuleNotFoundError) -- a failure inside it is reported as a skip reason, never as a parity failure.
"""

from __future__ import annotations

import time

import numpy as np

_MODEL_NAMES = ["SIMPLE_PINHOLE", "PINHOLE", "SIMPLE_RADIAL", "RADIAL", "OPENCV", "OPENCV_FISHEYE", "FULL_OPENCV", "FOV",
                "SIMPLE_RADIAL_FISHEYE", "RADIAL_FISHEYE", "THIN_PRISM_FISHEYE"]
_NUM_PARAMS = [3, 4, 4, 5, 8, 8, 12, 5, 4, 5, 12]
_LOSS_NAMES = {0: "TRIVIAL", 1: "SOFT_L1", 2: "CAUCHY", 3: "HUBER"}


def available() -> bool:
    try:
        import pyceres  # noqa: F401
        import pycolmap  # noqa: F401
    except Exception:
        return False
    return hasattr(pycolmap, "create_default_bundle_adjuster")


def build_reconstruction(problem):
    """A pycolmap.Reconstruction holding the BundleProblem's images, cameras, points3D and tracks (ids = index + 1)."""
    import pycolmap

    rec = pycolmap.Reconstruction()
    for k in range(problem.cam_model.shape[0]):
        m = int(problem.cam_model[k])
        rec.add_camera(pycolmap.Camera(camera_id=k + 1, model=_MODEL_NAMES[m], width=1600, height=1200,
                                       params=np.asarray(problem.cam_params[k, : _NUM_PARAMS[m]], dtype=np.float64)))
    per_image = [[] for _ in range(problem.num_images)]
    for o in range(problem.num_obs):
        per_image[int(problem.obs_img[o])].append(o)
    slot = np.empty(problem.num_obs, dtype=np.int64)
    for i in range(problem.num_images):
        obs = per_image[i]
        slot[obs] = np.arange(len(obs))
        pts2d = pycolmap.ListPoint2D([pycolmap.Point2D(problem.obs_xy[o]) for o in obs])
        q = problem.img_q[i] / np.linalg.norm(problem.img_q[i])
        pose = pycolmap.Rigid3d(pycolmap.Rotation3d(q), np.asarray(problem.img_t[i], dtype=np.float64))
        im = pycolmap.Image(image_id=i + 1, name=f"{i}.jpg", camera_id=int(problem.img_camera[i]) + 1,
                            cam_from_world=pose, points2D=pts2d)
        rec.add_image(im)
        if hasattr(rec, "register_image"):
            rec.register_image(i + 1)
    by_point = [[] for _ in range(problem.num_points)]
    for o in range(problem.num_obs):
        by_point[int(problem.obs_pt[o])].append(o)
    point_ids = np.zeros(problem.num_points, dtype=np.int64)
    for p in range(problem.num_points):
        track = pycolmap.Track()
        for o in by_point[p]:
            track.add_element(int(problem.obs_img[o]) + 1, int(slot[o]))
        point_ids[p] = rec.add_point3D(np.asarray(problem.pt_xyz[p], dtype=np.float64), track, np.zeros(3, dtype=np.uint8))
    return rec, point_ids


def solve(problem, num_threads: int = -1):
    """The reference's call sequence on real objects.  Returns a dict (final_cost, initial_cost, iterations, q, t, X,
    seconds, depth_blocks).  Supports the reference's own gauge (first image constant, second image tx constant,
    :111-122) and constant points via set_parameter_block_constant."""
    import pyceres
    import pycolmap

    rec, point_ids = build_reconstruction(problem)
    # :85-94
    ba_config = pycolmap.BundleAdjustmentConfig()
    for i in range(problem.num_images):
        ba_config.add_image(i + 1)
    for k in range(problem.cam_model.shape[0]):
        ba_config.set_constant_cam_intrinsics(k + 1)
    # :98-102 (loss_function_magnitude exists only in the fork; stock pycolmap has weight 1)
    kw = dict(loss_function_type=pycolmap.LossFunctionType(_LOSS_NAMES[int(problem.reproj_loss)]),
              loss_function_scale=float(problem.reproj_loss_scale))
    try:
        options = pycolmap.BundleAdjustmentOptions(loss_function_magnitude=float(problem.reproj_loss_weight), **kw)
    except (TypeError, AttributeError):
        if float(problem.reproj_loss_weight) != 1.0:
            raise RuntimeError("stock pycolmap has no loss_function_magnitude: needs reproj_loss_weight == 1")
        options = pycolmap.BundleAdjustmentOptions(**kw)
    bundler = pycolmap.create_default_bundle_adjuster(options, ba_config, rec)  # :104
    prob = bundler.problem
    # :111-122
    for i in range(problem.num_images):
        pose = rec.images[i + 1].cam_from_world
        mask = int(problem.img_free_mask[i])
        if mask == 0:
            prob.set_parameter_block_constant(pose.rotation.quat)
            prob.set_parameter_block_constant(pose.translation)
            continue
        fixed_t = [j for j in range(3) if not (mask >> (3 + j)) & 1]
        if fixed_t:
            prob.set_manifold(pose.translation, pyceres.SubsetManifold(3, fixed_t))
        prob.set_manifold(pose.rotation.quat, pyceres.EigenQuaternionManifold())
    for p in np.flatnonzero(np.asarray(problem.pt_const) != 0):
        prob.set_parameter_block_constant(rec.points3D[int(point_ids[p])].xyz)
    depth_blocks = 0
    if problem.depth_obs.shape[0] and hasattr(pycolmap, "create_depth_bundle_adjuster"):
        # :163-182 (fork only): one call per image with its p3Ds, depths, magnitudes (w), params (a)
        di = problem.depth_obs
        shift_scale = {}
        for i in range(problem.num_images):
            sel = np.flatnonzero(problem.obs_img[di] == i)
            if sel.size == 0:
                continue
            shift_scale[i] = np.array([0.0, 0.0])
            pycolmap.create_depth_bundle_adjuster(
                prob, i + 1, [int(point_ids[problem.obs_pt[di[s]]]) for s in sel], problem.depth_d[sel],
                problem.depth_w[sel], problem.depth_a[sel],
                pycolmap.LossFunctionType(_LOSS_NAMES[int(problem.depth_loss)]), shift_scale[i], rec, True, True, True)
            prob.set_manifold(shift_scale[i], pyceres.SubsetManifold(2, [0, 1]))
            depth_blocks += int(sel.size)
    # :285-293
    so = pyceres.SolverOptions()
    so.linear_solver_type = pyceres.LinearSolverType.SPARSE_SCHUR
    so.minimizer_progress_to_stdout = False
    so.num_threads = num_threads
    summary = pyceres.SolverSummary()
    t0 = time.perf_counter()
    pyceres.solve(so, prob, summary)
    seconds = time.perf_counter() - t0
    q = np.stack([np.asarray(rec.images[i + 1].cam_from_world.rotation.quat) for i in range(problem.num_images)])
    t = np.stack([np.asarray(rec.images[i + 1].cam_from_world.translation) for i in range(problem.num_images)])
    X = np.stack([np.asarray(rec.points3D[int(pid)].xyz) for pid in point_ids]) if problem.num_points else np.zeros((0, 3))
    return dict(final_cost=float(summary.final_cost), initial_cost=float(summary.initial_cost),
                iterations=int(summary.num_successful_steps + summary.num_unsuccessful_steps), q=q, t=t, X=X,
                seconds=seconds, depth_blocks=depth_blocks, report=summary.BriefReport())
