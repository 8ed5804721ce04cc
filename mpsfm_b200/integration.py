"""Depth-map integration with normal priors on the GPU: host side of ``mpba_integrate_depth`` (include/mpba.h).

Mirrors the reference's interface for this step (the one right before every bundle adjustment):

    reference                                                           here
    ------------------------------------------------------------------  ------------------------------------------
    MpsfmMapper.integrate_bundle(imids, int_covs)   base.py:619-631      DepthIntegrator.integrate_bundle(imids, ...)
    Image.integrate() -> Integration._integrate     integration.py:133,  one mpba_integrate_depth launch for ALL images
                                                    :383-520             of the bundle (one thread-block cluster each)
    _prepare_integration_variables                  :92-131              DepthIntegrator._prepare (same filters)
    IntVars state (integrated, energy_old, wu, wv)  :18-29               IntegrationState per image, kept by the class
    calculate_hessian + calculate_int_covs_at_kps   :522-616             mpba_integration_variances (y = H^-1 1)

``integrate_maps`` / ``propagate_variances`` are the array-level entry points (what the parity tests call); the class on
top walks the reference's objects (``mpsfm_rec.images[imid].depth / .normals / .camera``).  There is no CPU fallback.
"""

from __future__ import annotations

import ctypes as C
import dataclasses
from typing import Optional, Sequence

import numpy as np

from . import engine

MPBA_INT_MAX_IRLS = 32


class MpbaIntegrationConf(C.Structure):
    _fields_ = [
        ("max_iter", C.c_int32), ("cg_max_iter", C.c_int32), ("tol", C.c_double), ("step_size", C.c_double),
        ("cg_tol", C.c_double), ("lambda1", C.c_double), ("lambda2", C.c_double), ("k", C.c_double),
        ("large_number", C.c_double), ("depth_magnitude_multiplier", C.c_double),
        ("normals_magnitude_multiplier", C.c_double), ("scale_filter_factor", C.c_double),
    ]


class MpbaIntegrationJob(C.Structure):
    _fields_ = [
        ("H", C.c_int32), ("W", C.c_int32), ("num_sparse", C.c_int32), ("integrated", C.c_int32),
        ("K", C.c_double * 4),
        ("depth_prior", C.c_void_p), ("uncertainty", C.c_void_p), ("valid", C.c_void_p), ("normals", C.c_void_p),
        ("normals_uncertainty", C.c_void_p), ("sparse_xy", C.c_void_p), ("sparse_depth", C.c_void_p),
        ("sparse_zvar", C.c_void_p), ("depth", C.c_void_p), ("wu", C.c_void_p), ("wv", C.c_void_p),
        ("energy_old", C.c_double), ("variance", C.c_void_p),
    ]


class MpbaIntegrationResult(C.Structure):
    _fields_ = [
        ("updated", C.c_int32), ("integrated", C.c_int32), ("irls_steps", C.c_int32), ("cg_iterations", C.c_int32),
        ("energy_old", C.c_double), ("energy", C.c_double * (MPBA_INT_MAX_IRLS + 1)),
        ("cg_steps", C.c_int32 * MPBA_INT_MAX_IRLS),
    ]


INTEGRATION_EXPORTS = ("mpba_integration_default_conf", "mpba_integration_workspace_bytes", "mpba_integrate_depth",
                       "mpba_integration_variances")


def _lib():
    lib = engine.load_library()
    if not getattr(lib, "_integration_declared", False):
        lib.mpba_integration_default_conf.argtypes = [C.POINTER(MpbaIntegrationConf)]
        lib.mpba_integration_default_conf.restype = None
        lib.mpba_integration_workspace_bytes.argtypes = [C.c_int32, C.POINTER(MpbaIntegrationJob), C.POINTER(C.c_size_t)]
        lib.mpba_integrate_depth.argtypes = [C.c_void_p, C.c_int32, C.POINTER(MpbaIntegrationJob),
                                             C.POINTER(MpbaIntegrationConf), C.c_void_p, C.c_size_t,
                                             C.POINTER(MpbaIntegrationResult)]
        lib.mpba_integration_variances.argtypes = [C.c_void_p, C.c_int32, C.POINTER(MpbaIntegrationJob),
                                                   C.POINTER(MpbaIntegrationConf), C.c_int32, C.c_double, C.c_void_p,
                                                   C.c_size_t, C.POINTER(MpbaIntegrationResult)]
        lib._integration_declared = True
    return lib


def default_conf(**overrides) -> MpbaIntegrationConf:
    """Image.default_conf's integration keys (mpsfm/sfm/scene/image/base.py:36-53).  ``scale_filter=False`` maps to
    scale_filter_factor = 0."""
    conf = MpbaIntegrationConf()
    _lib().mpba_integration_default_conf(C.byref(conf))
    if overrides.pop("scale_filter", True) is False:
        conf.scale_filter_factor = 0.0
    for k, v in overrides.items():
        if not hasattr(conf, k):
            raise AttributeError(f"mpba_integration_conf has no field {k!r}")
        setattr(conf, k, v)
    return conf


@dataclasses.dataclass
class IntegrationState:
    """IntVars (integration.py:18-29): what the reference keeps on an Image between integrations."""

    integrated: bool = False
    energy_old: float = 0.0
    wu: Optional[np.ndarray] = None
    wv: Optional[np.ndarray] = None


@dataclasses.dataclass
class MapInputs:
    """One image's inputs as arrays (what ``_integrate`` reads off the Image, :386-389)."""

    depth_prior: np.ndarray          # [H,W]  depth.data_prior
    uncertainty: np.ndarray          # [H,W]  depth.uncertainty
    valid: np.ndarray                # [H,W]  depth.valid
    normals: np.ndarray              # [H,W,3]
    normals_uncertainty: np.ndarray  # [H,W,3,3]
    depth: np.ndarray                # [H,W]  depth.data (start)
    K: Sequence[float]               # [K11 sy, K00 sx, K12 sy, K02 sx]
    kps: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros((0, 2), dtype=np.int64))  # integer pixels (x, y)
    depth3d: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(0))
    zvars3d: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(0))


@dataclasses.dataclass
class IntegrationOutcome:
    updated: bool
    depth: Optional[np.ndarray]   # refined depth when updated
    state: IntegrationState
    energies: list
    cg_iterations: list


class _Batch:
    """Device buffers + job structs of a batch of images; keeps the torch tensors alive for the call."""

    def __init__(self, inputs: Sequence[MapInputs], states: Sequence[Optional[IntegrationState]], device: int,
                 variances: bool):
        import torch

        if not torch.cuda.is_available():
            raise engine.EngineUnavailable("no CUDA device: mpsfm_b200 has no CPU fallback")
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.keep = []
        self.jobs = (MpbaIntegrationJob * max(len(inputs), 1))()
        self.out = []
        for k, (m, st) in enumerate(zip(inputs, states)):
            H, W = m.depth_prior.shape
            n = H * W
            j = self.jobs[k]
            j.H, j.W = H, W
            kps = np.asarray(m.kps).reshape(-1, 2)
            j.num_sparse = int(kps.shape[0])
            for q in range(4):
                j.K[q] = float(m.K[q])

            def put(a, dt):
                t = torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(self.dev)
                self.keep.append(t)
                return t

            j.depth_prior = put(m.depth_prior, np.float64).data_ptr()
            j.uncertainty = put(m.uncertainty, np.float64).data_ptr()
            j.valid = put(np.asarray(m.valid).astype(np.uint8), np.uint8).data_ptr()
            j.normals = put(m.normals, np.float64).data_ptr()
            j.normals_uncertainty = put(m.normals_uncertainty, np.float64).data_ptr()
            if j.num_sparse:
                j.sparse_xy = put(kps, np.int32).data_ptr()
                j.sparse_depth = put(m.depth3d, np.float64).data_ptr()
                j.sparse_zvar = put(m.zvars3d, np.float64).data_ptr()
            depth = put(m.depth, np.float64)
            j.depth = depth.data_ptr()
            st = st or IntegrationState()
            j.integrated = 1 if st.integrated else 0
            j.energy_old = float(st.energy_old)
            wu = put(st.wu if st.integrated else np.zeros(n), np.float64)
            wv = put(st.wv if st.integrated else np.zeros(n), np.float64)
            j.wu, j.wv = wu.data_ptr(), wv.data_ptr()
            var = None
            if variances:
                var = torch.empty(n, dtype=torch.float64, device=self.dev)
                self.keep.append(var)
                j.variance = var.data_ptr()
            self.out.append((depth, wu, wv, var, (H, W), st))
        self.num = len(inputs)
        nbytes = C.c_size_t()
        rc = _lib().mpba_integration_workspace_bytes(self.num, self.jobs, C.byref(nbytes))
        if rc != 0:
            raise engine.EngineError(f"mpba_integration_workspace_bytes failed with code {rc}")
        self.workspace = torch.empty(max(int(nbytes.value), 256), dtype=torch.uint8, device=self.dev)
        self.results = (MpbaIntegrationResult * max(self.num, 1))()


def _with_handle(device, fn):
    lib = _lib()
    free = engine._FREE_HANDLES.setdefault(int(device), [])
    handle = free.pop() if free else C.c_void_p()
    if not handle:
        rc = lib.mpba_create(C.byref(handle), device, None)
        if rc != 0:
            raise engine.EngineUnavailable(f"mpba_create failed with code {rc} (no usable CUDA device {device})")
    try:
        rc = fn(lib, handle)
        if rc != 0:
            msg = lib.mpba_last_error(handle)
            raise engine.EngineError(f"libmpba error {rc}: {msg.decode() if msg else ''}")
    finally:
        if len(free) < 4:
            free.append(handle)
        else:
            lib.mpba_destroy(handle)


def integrate_maps(inputs: Sequence[MapInputs], conf: Optional[MpbaIntegrationConf] = None,
                   states: Optional[Sequence[Optional[IntegrationState]]] = None, device: int = 0):
    """``Integration._integrate`` for a batch of images in one launch.  Returns a list of IntegrationOutcome."""
    conf = conf or default_conf()
    states = list(states) if states is not None else [None] * len(inputs)
    if not inputs:
        return []
    b = _Batch(inputs, states, device, variances=False)
    _with_handle(device, lambda lib, h: lib.mpba_integrate_depth(h, b.num, b.jobs, C.byref(conf), b.workspace.data_ptr(),
                                                                 b.workspace.numel(), b.results))
    outs = []
    for k in range(b.num):
        r = b.results[k]
        depth, wu, wv, _, (H, W), st_in = b.out[k]
        skipped = bool(st_in.integrated) and r.irls_steps == 0 and not r.updated
        if skipped:
            state = IntegrationState(True, float(r.energy_old), st_in.wu, st_in.wv)
        else:
            state = IntegrationState(bool(r.integrated), float(r.energy_old), wu.cpu().numpy(), wv.cpu().numpy())
        outs.append(IntegrationOutcome(
            updated=bool(r.updated), depth=depth.cpu().numpy().reshape(H, W) if r.updated else None, state=state,
            energies=[r.energy[i] for i in range(r.irls_steps + 1)], cg_iterations=[r.cg_steps[i] for i in range(r.irls_steps)]))
    return outs


def propagate_variances(inputs: Sequence[MapInputs], conf: Optional[MpbaIntegrationConf] = None, ignore_depths: bool = True,
                        tol: float = 1e-10, max_iter: int = 200000, device: int = 0):
    """``calculate_hessian`` + ``IntegrationUncertainty.solve`` (:522-574, :51-79) for a batch: per image the vector
    y = H^-1 1 over the grid ([H,W]); the propagated log-depth variance at a pixel (x, y) is y[round(y), round(x)]
    (the reference's one-hot solves followed by ``x.sum(0)`` compute exactly these entries, H being symmetric).
    Returns (list of [H,W] arrays, list of PCG iteration counts)."""
    conf = conf or default_conf()
    cf = MpbaIntegrationConf.from_buffer_copy(conf)
    cf.cg_max_iter = int(max_iter)
    if not inputs:
        return [], []
    b = _Batch(inputs, [None] * len(inputs), device, variances=True)
    _with_handle(device, lambda lib, h: lib.mpba_integration_variances(h, b.num, b.jobs, C.byref(cf), 1 if ignore_depths else 0,
                                                                       float(tol), b.workspace.data_ptr(), b.workspace.numel(),
                                                                       b.results))
    ys = [b.out[k][3].cpu().numpy().reshape(b.out[k][4]) for k in range(b.num)]
    return ys, [int(b.results[k].cg_iterations) for k in range(b.num)]


class DepthIntegrator:
    """The mapper-facing half: ``MpsfmMapper.integrate_bundle`` (mpsfm/sfm/mapper/base.py:619-631) over the reference's
    own objects.  ``mpsfm_rec.images[imid]`` must offer what ``Integration`` reads: ``.depth.{data, data_prior,
    uncertainty, valid, activated, uncertainty_update, data_prior_at_kps}``, ``.normals.{data, uncertainty}``,
    ``.camera.{sx, sy, calibration_matrix()}``, ``.conf`` (integration keys) and ``mpsfm_rec.project_image_3d_points``,
    ``mpsfm_rec.point_covs.points_zvars``, ``mpsfm_rec.find_points3D_with_small_triangulation_angle``,
    ``mpsfm_rec.keypoints``.  State (IntVars) is kept per image id by this object."""

    def __init__(self, mpsfm_rec, device: int = 0):
        self.mpsfm_rec = mpsfm_rec
        self.device = device
        self.state = {}
        self.first_refinement = True

    # _prepare_integration_variables (:92-131)
    def _prepare(self, imid):
        rec = self.mpsfm_rec
        image = rec.images[imid]
        conf = image.conf
        _, pts3dids, kps, depth3d, success = rec.project_image_3d_points(imid)
        if not success:
            return None
        pts3dids = np.array(pts3dids)
        kps = np.array(kps, dtype=np.float64).reshape(-1, 2)
        depth3d = np.array(depth3d, dtype=np.float64)
        rt = getattr(conf, "robust_triangles", None)
        if rt is not None and len(pts3dids):
            safe = ~np.asarray(rec.find_points3D_with_small_triangulation_angle(min_angle=rt, point3D_ids=pts3dids), dtype=bool)
            pts3dids, kps, depth3d = pts3dids[safe], kps[safe], depth3d[safe]
        cam = image.camera
        kps = kps * np.array([cam.sx, cam.sy])
        kps = (kps + 0.5).astype(int)
        H, W = image.depth.data.shape
        if len(pts3dids) == 0:
            zvars = np.zeros(0)
            mask = slice(None)
        else:
            _, zvars = rec.point_covs.points_zvars(image, pts3dids)
            zvars = np.asarray(zvars, dtype=np.float64)
            x, y = kps.T
            mask = (x >= 0) & (x < W) & (y >= 0) & (y < H)
        Km = cam.calibration_matrix()
        K = [Km[1, 1] * cam.sy, Km[0, 0] * cam.sx, Km[1, 2] * cam.sy, Km[0, 2] * cam.sx]
        return MapInputs(depth_prior=image.depth.data_prior, uncertainty=image.depth.uncertainty, valid=image.depth.valid,
                         normals=image.normals.data, normals_uncertainty=image.normals.uncertainty, depth=image.depth.data,
                         K=K, kps=kps[mask], depth3d=depth3d[mask], zvars3d=zvars[mask])

    @staticmethod
    def _conf_of(image) -> MpbaIntegrationConf:
        c = image.conf
        keys = ("max_iter", "cg_max_iter", "tol", "step_size", "cg_tol", "lambda1", "lambda2", "k", "large_number",
                "depth_magnitude_multiplier", "normals_magnitude_multiplier")
        kw = {k: getattr(c, k) for k in keys if hasattr(c, k)}
        if getattr(c, "scale_filter", True):
            kw["scale_filter_factor"] = float(getattr(c, "scale_filter_factor", 1.5))
        else:
            kw["scale_filter"] = False
        return default_conf(**kw)

    def integrate_bundle(self, imids, int_covs=True, int_covs_every_iter=False, **kwargs):
        """Integrates bundle for given image ids (base.py:619-631): every image whose energy moved is refined, all in
        ONE launch; then (first refinement only, :623-627) the propagated variances of the changed images.
        Returns {imid: changed}."""
        rec = self.mpsfm_rec
        ids, inputs, states, confs = [], [], [], []
        for imid in imids:
            image = rec.images[imid]
            assert image.depth.activated, "Image not registered or depth map not activated"  # :135
            m = self._prepare(imid)
            if m is None:
                continue
            ids.append(imid)
            inputs.append(m)
            states.append(self.state.get(imid))
            confs.append(self._conf_of(image))
        changed = {imid: False for imid in imids}
        # images sharing a conf go in one launch (the mapper uses one conf for all of them)
        groups = {}
        for k, cf in enumerate(confs):
            groups.setdefault(bytes(cf), []).append(k)
        for key, members in groups.items():
            outs = integrate_maps([inputs[k] for k in members], confs[members[0]], [states[k] for k in members], self.device)
            for k, out in zip(members, outs):
                imid = ids[k]
                self.state[imid] = out.state
                if out.updated:
                    rec.images[imid].depth.data = out.depth.astype(np.float64)  # :515
                    changed[imid] = True
        if int_covs and self.first_refinement:
            todo = [k for k in range(len(ids)) if changed[ids[k]]]
            if todo:
                self.calculate_int_covs_at_kps([ids[k] for k in todo])
        if not int_covs_every_iter:
            self.first_refinement = False
        return changed

    def calculate_int_covs_at_kps(self, imids, ignore_depths=None):
        """calculate_hessian + calculate_int_covs_at_kps (:522-616) at the maps' own resolution (``downscaled`` needs the
        caller's cv2-resized maps: pass them as the image's maps).  Writes depth.uncertainty_update in place."""
        rec = self.mpsfm_rec
        inputs, confs = [], []
        for imid in imids:
            image = rec.images[imid]
            m = self._prepare(imid)
            inputs.append(m)
            confs.append(self._conf_of(image))
        ign = ignore_depths if ignore_depths is not None else bool(getattr(rec.images[imids[0]].conf, "ignore_depths", True))
        ys, _ = propagate_variances(inputs, confs[0], ignore_depths=ign, device=self.device)
        out = {}
        for imid, y in zip(imids, ys):
            image = rec.images[imid]
            kps = np.asarray(rec.keypoints(imid), dtype=np.float64)
            cam = image.camera
            xy = np.round(kps * np.array([cam.sx, cam.sy])).astype(int)
            log_uncert = y[xy[:, 1], xy[:, 0]]
            uncert = log_uncert * np.asarray(image.depth.data_prior_at_kps(kps)) ** 2  # :614
            image.depth.uncertainty_update[np.arange(len(kps))] = uncert
            out[imid] = uncert
        return out
