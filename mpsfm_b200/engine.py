"""ctypes binding of ``libmpba.so`` (C ABI in ``include/mpba.h``) and the host-side problem layout.

PyTorch is used for one thing only: owning the device buffers whose raw pointers go into
``mpba_problem``.  All arithmetic of a solve runs inside the library's sm_100a kernels.  There is
no CPU path: every compute entry point raises :class:`EngineUnavailable` when the shared library
or a CUDA device is missing.

Layout built here:

* *track order* -- observations grouped by point3D (``pt_ptr``), sorted by image inside a track: feeds
  the point-keyed reductions, the Schur complement and back-substitution,
* *image order* -- observations grouped by image (``img_ptr``) in chunks of <= 1024: feeds the
  image-keyed reductions,
* the block-CSR structure of the reduced camera system over co-visible image pairs and, for every
  track, the block index of each of its image pairs.
"""

from __future__ import annotations

import atexit
import ctypes as C
import dataclasses
import os
from typing import Optional

import numpy as np

from .problem import MAX_CAMERA_PARAMS, BundleProblem

_LIB_PATH = os.environ.get("MPBA_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libmpba.so")
MPBA_CHUNK = 1024


def chunk_size(num_obs: int) -> int:
    """mpba_chunk_size of include/mpba.h: image-order chunk length for a bundle of ``num_obs`` observations."""
    chunk = MPBA_CHUNK
    while chunk > 128 and num_obs // chunk < 592:
        chunk //= 2
    return chunk
MPBA_NUM_KERNELS = 14
MPBA_UNIQUE_ID_BYTES = 128
FLAG_STRUCTURE_ONLY = 1

TERMINATION = {0: "CONVERGENCE", 1: "NO_CONVERGENCE", 2: "FAILURE"}


class EngineUnavailable(RuntimeError):
    """libmpba.so or a CUDA device is missing -- there is deliberately no fallback."""


class EngineError(RuntimeError):
    pass


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_bp = C.POINTER(C.c_uint8)


class MpbaProblem(C.Structure):
    _fields_ = [
        ("num_images", C.c_int32), ("num_cameras", C.c_int32), ("num_points", C.c_int32), ("num_obs", C.c_int32),
        ("img_pose", C.c_void_p), ("img_camera", C.c_void_p), ("img_free_mask", C.c_void_p),
        ("img_shift_scale", C.c_void_p), ("cam_model", C.c_void_p), ("cam_params", C.c_void_p),
        ("pt_xyz", C.c_void_p), ("pt_const", C.c_void_p),
        ("pt_ptr", C.c_void_p), ("pm_img", C.c_void_p), ("pm_xy", C.c_void_p), ("pm_depth", C.c_void_p),
        ("pm_src", C.c_void_p),
        ("img_ptr", C.c_void_p), ("cm_pt", C.c_void_p), ("cm_xy", C.c_void_p), ("cm_depth", C.c_void_p),
        ("num_chunks", C.c_int32),
        ("chunk_img", C.c_void_p), ("chunk_begin", C.c_void_p), ("chunk_end", C.c_void_p),
        ("img_chunk_ptr", C.c_void_p),
        ("num_blocks", C.c_int32),
        ("s_row_ptr", C.c_void_p), ("s_col", C.c_void_p), ("s_diag", C.c_void_p), ("s_transpose", C.c_void_p),
        ("s_row", C.c_void_p), ("num_upper", C.c_int32),
        ("u_block", C.c_void_p), ("u_item_ptr", C.c_void_p), ("item_a", C.c_void_p), ("item_b", C.c_void_p),
        ("reproj_loss", C.c_int32), ("depth_loss", C.c_int32),
        ("reproj_loss_scale", C.c_double), ("reproj_loss_weight", C.c_double),
        ("flags", C.c_int32), ("s_max_row_blocks", C.c_int32),
        ("num_track_warps", C.c_int32), ("reserved", C.c_int32), ("track_warp_first", C.c_void_p),
        # normal-prior rows (extension, include/mpba.h); NULL / 0: none
        ("pm_plane", C.c_void_p), ("cm_plane", C.c_void_p), ("plane_loss_scale", C.c_double),
    ]


class MpbaShard(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32),
        ("pt_begin", C.c_int32), ("pt_end", C.c_int32), ("pt_stride", C.c_int32),
        ("img_begin", C.c_int32), ("img_end", C.c_int32), ("reserved", C.c_int32),
    ]


class MpbaOptions(C.Structure):
    _fields_ = [
        ("max_num_iterations", C.c_int32), ("max_num_consecutive_invalid_steps", C.c_int32),
        ("jacobi_scaling", C.c_int32), ("pcg_max_iterations", C.c_int32),
        ("initial_trust_region_radius", C.c_double), ("max_trust_region_radius", C.c_double),
        ("min_trust_region_radius", C.c_double), ("min_relative_decrease", C.c_double),
        ("min_lm_diagonal", C.c_double), ("max_lm_diagonal", C.c_double),
        ("function_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
        ("parameter_tolerance", C.c_double), ("pcg_tolerance", C.c_double),
        ("profile", C.c_int32), ("verbose", C.c_int32), ("pcg_kernel", C.c_int32), ("pcg_coarse", C.c_int32),
        ("behind_camera", C.c_int32), ("coarse_lag", C.c_int32), ("lm_loop", C.c_int32), ("reserved0", C.c_int32),
    ]


class MpbaIteration(C.Structure):
    _fields_ = [
        ("cost", C.c_double), ("cost_change", C.c_double), ("gradient_max_norm", C.c_double),
        ("step_norm", C.c_double), ("relative_decrease", C.c_double), ("trust_region_radius", C.c_double),
        ("step_is_valid", C.c_int32), ("step_is_successful", C.c_int32),
        ("pcg_iterations", C.c_int32), ("reserved", C.c_int32),
    ]


class MpbaSummary(C.Structure):
    _fields_ = [
        ("initial_cost", C.c_double), ("final_cost", C.c_double), ("fixed_cost", C.c_double),
        ("termination", C.c_int32), ("num_iterations", C.c_int32),
        ("num_successful_steps", C.c_int32), ("num_unsuccessful_steps", C.c_int32),
        ("total_pcg_iterations", C.c_int64), ("kernel_launches", C.c_int64),
        ("solve_ms", C.c_double),
        ("iterations", C.POINTER(MpbaIteration)), ("capacity", C.c_int32), ("num_behind_camera", C.c_int32),
        ("message", C.c_char * 128),
    ]


EXPORTS = (
    "mpba_version", "mpba_chunk_size", "mpba_create", "mpba_destroy", "mpba_last_error", "mpba_default_options",
    "mpba_workspace_bytes", "mpba_evaluate", "mpba_linearize", "mpba_reduced_system", "mpba_solve",
    "mpba_kernel_times", "mpba_kernel_name", "mpba_point_covariances", "mpba_filter_tracks",
    "mpba_pack_tracks", "mpba_sample_maps", "mpba_depth_prior_terms", "mpba_segmented_median_workspace_bytes",
    "mpba_segmented_median", "mpba_segmented_median_static", "mpba_gather_observations", "mpba_host_layout_create", "mpba_host_layout_destroy",
    "mpba_host_gather", "mpba_comm_unique_id", "mpba_comm_init", "mpba_workspace_bytes_sharded", "mpba_solve_sharded",
    "mpba_integration_default_conf", "mpba_integration_workspace_bytes", "mpba_integrate_depth",
    "mpba_integration_variances",
)

_lib = None


def load_library(path: Optional[str] = None):
    """dlopen libmpba.so and declare the prototypes of include/mpba.h.  Needs no GPU."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or _LIB_PATH
    if not os.path.exists(p):
        raise EngineUnavailable(f"{p} not found -- build it with `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(p)
    lib.mpba_version.restype = C.c_int
    lib.mpba_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_void_p]
    lib.mpba_destroy.argtypes = [C.c_void_p]
    lib.mpba_destroy.restype = None
    lib.mpba_last_error.argtypes = [C.c_void_p]
    lib.mpba_last_error.restype = C.c_char_p
    lib.mpba_default_options.argtypes = [C.POINTER(MpbaOptions)]
    lib.mpba_default_options.restype = None
    lib.mpba_workspace_bytes.argtypes = [C.POINTER(MpbaProblem), C.POINTER(C.c_size_t)]
    lib.mpba_evaluate.argtypes = [C.c_void_p, C.POINTER(MpbaProblem), C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, _dp]
    lib.mpba_linearize.argtypes = [C.c_void_p, C.POINTER(MpbaProblem), C.c_void_p, C.c_size_t,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, _dp]
    lib.mpba_reduced_system.argtypes = [C.c_void_p, C.POINTER(MpbaProblem), C.c_void_p, C.c_size_t,
                                        C.POINTER(MpbaOptions), C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, _ip]
    lib.mpba_point_covariances.argtypes = [C.c_void_p, C.POINTER(MpbaProblem), C.c_void_p, C.c_size_t, C.c_void_p]
    lib.mpba_filter_tracks.argtypes = [C.c_void_p, C.POINTER(MpbaProblem), C.c_void_p, C.c_size_t, C.c_double, C.c_double,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
    lib.mpba_solve.argtypes = [C.c_void_p, C.POINTER(MpbaProblem), C.c_void_p, C.c_size_t,
                               C.POINTER(MpbaOptions), C.POINTER(MpbaSummary)]
    lib.mpba_sample_maps.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpba_depth_prior_terms.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_double, C.c_int32, C.c_double, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpba_segmented_median_workspace_bytes.argtypes = [C.c_int32, C.c_int32, C.POINTER(C.c_size_t)]
    lib.mpba_segmented_median_static.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                                 C.c_void_p, C.c_void_p]
    lib.mpba_segmented_median.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t,
                                          C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpba_host_layout_create.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                            C.POINTER(C.c_void_p), C.c_void_p]
    lib.mpba_host_layout_destroy.argtypes = [C.c_void_p]
    lib.mpba_host_layout_destroy.restype = None
    lib.mpba_host_gather.argtypes = [C.c_int32] + [C.c_void_p] * 7
    lib.mpba_gather_observations.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
    lib.mpba_pack_tracks.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_int32)]
    lib.mpba_comm_unique_id.argtypes = [C.c_void_p, C.c_char_p]
    lib.mpba_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_char_p]
    lib.mpba_workspace_bytes_sharded.argtypes = [C.POINTER(MpbaProblem), C.POINTER(MpbaShard), C.POINTER(C.c_size_t)]
    lib.mpba_solve_sharded.argtypes = [C.c_void_p, C.POINTER(MpbaProblem), C.POINTER(MpbaShard), C.c_void_p, C.c_size_t,
                                       C.POINTER(MpbaOptions), C.POINTER(MpbaSummary)]
    lib.mpba_kernel_times.argtypes = [C.c_void_p, _dp, C.POINTER(C.c_int64)]
    lib.mpba_kernel_name.argtypes = [C.c_int]
    lib.mpba_kernel_name.restype = C.c_char_p
    if path is None:
        _lib = lib
    return lib


def default_options(**overrides) -> MpbaOptions:
    opt = MpbaOptions()
    load_library().mpba_default_options(C.byref(opt))
    for k, v in overrides.items():
        if not hasattr(opt, k):
            raise AttributeError(f"mpba_options has no field {k!r}")
        setattr(opt, k, v)
    return opt


# ----------------------------------------------------------------------------------------------
# host-side layout (numpy only; runs anywhere)
# ----------------------------------------------------------------------------------------------
@dataclasses.dataclass
class HostLayout:
    C: int
    K: int
    P: int
    N: int
    img_pose: np.ndarray  # [C,8]
    img_camera: np.ndarray
    img_free_mask: np.ndarray
    img_shift_scale: np.ndarray  # [C,2]
    cam_model: np.ndarray
    cam_params: np.ndarray  # [K,12]
    pt_xyz: np.ndarray  # [P,4]
    pt_const: np.ndarray
    pt_ptr: np.ndarray
    pm_img: np.ndarray
    pm_xy: np.ndarray
    pm_depth: np.ndarray  # [3,N]
    pm_src: np.ndarray
    img_ptr: np.ndarray
    cm_pt: np.ndarray
    cm_xy: np.ndarray
    cm_depth: np.ndarray
    cm_src: np.ndarray
    chunk_img: np.ndarray
    chunk_begin: np.ndarray
    chunk_end: np.ndarray
    img_chunk_ptr: np.ndarray
    s_row_ptr: np.ndarray
    s_col: np.ndarray
    s_diag: np.ndarray
    s_transpose: np.ndarray
    s_row: np.ndarray
    u_block: np.ndarray
    u_item_ptr: np.ndarray
    item_a: np.ndarray
    item_b: np.ndarray
    track_warp_first: np.ndarray  # [num_track_warps + 1] (mpba_pack_tracks)
    reproj_loss: int
    depth_loss: int
    reproj_loss_scale: float
    reproj_loss_weight: float
    flags: int
    cm_pm: Optional[np.ndarray] = None  # [N] track-order slot of every image-order slot (unsharded layouts)
    # normal-prior rows (extension, mpsfm_b200/normal_prior.py): {m, c} per slot in both orders, or None
    pm_plane: Optional[np.ndarray] = None  # [N,4]
    cm_plane: Optional[np.ndarray] = None  # [N,4]
    plane_loss_scale: float = 1.0

    @property
    def num_blocks(self) -> int:
        return int(self.s_col.shape[0])

    @property
    def num_chunks(self) -> int:
        return int(self.chunk_img.shape[0])


def pack_tracks(pt_ptr: np.ndarray, pt_begin: int, pt_end: int) -> np.ndarray:
    """Warp groups of whole tracks for the track-order passes (host helper of libmpba.so: a sequential greedy
    scan, <= 32 observations and <= 32 tracks per group)."""
    lib = load_library()
    ptr = np.ascontiguousarray(pt_ptr, dtype=np.int32)
    first = np.empty(pt_end - pt_begin + 1, dtype=np.int32)
    nw = C.c_int32()
    rc = lib.mpba_pack_tracks(ptr.ctypes.data, int(pt_begin), int(pt_end), first.ctypes.data, C.byref(nw))
    if rc != 0:
        raise EngineError(f"mpba_pack_tracks failed with code {rc}")
    return first[: nw.value + 1].copy()


def _ptr_from_counts(counts: np.ndarray) -> np.ndarray:
    out = np.zeros(counts.shape[0] + 1, dtype=np.int64)
    np.cumsum(counts, out=out[1:])
    if out[-1] >= 2**31:
        raise ValueError("more than 2^31-1 entries: shard the problem across GPUs")
    return out.astype(np.int32)


class MpbaHostLayoutView(C.Structure):
    _fields_ = ([(n, C.c_int32) for n in ("num_images", "num_points", "num_obs", "num_chunks", "num_blocks", "num_upper",
                                         "num_track_warps")]
                + [("num_items", C.c_int64)]
                + [(n, C.POINTER(C.c_int32)) for n in (
                    "pt_ptr", "pm_img", "pm_src", "img_ptr", "cm_pt", "cm_src", "cm_pm", "chunk_img", "chunk_begin",
                    "chunk_end", "img_chunk_ptr", "s_row_ptr", "s_col", "s_diag", "s_transpose", "s_row", "u_block",
                    "u_item_ptr", "item_a", "item_b", "track_warp_first")])


class _NativeLayoutOwner:
    """Keeps a mpba_host_layout alive while numpy views of its buffers exist."""

    def __init__(self, lib, handle):
        self._lib, self._handle = lib, handle

    def __del__(self):
        if self._handle is not None and self._handle.value:
            self._lib.mpba_host_layout_destroy(self._handle)
            self._handle = None


_SCRATCH = {}


def _scratch(name: str, n: int) -> np.ndarray:
    buf = _SCRATCH.get(name)
    if buf is None or buf.shape[0] < n:
        buf = _SCRATCH[name] = np.empty(max(n, 1024) * 5 // 4, dtype=np.float64)
    return buf[:n]


def _plane_slots(pr: BundleProblem, pm_src, cm_src) -> dict:
    """HostLayout fields of the normal-prior rows (extension): the {m, c} record of every slot in both orders."""
    if not pr.num_plane:
        return {}
    full = np.zeros((pr.num_obs, 4))
    full[pr.plane_obs, :3], full[pr.plane_obs, 3] = pr.plane_m, pr.plane_c
    return {"pm_plane": np.ascontiguousarray(full[pm_src]), "cm_plane": np.ascontiguousarray(full[cm_src]),
            "plane_loss_scale": float(pr.plane_loss_scale)}


def build_layout(problem: BundleProblem) -> HostLayout:
    """Sort the observations into the two orders and build the reduced-system structure: the native, multi-threaded
    builder of libmpba.so (mpba_host_layout_create / mpba_host_gather, csrc/host_layout.cpp).

    The natural observation index of every slot is kept (``pm_src`` / ``cm_src``): per-residual outputs are mapped
    back through it, which is the "correspondence / track indexing bit-exact" gate of BASELINE.json.
    ``build_layout_numpy`` states the same layout in numpy; tests/test_host_layout.py holds the two equal."""
    lib = load_library()
    pr = problem.normalized()
    pr.validate()
    Cn, P, N = pr.num_images, pr.num_points, pr.num_obs
    K = int(pr.cam_model.shape[0])
    if pr.num_depth and not np.all(pr.depth_w > 0):
        raise ValueError("depth prior weights must be > 0 (w == 0 marks 'no prior' in the device layout)")
    obs_img = np.ascontiguousarray(pr.obs_img, dtype=np.int32)
    obs_pt = np.ascontiguousarray(pr.obs_pt, dtype=np.int32)
    pt_const = np.ascontiguousarray(pr.pt_const, dtype=np.uint8)
    structure_only = bool(np.all(pr.img_free_mask == 0))
    view, handle = MpbaHostLayoutView(), C.c_void_p()
    rc = lib.mpba_host_layout_create(Cn, P, N, obs_img.ctypes.data, obs_pt.ctypes.data, pt_const.ctypes.data,
                                     int(structure_only), C.byref(handle), C.byref(view))
    if rc != 0:
        raise ValueError(f"mpba_host_layout_create failed with code {rc} (index out of range, or more than 2^31-1 "
                         "Schur items: shard the problem across GPUs)")
    # the index arrays are VIEWS of the native builder's buffers (no copy: at the mapper's bundle sizes the copies cost
    # as much as building them); the owner object frees the native layout when the last array is gone
    owner = _NativeLayoutOwner(lib, handle)

    def take(name, n):
        if not n:
            return np.zeros(0, dtype=np.int32)
        # the ctypes array is the base object of every numpy view derived from it, and it holds the owner
        mem = (C.c_int32 * n).from_address(C.cast(getattr(view, name), C.c_void_p).value)
        mem._owner = owner
        return np.frombuffer(mem, dtype=np.int32)

    B, U = view.num_blocks, view.num_upper
    idx = {name: take(name, n) for name, n in (
        ("pt_ptr", P + 1), ("pm_img", N), ("pm_src", N), ("img_ptr", Cn + 1), ("cm_pt", N), ("cm_src", N), ("cm_pm", N),
        ("chunk_img", view.num_chunks), ("chunk_begin", view.num_chunks), ("chunk_end", view.num_chunks),
        ("img_chunk_ptr", Cn + 1), ("s_row_ptr", Cn + 1), ("s_col", B), ("s_diag", Cn), ("s_transpose", B),
        ("s_row", B), ("u_block", U), ("u_item_ptr", U + 1), ("item_a", view.num_items),
        ("item_b", view.num_items), ("track_warp_first", view.num_track_warps + 1))}
    # d / w / a in natural order are temporaries of this call: grow-only scratch instead of 24 N bytes of fresh pages
    d, w, a = _scratch("d", N), _scratch("w", N), _scratch("a", N)
    d.fill(1.0), w.fill(0.0), a.fill(1.0)
    d[pr.depth_obs], w[pr.depth_obs], a[pr.depth_obs] = pr.depth_d, pr.depth_w, pr.depth_a
    obs_xy = np.ascontiguousarray(pr.obs_xy, dtype=np.float64)
    vals = {}
    for order in ("pm", "cm"):
        xy, dep = np.empty((N, 2)), np.empty((3, N))
        rc = lib.mpba_host_gather(N, idx[order + "_src"].ctypes.data, obs_xy.ctypes.data, d.ctypes.data, w.ctypes.data,
                                  a.ctypes.data, xy.ctypes.data, dep.ctypes.data)
        if rc != 0:
            raise EngineError(f"mpba_host_gather failed with code {rc}")
        vals[order] = (xy, dep)
    pose = np.zeros((Cn, 8))
    pose[:, :4], pose[:, 4:7] = pr.img_q, pr.img_t
    xyz = np.zeros((P, 4))
    xyz[:, :3] = pr.pt_xyz
    params = np.zeros((K, MAX_CAMERA_PARAMS))
    params[:, : pr.cam_params.shape[1]] = pr.cam_params
    return HostLayout(
        C=Cn, K=K, P=P, N=N,
        img_pose=pose, img_camera=pr.img_camera, img_free_mask=pr.img_free_mask,
        img_shift_scale=np.ascontiguousarray(pr.shift_scale), cam_model=pr.cam_model, cam_params=params,
        pt_xyz=xyz, pt_const=pr.pt_const,
        pm_xy=vals["pm"][0], pm_depth=vals["pm"][1], cm_xy=vals["cm"][0], cm_depth=vals["cm"][1],
        reproj_loss=int(pr.reproj_loss), depth_loss=int(pr.depth_loss),
        reproj_loss_scale=float(pr.reproj_loss_scale), reproj_loss_weight=float(pr.reproj_loss_weight),
        flags=FLAG_STRUCTURE_ONLY if structure_only else 0, **_plane_slots(pr, idx["pm_src"], idx["cm_src"]), **idx)


def build_layout_numpy(problem: BundleProblem) -> HostLayout:
    """The same layout stated in numpy (general sorts): the checker of the native builder."""
    pr = problem.normalized()
    pr.validate()
    Cn, P, N = pr.num_images, pr.num_points, pr.num_obs
    K = int(pr.cam_model.shape[0])
    oi = pr.obs_img.astype(np.int64)
    op = pr.obs_pt.astype(np.int64)

    d = np.ones(N)
    w = np.zeros(N)
    a = np.ones(N)
    d[pr.depth_obs] = pr.depth_d
    w[pr.depth_obs] = pr.depth_w
    a[pr.depth_obs] = pr.depth_a
    if pr.num_depth and not np.all(pr.depth_w > 0):
        raise ValueError("depth prior weights must be > 0 (w == 0 marks 'no prior' in the device layout)")

    # ---- track order: by point, then image
    pm_src = np.argsort(op * Cn + oi, kind="stable").astype(np.int32)
    pt_cnt = np.bincount(op, minlength=P)
    pt_ptr = _ptr_from_counts(pt_cnt)
    pm_img = pr.obs_img[pm_src]
    pm_xy = np.ascontiguousarray(pr.obs_xy[pm_src])
    pm_depth = np.ascontiguousarray(np.stack([d[pm_src], w[pm_src], a[pm_src]], axis=0))

    # ---- image order: by image, then point (ascending gather addresses)
    cm_src = np.argsort(oi * max(P, 1) + op, kind="stable").astype(np.int32)
    img_cnt = np.bincount(oi, minlength=Cn)
    img_ptr = _ptr_from_counts(img_cnt)
    cm_pt = pr.obs_pt[cm_src]
    cm_xy = np.ascontiguousarray(pr.obs_xy[cm_src])
    cm_depth = np.ascontiguousarray(np.stack([d[cm_src], w[cm_src], a[cm_src]], axis=0))

    CHUNK = chunk_size(int(img_ptr[-1]))
    nch = (img_cnt + CHUNK - 1) // CHUNK
    img_chunk_ptr = _ptr_from_counts(nch)
    chunk_img = np.repeat(np.arange(Cn, dtype=np.int32), nch)
    local = np.arange(chunk_img.shape[0], dtype=np.int64) - img_chunk_ptr[chunk_img]
    chunk_begin = (img_ptr[chunk_img] + local * CHUNK).astype(np.int32)
    chunk_end = np.minimum(chunk_begin.astype(np.int64) + CHUNK, img_ptr[chunk_img + 1]).astype(np.int32)

    # ---- reduced camera system structure and the Schur work lists (one unit per upper block)
    structure_only = bool(np.all(pr.img_free_mask == 0))
    active = (pr.pt_const == 0)
    L = np.where(active, pt_cnt, 0).astype(np.int64)
    npairs = L * (L + 1) // 2
    pair_ptr = _ptr_from_counts(npairs)
    n_items = 0 if structure_only else int(pair_ptr[-1])
    pair_key = np.empty(n_items, dtype=np.int64)
    item_a = np.empty(n_items, dtype=np.int32)
    item_b = np.empty(n_items, dtype=np.int32)
    if not structure_only:
        for ell in np.unique(L):
            ell = int(ell)
            if ell == 0:
                continue
            pts = np.flatnonzero(L == ell)
            slots = pt_ptr[pts].astype(np.int64)[:, None] + np.arange(ell)[None, :]
            imgs = pm_img[slots].astype(np.int64)  # [n, ell], ascending in every row
            ia, ib = np.triu_indices(ell)  # row-major over a, then b >= a
            dst = (pair_ptr[pts].astype(np.int64)[:, None] + np.arange(ia.shape[0])[None, :]).reshape(-1)
            pair_key[dst] = (imgs[:, ia] * Cn + imgs[:, ib]).reshape(-1)
            item_a[dst] = slots[:, ia].reshape(-1)
            item_b[dst] = slots[:, ib].reshape(-1)
    diag_keys = np.arange(Cn, dtype=np.int64) * (Cn + 1)
    if structure_only:
        all_keys = diag_keys
        item_block = np.zeros(0, dtype=np.int32)
    else:
        upper = np.unique(pair_key)
        ui, uj = upper // Cn, upper % Cn
        lower = (uj * Cn + ui)[ui != uj]
        all_keys = np.unique(np.concatenate([upper, lower, diag_keys]))
        item_block = np.searchsorted(all_keys, pair_key).astype(np.int32)
    del pair_key
    rows, cols = all_keys // Cn, all_keys % Cn
    s_row_ptr = _ptr_from_counts(np.bincount(rows, minlength=Cn))
    s_col = cols.astype(np.int32)
    s_row = rows.astype(np.int32)
    s_diag = np.searchsorted(all_keys, diag_keys).astype(np.int32)
    s_transpose = np.searchsorted(all_keys, cols * Cn + rows).astype(np.int32)
    u_block = np.flatnonzero(rows <= cols).astype(np.int32)
    # items were generated track by track, so a stable sort by block leaves every unit sorted by track
    order = np.argsort(item_block, kind="stable")
    item_a, item_b = item_a[order], item_b[order]
    u_item_ptr = _ptr_from_counts(np.bincount(item_block, minlength=all_keys.shape[0])[u_block])
    del order, item_block

    pose = np.zeros((Cn, 8))
    pose[:, :4] = pr.img_q
    pose[:, 4:7] = pr.img_t
    xyz = np.zeros((P, 4))
    xyz[:, :3] = pr.pt_xyz
    params = np.zeros((K, MAX_CAMERA_PARAMS))
    params[:, : pr.cam_params.shape[1]] = pr.cam_params
    pm_pos = np.empty(N, dtype=np.int32)
    pm_pos[pm_src] = np.arange(N, dtype=np.int32)

    return HostLayout(
        cm_pm=pm_pos[cm_src], **_plane_slots(pr, pm_src, cm_src),
        C=Cn, K=K, P=P, N=N,
        img_pose=pose, img_camera=pr.img_camera, img_free_mask=pr.img_free_mask,
        img_shift_scale=np.ascontiguousarray(pr.shift_scale), cam_model=pr.cam_model, cam_params=params,
        pt_xyz=xyz, pt_const=pr.pt_const,
        pt_ptr=pt_ptr, pm_img=np.ascontiguousarray(pm_img), pm_xy=pm_xy, pm_depth=pm_depth, pm_src=pm_src,
        img_ptr=img_ptr, cm_pt=np.ascontiguousarray(cm_pt), cm_xy=cm_xy, cm_depth=cm_depth, cm_src=cm_src,
        chunk_img=chunk_img, chunk_begin=chunk_begin, chunk_end=chunk_end, img_chunk_ptr=img_chunk_ptr,
        s_row_ptr=s_row_ptr, s_col=s_col, s_diag=s_diag, s_transpose=s_transpose,
        s_row=s_row, u_block=u_block, u_item_ptr=u_item_ptr, item_a=item_a, item_b=item_b,
        track_warp_first=pack_tracks(pt_ptr, 0, P),
        reproj_loss=int(pr.reproj_loss), depth_loss=int(pr.depth_loss),
        reproj_loss_scale=float(pr.reproj_loss_scale), reproj_loss_weight=float(pr.reproj_loss_weight),
        flags=FLAG_STRUCTURE_ONLY if structure_only else 0,
    )


# ----------------------------------------------------------------------------------------------
# sharding (include/mpba.h: mpba_shard; DESIGN.md "Multi-GPU")
# ----------------------------------------------------------------------------------------------
@dataclasses.dataclass
class ShardSpec:
    rank: int
    world: int
    pt_begin: int
    pt_end: int
    pt_stride: int
    img_begin: int
    img_end: int


def shard_ranges(lay: HostLayout, world: int):
    """Owned ranges of every rank: images in contiguous ranges balanced by observation count (north star:
    "observations are sharded by camera"), points in equal contiguous ranges (equal counts for the all-gather)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    targets = (np.arange(world + 1, dtype=np.int64) * int(lay.N)) // world
    ib = np.searchsorted(lay.img_ptr.astype(np.int64), targets, side="left").astype(np.int64)
    ib[0], ib[-1] = 0, lay.C
    ib = np.maximum.accumulate(np.clip(ib, 0, lay.C))
    stride = (lay.P + world - 1) // world if lay.P else 0
    out = []
    for r in range(world):
        p0 = min(r * stride, lay.P)
        out.append(ShardSpec(rank=r, world=world, pt_begin=p0, pt_end=min(p0 + stride, lay.P), pt_stride=stride,
                             img_begin=int(ib[r]), img_end=int(ib[r + 1])))
    return out


def shard_layout(lay: HostLayout, rank: int, world: int):
    """The rank's share of a layout: image-order chunks of its own images, tracks (and Schur items) of its own
    points; poses, points and the reduced-system structure stay global.  Returns (HostLayout, ShardSpec)."""
    spec = shard_ranges(lay, world)[rank]
    s0, s1 = int(lay.pt_ptr[spec.pt_begin]), int(lay.pt_ptr[spec.pt_end])
    c0, c1 = int(lay.img_ptr[spec.img_begin]), int(lay.img_ptr[spec.img_end])
    n_loc = max(s1 - s0, c1 - c0, 1)  # one stride for both orders (the [3][N] depth rows use num_obs)

    def pad1(a, fill=0):
        out = np.full((n_loc,) + a.shape[1:], fill, dtype=a.dtype)
        out[: a.shape[0]] = a
        return out

    def pad_rows(a):
        out = np.zeros((a.shape[0], n_loc), dtype=a.dtype)
        out[:, : a.shape[1]] = a
        out[0, a.shape[1]:] = 1.0  # d = 1 on padding (never read: no slot points there)
        return out

    ch0, ch1 = int(lay.img_chunk_ptr[spec.img_begin]), int(lay.img_chunk_ptr[spec.img_end])
    U = int(lay.u_block.shape[0])
    if lay.item_a.shape[0]:
        unit_of_item = np.repeat(np.arange(U, dtype=np.int64), np.diff(lay.u_item_ptr.astype(np.int64)))
        keep = (lay.item_a >= s0) & (lay.item_a < s1)
        item_a = (lay.item_a[keep].astype(np.int64) - s0).astype(np.int32)
        item_b = (lay.item_b[keep].astype(np.int64) - s0).astype(np.int32)
        u_item_ptr = _ptr_from_counts(np.bincount(unit_of_item[keep], minlength=U))
    else:
        item_a, item_b = lay.item_a, lay.item_b
        u_item_ptr = np.zeros(U + 1, dtype=np.int32)
    out = dataclasses.replace(
        lay, N=n_loc,
        pt_ptr=np.clip(lay.pt_ptr.astype(np.int64) - s0, 0, s1 - s0).astype(np.int32),
        pm_img=pad1(lay.pm_img[s0:s1]), pm_xy=pad1(lay.pm_xy[s0:s1]), pm_depth=pad_rows(lay.pm_depth[:, s0:s1]),
        pm_src=pad1(lay.pm_src[s0:s1], -1),
        img_ptr=np.clip(lay.img_ptr.astype(np.int64) - c0, 0, c1 - c0).astype(np.int32),
        cm_pt=pad1(lay.cm_pt[c0:c1]), cm_xy=pad1(lay.cm_xy[c0:c1]), cm_depth=pad_rows(lay.cm_depth[:, c0:c1]),
        cm_src=pad1(lay.cm_src[c0:c1], -1),
        chunk_img=lay.chunk_img[ch0:ch1].copy(),
        chunk_begin=(lay.chunk_begin[ch0:ch1].astype(np.int64) - c0).astype(np.int32),
        chunk_end=(lay.chunk_end[ch0:ch1].astype(np.int64) - c0).astype(np.int32),
        img_chunk_ptr=np.clip(lay.img_chunk_ptr.astype(np.int64) - ch0, 0, ch1 - ch0).astype(np.int32),
        u_item_ptr=u_item_ptr, item_a=item_a, item_b=item_b, cm_pm=None,
        pm_plane=None if lay.pm_plane is None else pad1(lay.pm_plane[s0:s1]),
        cm_plane=None if lay.cm_plane is None else pad1(lay.cm_plane[c0:c1]),
    )
    out.track_warp_first = pack_tracks(out.pt_ptr, spec.pt_begin, spec.pt_end)
    return out, spec


def broadcast_bytes(payload: bytes, src: int = 0) -> bytes:
    """Carry a small host blob (the NCCL id) from ``src`` to every rank over torch.distributed (any backend)."""
    import torch.distributed as dist

    box = [payload]
    dist.broadcast_object_list(box, src=src)
    return box[0]


# ----------------------------------------------------------------------------------------------
# device side
# ----------------------------------------------------------------------------------------------
_DEVICE_FIELDS = (
    "img_pose", "img_camera", "img_free_mask", "img_shift_scale", "cam_model", "cam_params", "pt_xyz", "pt_const",
    "pt_ptr", "pm_img", "pm_xy", "pm_depth", "pm_src", "img_ptr", "cm_pt", "cm_xy", "cm_depth",
    "chunk_img", "chunk_begin", "chunk_end", "img_chunk_ptr", "s_row_ptr", "s_col", "s_diag", "s_transpose",
    "s_row", "u_block", "u_item_ptr", "item_a", "item_b", "track_warp_first",
)


@dataclasses.dataclass
class SolveSummary:
    initial_cost: float
    final_cost: float
    fixed_cost: float
    termination: str
    message: str
    num_iterations: int
    num_successful_steps: int
    num_unsuccessful_steps: int
    total_pcg_iterations: int
    kernel_launches: int
    solve_ms: float
    iterations: list  # of dicts
    num_behind_camera: int = 0  # observations at z < epsilon in their camera frame at the returned state

    def brief_report(self) -> str:
        return (f"mpba: iterations {self.num_iterations}, initial cost {self.initial_cost:.6e}, final cost "
                f"{self.final_cost:.6e}, termination {self.termination} ({self.message})")


# handles of closed (unsharded) BundleAdjusters, per device: a mapper solves hundreds of bundles one after the other,
# and a handle (pinned scalar block, events, pinned plan staging) carries no per-problem state
_FREE_HANDLES = {}


def _release_handles():
    lib = _lib
    for handles in _FREE_HANDLES.values():
        while handles:
            h = handles.pop()
            if lib is not None:
                lib.mpba_destroy(h)


atexit.register(_release_handles)


class BundleAdjuster:
    """One problem resident on one GPU: buffers (torch), handle and workspace.

    ``shard=(rank, world)``: this process holds one shard of the problem (``shard_layout``) and ``solve`` runs
    ``mpba_solve_sharded`` together with the other ranks; ``torch.distributed`` must be initialised (any backend:
    it only carries the 128-byte NCCL id from rank 0 to the others)."""

    def __init__(self, problem, device: int = 0, layout: Optional[HostLayout] = None, pinned: bool = False,
                 shard: Optional[tuple] = None):
        import torch

        self.lib = load_library()
        if not torch.cuda.is_available():
            raise EngineUnavailable("no CUDA device: mpsfm_b200 has no CPU fallback")
        self.torch = torch
        self.device = torch.device("cuda", device)
        self.layout = layout if layout is not None else build_layout(problem)
        self.full_layout = self.layout
        self.shard = None
        if shard is not None:  # world == 1 is allowed: the sharded code path on one GPU (tests)
            self.layout, spec = shard_layout(self.full_layout, int(shard[0]), int(shard[1]))
            self.shard = MpbaShard(rank=spec.rank, world=spec.world, pt_begin=spec.pt_begin, pt_end=spec.pt_end,
                                   pt_stride=spec.pt_stride, img_begin=spec.img_begin, img_end=spec.img_end)
        lay = self.layout
        self.buf = {}
        self.host = {}  # pinned staging copies of the problem arrays (pinned=True): source of upload()
        self.h2d_bytes = 0
        # an unsharded problem holds the same observations in both orders: upload() sends the track order and a
        # permutation, the image-order xy / depth arrays are gathered on the device (mpba_gather_observations)
        self._derive_cm = bool(pinned and self.shard is None and lay.N > 0)
        with torch.cuda.device(self.device):
            for name in _DEVICE_FIELDS:
                arr = np.ascontiguousarray(getattr(lay, name))
                t = torch.from_numpy(arr)
                if pinned and not (self._derive_cm and name in ("cm_xy", "cm_depth")):
                    t = t.pin_memory()
                    self.host[name] = t
                self.buf[name] = t.to(self.device, non_blocking=pinned)
                self.h2d_bytes += arr.nbytes
            if lay.pm_plane is not None:
                for name in ("pm_plane", "cm_plane"):
                    t = torch.from_numpy(np.ascontiguousarray(getattr(lay, name)))
                    if pinned:
                        t = t.pin_memory()
                        self.host[name] = t
                    self.buf[name] = t.to(self.device, non_blocking=pinned)
                    self.h2d_bytes += t.numel() * 8
            if self._derive_cm:
                if lay.cm_pm is not None:
                    cm_pm = lay.cm_pm
                else:
                    pm_pos = np.empty(lay.N, dtype=np.int32)
                    pm_pos[lay.pm_src] = np.arange(lay.N, dtype=np.int32)
                    cm_pm = pm_pos[lay.cm_src]
                self.host["cm_pm"] = torch.from_numpy(np.ascontiguousarray(cm_pm, dtype=np.int32)).pin_memory()
                self.buf["cm_pm"] = self.host["cm_pm"].to(self.device, non_blocking=True)
            if pinned:
                self._pose_host = torch.empty((lay.C, 8), dtype=torch.float64).pin_memory()
                self._xyz_host = torch.empty((lay.P, 4), dtype=torch.float64).pin_memory()
            for name in ("item_a", "item_b", "chunk_img", "chunk_begin", "chunk_end"):
                if self.buf[name].numel() == 0:
                    self.buf[name] = torch.zeros(1, dtype=torch.int32, device=self.device)
            # handles are pooled per device; sharded ones per (device, world, rank) with their NCCL communicator, so the
            # hundreds of bundles of a mapper run pay mpba_create / ncclCommInitRank once
            self._pool_key = int(device) if self.shard is None else (int(device), int(self.shard.world), int(self.shard.rank))
            free = _FREE_HANDLES.setdefault(self._pool_key, [])
            self._comm_ready = False
            if free:
                self.handle = free.pop()
                self._comm_ready = self.shard is not None
            else:
                self.handle = C.c_void_p()
                rc = self.lib.mpba_create(C.byref(self.handle), device, None)
                if rc != 0:
                    raise EngineUnavailable(f"mpba_create failed with code {rc} (no usable CUDA device {device})")
            self.prob = MpbaProblem()
            self.prob.num_images, self.prob.num_cameras = lay.C, lay.K
            self.prob.num_points, self.prob.num_obs = lay.P, lay.N
            for name in _DEVICE_FIELDS:
                setattr(self.prob, name, self.buf[name].data_ptr())
            self.prob.num_chunks = lay.num_chunks
            self.prob.num_blocks = lay.num_blocks
            self.prob.num_upper = int(lay.u_block.shape[0])
            self.prob.num_track_warps = int(lay.track_warp_first.shape[0]) - 1
            self.prob.reproj_loss, self.prob.depth_loss = lay.reproj_loss, lay.depth_loss
            self.prob.reproj_loss_scale, self.prob.reproj_loss_weight = lay.reproj_loss_scale, lay.reproj_loss_weight
            self.prob.flags = lay.flags
            if lay.pm_plane is not None:
                self.prob.pm_plane, self.prob.cm_plane = self.buf["pm_plane"].data_ptr(), self.buf["cm_plane"].data_ptr()
                self.prob.plane_loss_scale = lay.plane_loss_scale
            self.prob.s_max_row_blocks = int(np.max(np.diff(lay.s_row_ptr))) if lay.C else 0
            nbytes = C.c_size_t()
            if self.shard is not None:
                if not self._comm_ready:
                    self._init_comm()
                self._check(self.lib.mpba_workspace_bytes_sharded(C.byref(self.prob), C.byref(self.shard), C.byref(nbytes)))
            else:
                self._check(self.lib.mpba_workspace_bytes(C.byref(self.prob), C.byref(nbytes)))
            self.workspace = torch.empty(max(int(nbytes.value), 256), dtype=torch.uint8, device=self.device)
            self.workspace_bytes = int(nbytes.value)
        self._pose0 = self.buf["img_pose"].clone()
        self._xyz0 = self.buf["pt_xyz"].clone()

    # -- plumbing
    def _init_comm(self):
        """NCCL communicator of the library: rank 0 draws the id, torch.distributed carries it to the others."""
        import torch.distributed as dist

        have_group = dist.is_available() and dist.is_initialized()
        if self.shard.world > 1 and not have_group:
            raise EngineError("sharded solve needs an initialised torch.distributed process group")
        ident = (C.c_uint8 * MPBA_UNIQUE_ID_BYTES)()
        if self.shard.rank == 0:
            rc = self.lib.mpba_comm_unique_id(ident, None)
            if rc != 0:
                raise EngineUnavailable(f"mpba_comm_unique_id failed with code {rc} (libnccl not loadable)")
        if self.shard.world > 1:
            ident = (C.c_uint8 * MPBA_UNIQUE_ID_BYTES).from_buffer_copy(broadcast_bytes(bytes(ident), src=0))
        self._check(self.lib.mpba_comm_init(self.handle, self.shard.world, self.shard.rank, ident, None))

    def _require_unsharded(self, what: str):
        if self.shard is not None:
            raise EngineError(f"{what} is a single-GPU entry point; a sharded BundleAdjuster only solves")

    def _check(self, rc: int):
        if rc != 0:
            msg = self.lib.mpba_last_error(self.handle)
            raise EngineError(f"libmpba error {rc}: {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "handle", None) is not None and self.handle.value:
            free = _FREE_HANDLES.setdefault(getattr(self, "_pool_key", int(self.device.index)), [])
            if len(free) < 4:
                free.append(self.handle)  # reused by the next BundleAdjuster on this device (and rank of this world)
            else:
                self.lib.mpba_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self):
        """Host -> device copy of the whole problem (every array of mpba_problem, initial poses and points
        included) from the pinned staging buffers, on the current stream: the H2D leg of an end-to-end solve."""
        if not self.host:
            raise EngineError("upload() needs a BundleAdjuster created with pinned=True")
        n = 0
        for name, t in self.host.items():
            if t.numel():
                self.buf[name].copy_(t, non_blocking=True)
                n += t.numel() * t.element_size()
        if self._derive_cm:
            self._check(self.lib.mpba_gather_observations(
                self.handle, self.layout.N, self.buf["cm_pm"].data_ptr(), self.buf["pm_xy"].data_ptr(),
                self.buf["pm_depth"].data_ptr(), self.buf["cm_xy"].data_ptr(), self.buf["cm_depth"].data_ptr(),
                C.c_void_p(self.torch.cuda.current_stream().cuda_stream)))
        return n

    def reset_state(self):
        """Restore the poses / points the problem was created with (device-to-device)."""
        self.buf["img_pose"].copy_(self._pose0)
        self.buf["pt_xyz"].copy_(self._xyz0)

    def set_state(self, img_q, img_t, pt_xyz):
        torch = self.torch
        pose = np.zeros((self.layout.C, 8))
        pose[:, :4], pose[:, 4:7] = img_q, img_t
        xyz = np.zeros((self.layout.P, 4))
        xyz[:, :3] = pt_xyz
        self.buf["img_pose"].copy_(torch.from_numpy(pose))
        self.buf["pt_xyz"].copy_(torch.from_numpy(xyz))

    def state(self, copy: bool = True):
        """(img_q [C,4], img_t [C,3], pt_xyz [P,3]) as numpy (device -> host).

        ``copy=False`` (adjusters created with ``pinned=True``): the arrays are VIEWS of the adjuster's pinned staging
        buffers -- one DMA per buffer, no host-side repacking; they stay valid until the next ``state()`` / ``close()``."""
        if self.host:
            # pinned result buffers: the device records ([C,8] poses, [P,4] points) cross PCIe as they are, one
            # contiguous DMA each; the host only slices
            self._pose_host.copy_(self.buf["img_pose"], non_blocking=True)
            self._xyz_host.copy_(self.buf["pt_xyz"], non_blocking=True)
            self.torch.cuda.current_stream().synchronize()
            pose, xyz = self._pose_host.numpy(), self._xyz_host.numpy()
            if copy:
                return pose[:, :4].copy(), pose[:, 4:7].copy(), xyz[:, :3].copy()
            return pose[:, :4], pose[:, 4:7], xyz[:, :3]
        else:
            pose = self.buf["img_pose"].cpu().numpy()
            xyz = self.buf["pt_xyz"].cpu().numpy()
        return pose[:, :4].copy(), pose[:, 4:7].copy(), xyz[:, :3].copy()

    def _ws(self):
        return C.c_void_p(self.workspace.data_ptr()), C.c_size_t(self.workspace.numel())

    # -- entry points of include/mpba.h
    def evaluate(self):
        """Raw residuals in NATURAL observation order: (reproj [N,2], depth [N] (nan where no prior),
        cost, fixed_cost)."""
        self._require_unsharded("evaluate")
        torch, lay = self.torch, self.layout
        rr = torch.empty((lay.N, 2), dtype=torch.float64, device=self.device)
        rd = torch.empty((lay.N,), dtype=torch.float64, device=self.device)
        cost = (C.c_double * 2)()
        ws, wsn = self._ws()
        self._check(self.lib.mpba_evaluate(self.handle, C.byref(self.prob), ws, wsn, rr.data_ptr(), rd.data_ptr(), cost))
        src = lay.pm_src.astype(np.int64)
        r_reproj = np.empty((lay.N, 2))
        r_depth = np.empty(lay.N)
        r_reproj[src] = rr.cpu().numpy()
        r_depth[src] = rd.cpu().numpy()
        has = np.empty(lay.N, dtype=bool)
        has[src] = lay.pm_depth[1] > 0
        r_depth[~has] = np.nan
        return r_reproj, r_depth, float(cost[0]), float(cost[1])

    def linearize(self):
        """(Hcc [C,6,6], gc [C,6], Hpp [P,3,3], gp [P,3], cost, fixed_cost) at the current state."""
        self._require_unsharded("linearize")
        torch, lay = self.torch, self.layout
        Hcc = torch.empty((lay.C, 36), dtype=torch.float64, device=self.device)
        gc = torch.empty((lay.C, 6), dtype=torch.float64, device=self.device)
        Hpp = torch.empty((max(lay.P, 1), 6), dtype=torch.float64, device=self.device)
        gp = torch.empty((max(lay.P, 1), 3), dtype=torch.float64, device=self.device)
        cost = (C.c_double * 2)()
        ws, wsn = self._ws()
        self._check(self.lib.mpba_linearize(self.handle, C.byref(self.prob), ws, wsn, Hcc.data_ptr(), gc.data_ptr(),
                                            Hpp.data_ptr(), gp.data_ptr(), cost))
        h6 = Hpp.cpu().numpy()[: lay.P]
        H3 = np.empty((lay.P, 3, 3))
        H3[:, 0, 0], H3[:, 0, 1], H3[:, 0, 2] = h6[:, 0], h6[:, 1], h6[:, 2]
        H3[:, 1, 0], H3[:, 1, 1], H3[:, 1, 2] = h6[:, 1], h6[:, 3], h6[:, 4]
        H3[:, 2, 0], H3[:, 2, 1], H3[:, 2, 2] = h6[:, 2], h6[:, 4], h6[:, 5]
        return (Hcc.cpu().numpy().reshape(lay.C, 6, 6), gc.cpu().numpy(), H3, gp.cpu().numpy()[: lay.P],
                float(cost[0]), float(cost[1]))

    def filter_tracks(self, max_reproj_error: float, min_tri_angle: float, select=None):
        """Post-BA track filter at the current state: pycolmap's ``obs.filter_points3D(max_reproj_error,
        min_tri_angle, point3D_ids)`` (mpsfm/sfm/mapper/base.py:700-711, 772-797) as one pass over the resident arrays.
        ``select`` [P] bool = the point3D_ids argument (None: every point).  Returns (obs_keep [N] bool in NATURAL
        observation order, pt_keep [P] bool, pt_error [P] (mean kept reprojection error, -1 for removed / unselected
        points), num_filtered)."""
        self._require_unsharded("filter_tracks")
        torch, lay = self.torch, self.layout
        ok = torch.empty((max(lay.N, 1),), dtype=torch.uint8, device=self.device)
        pk = torch.empty((max(lay.P, 1),), dtype=torch.uint8, device=self.device)
        pe = torch.empty((max(lay.P, 1),), dtype=torch.float64, device=self.device)
        sel = None
        if select is not None:
            sel_h = np.ascontiguousarray(np.asarray(select).astype(np.uint8).reshape(-1))
            if sel_h.shape[0] != lay.P:
                raise ValueError("select must have one entry per point")
            sel = torch.from_numpy(sel_h if lay.P else np.zeros(1, np.uint8)).to(self.device)
        nf = C.c_int64(0)
        ws, wsn = self._ws()
        self._check(self.lib.mpba_filter_tracks(self.handle, C.byref(self.prob), ws, wsn, float(max_reproj_error),
                                                float(min_tri_angle), sel.data_ptr() if sel is not None else None,
                                                ok.data_ptr(), pk.data_ptr(), pe.data_ptr(), C.byref(nf)))
        obs_keep = np.empty(lay.N, dtype=bool)
        obs_keep[lay.pm_src.astype(np.int64)] = ok.cpu().numpy()[: lay.N] != 0
        return obs_keep, pk.cpu().numpy()[: lay.P] != 0, pe.cpu().numpy()[: lay.P].copy(), int(nf.value)

    def point_covariances(self):
        """[P,3,3] covariance of every variable point conditioned on the poses (zeros for constant points)."""
        self._require_unsharded("point_covariances")
        torch, lay = self.torch, self.layout
        cov = torch.empty((max(lay.P, 1), 6), dtype=torch.float64, device=self.device)
        ws, wsn = self._ws()
        self._check(self.lib.mpba_point_covariances(self.handle, C.byref(self.prob), ws, wsn, cov.data_ptr()))
        c6 = cov.cpu().numpy()[: lay.P]
        out = np.empty((lay.P, 3, 3))
        out[:, 0, 0], out[:, 0, 1], out[:, 0, 2] = c6[:, 0], c6[:, 1], c6[:, 2]
        out[:, 1, 0], out[:, 1, 1], out[:, 1, 2] = c6[:, 1], c6[:, 3], c6[:, 4]
        out[:, 2, 0], out[:, 2, 1], out[:, 2, 2] = c6[:, 2], c6[:, 4], c6[:, 5]
        return out

    def reduced_system(self, radius: float = 1e4, options: Optional[MpbaOptions] = None):
        """(S blocks [B,6,6], rhs [C,6], y [C,6], pcg_iterations) of the damped normal equations."""
        self._require_unsharded("reduced_system")
        torch, lay = self.torch, self.layout
        S = torch.empty((lay.num_blocks, 36), dtype=torch.float64, device=self.device)
        rhs = torch.empty((lay.C, 6), dtype=torch.float64, device=self.device)
        y = torch.empty((lay.C, 6), dtype=torch.float64, device=self.device)
        it = C.c_int32()
        opt = options or default_options()
        ws, wsn = self._ws()
        self._check(self.lib.mpba_reduced_system(self.handle, C.byref(self.prob), ws, wsn, C.byref(opt), radius,
                                                 S.data_ptr(), rhs.data_ptr(), y.data_ptr(), C.byref(it)))
        return S.cpu().numpy().reshape(-1, 6, 6), rhs.cpu().numpy(), y.cpu().numpy(), int(it.value)

    def solve(self, options: Optional[MpbaOptions] = None) -> SolveSummary:
        """Run the LM loop; poses / points on the device are updated in place."""
        opt = options or default_options()
        cap = int(opt.max_num_iterations) + 2
        rows = (MpbaIteration * cap)()
        sm = MpbaSummary()
        sm.iterations = C.cast(rows, C.POINTER(MpbaIteration))
        sm.capacity = cap
        ws, wsn = self._ws()
        if self.shard is not None:
            self._check(self.lib.mpba_solve_sharded(self.handle, C.byref(self.prob), C.byref(self.shard), ws, wsn,
                                                    C.byref(opt), C.byref(sm)))
        else:
            self._check(self.lib.mpba_solve(self.handle, C.byref(self.prob), ws, wsn, C.byref(opt), C.byref(sm)))
        n = min(sm.num_iterations + 1, cap)
        its = []
        for k in range(n):
            r = rows[k]
            its.append({f: getattr(r, f) for f, _ in MpbaIteration._fields_ if f != "reserved"})
        return SolveSummary(
            initial_cost=sm.initial_cost, final_cost=sm.final_cost, fixed_cost=sm.fixed_cost,
            termination=TERMINATION.get(sm.termination, str(sm.termination)), message=sm.message.decode(),
            num_iterations=sm.num_iterations, num_successful_steps=sm.num_successful_steps,
            num_unsuccessful_steps=sm.num_unsuccessful_steps, total_pcg_iterations=int(sm.total_pcg_iterations),
            kernel_launches=int(sm.kernel_launches), solve_ms=sm.solve_ms, iterations=its,
            num_behind_camera=int(sm.num_behind_camera),
        )

    def kernel_times(self):
        """{kernel: (summed ms, launches)} of the last call made with options.profile = 1."""
        ms = (C.c_double * MPBA_NUM_KERNELS)()
        n = (C.c_int64 * MPBA_NUM_KERNELS)()
        self._check(self.lib.mpba_kernel_times(self.handle, ms, n))
        return {self.lib.mpba_kernel_name(i).decode(): (ms[i], int(n[i])) for i in range(MPBA_NUM_KERNELS) if n[i]}


def sample_maps(kps_per_map, maps, scales, device: int = 0):
    """GPU ``PriorUtils._data_at_kps`` for a whole bundle in one launch (mpba_sample_maps).

    kps_per_map: list of [n_m, 2] keypoints (original pixel scale), maps: list of [H_m, W_m] arrays (depth, variance
    or 0/1 validity), scales: list of (sx, sy).  Returns the list of [n_m] sampled values (float64)."""
    import torch

    lib = load_library()
    if not torch.cuda.is_available():
        raise EngineUnavailable("no CUDA device: mpsfm_b200 has no CPU fallback")
    if not (len(kps_per_map) == len(maps) == len(scales)):
        raise ValueError("kps_per_map, maps and scales must have one entry per map")
    M = len(maps)
    counts = [int(np.asarray(k).reshape(-1, 2).shape[0]) for k in kps_per_map]
    n = int(sum(counts))
    if n == 0:
        return [np.zeros(0) for _ in range(M)]
    kps = np.ascontiguousarray(np.concatenate([np.asarray(k, dtype=np.float64).reshape(-1, 2) for k in kps_per_map]))
    map_of = np.repeat(np.arange(M, dtype=np.int32), counts)
    flat = [np.ascontiguousarray(np.asarray(m, dtype=np.float64)) for m in maps]
    hw = np.array([m.shape[:2] for m in flat], dtype=np.int32)
    offset = np.zeros(M, dtype=np.int64)
    offset[1:] = np.cumsum([m.size for m in flat])[:-1]
    dev = torch.device("cuda", device)
    with torch.cuda.device(dev):
        t = {k: torch.from_numpy(v).to(dev) for k, v in dict(
            kps=kps, map_of=map_of, offset=offset, hw=hw, scale=np.asarray(scales, dtype=np.float64).reshape(M, 2),
            maps=np.concatenate([m.reshape(-1) for m in flat])).items()}
        out = torch.empty(n, dtype=torch.float64, device=dev)
        rc = lib.mpba_sample_maps(None, n, t["kps"].data_ptr(), t["map_of"].data_ptr(), M, t["offset"].data_ptr(),
                                  t["hw"].data_ptr(), t["scale"].data_ptr(), t["maps"].data_ptr(), out.data_ptr(),
                                  C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            raise EngineError(f"mpba_sample_maps failed with code {rc}")
        res = out.cpu().numpy()
    return np.split(res, np.cumsum(counts)[:-1])


PRIOR_TERMS = ("z", "mask", "w", "a", "logratio", "whitened")


def depth_prior_terms(img_of, row3, xyz, depth, variance, scale_filter_factor: float = 0.0, gross_outliers: bool = False,
                      loss_multiplier: float = 1.0, want=PRIOR_TERMS, device: int = 0, as_torch: bool = False):
    """Depth-prior terms of every sampled observation of a bundle in one launch (mpba_depth_prior_terms): z-depth,
    selection mask, ScaledLoss weight ``w``, loss scale ``a`` (bundle_adjustment.py:136-161), the log ratio of
    optimize_prior_shiftscale (:241) and the whitened log discrepancy of update_truncation_multiplier (:326-330).

    img_of [n] int32: image row of each sample; row3 [num_images, 4]: third row of cam_from_world.matrix();
    xyz [n, 3]; depth, variance [n].  Returns {name: array} for the names in ``want``."""
    import torch

    lib = load_library()
    if not torch.cuda.is_available():
        raise EngineUnavailable("no CUDA device: mpsfm_b200 has no CPU fallback")
    bad = [w for w in want if w not in PRIOR_TERMS]
    if bad:
        raise ValueError(f"unknown terms {bad}")
    dev = torch.device("cuda", device)

    def put(a, dt, shape):
        if isinstance(a, torch.Tensor):
            return a.to(device=dev, dtype=dt).reshape(shape).contiguous()
        return torch.from_numpy(np.ascontiguousarray(np.asarray(a), dtype={torch.float64: np.float64, torch.int32: np.int32}[dt]).reshape(shape)).to(dev)

    with torch.cuda.device(dev):
        t_img = put(img_of, torch.int32, (-1,))
        n = int(t_img.numel())
        t_row = put(row3, torch.float64, (-1, 4))
        t_xyz, t_d, t_v = put(xyz, torch.float64, (n, 3)), put(depth, torch.float64, (n,)), put(variance, torch.float64, (n,))
        out = {w: torch.empty(n, dtype=torch.uint8 if w == "mask" else torch.float64, device=dev) for w in want}
        ptr = lambda w: out[w].data_ptr() if w in out else None  # noqa: E731
        rc = lib.mpba_depth_prior_terms(None, n, t_img.data_ptr(), int(t_row.shape[0]), t_row.data_ptr(), t_xyz.data_ptr(),
                                        t_d.data_ptr(), t_v.data_ptr(), float(scale_filter_factor), int(bool(gross_outliers)),
                                        float(loss_multiplier), ptr("z"), ptr("mask"), ptr("w"), ptr("a"), ptr("logratio"),
                                        ptr("whitened"), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            raise EngineError(f"mpba_depth_prior_terms failed with code {rc}")
        if as_torch:
            return out
        res = {w: v.cpu().numpy() for w, v in out.items()}
    if "mask" in res:
        res["mask"] = res["mask"].astype(bool)
    return res


def segmented_median(values, seg_ptr, mad: bool = False, device: int = 0):
    """Exact ``np.median`` of every segment ``values[seg_ptr[s]:seg_ptr[s+1]]`` on the GPU (mpba_segmented_median);
    with ``mad`` also ``np.median(np.abs(x - median))`` per segment (fit_robust_gaussian_mad,
    bundle_adjustment.py:10-15).  values: numpy array or CUDA tensor (float64).  Returns median [S] (and mad [S])."""
    import torch

    lib = load_library()
    if not torch.cuda.is_available():
        raise EngineUnavailable("no CUDA device: mpsfm_b200 has no CPU fallback")
    dev = torch.device("cuda", device)
    seg = np.ascontiguousarray(np.asarray(seg_ptr, dtype=np.int64))
    if seg.ndim != 1 or seg.size < 1 or np.any(np.diff(seg) < 0) or seg[0] < 0:
        raise ValueError("seg_ptr must be a non-decreasing offset array")
    S = int(seg.size - 1)
    with torch.cuda.device(dev):
        if isinstance(values, torch.Tensor):
            t_val = values.to(device=dev, dtype=torch.float64).reshape(-1).contiguous()
        else:
            t_val = torch.from_numpy(np.ascontiguousarray(np.asarray(values, dtype=np.float64).reshape(-1))).to(dev)
        n = int(t_val.numel())
        if seg[-1] > n or n >= 2**31:
            raise ValueError("seg_ptr runs past the values")
        t_seg = torch.from_numpy(seg.astype(np.int32)).to(dev)
        nbytes = C.c_size_t(0)
        rc = lib.mpba_segmented_median_workspace_bytes(n, S, C.byref(nbytes))
        if rc != 0:
            raise EngineError(f"mpba_segmented_median_workspace_bytes failed with code {rc}")
        ws = torch.empty(int(nbytes.value) + 256, dtype=torch.uint8, device=dev)
        ws_ptr = (ws.data_ptr() + 255) & ~255
        med = torch.empty(S, dtype=torch.float64, device=dev)
        t_mad = torch.empty(S, dtype=torch.float64, device=dev) if mad else None
        rc = lib.mpba_segmented_median(None, n, t_val.data_ptr() if n else None, S, t_seg.data_ptr(), ws_ptr,
                                       int(nbytes.value), med.data_ptr(), t_mad.data_ptr() if mad else None,
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            raise EngineError(f"mpba_segmented_median failed with code {rc}")
        if mad:
            return med.cpu().numpy(), t_mad.cpu().numpy()
        return med.cpu().numpy()


def dense_reduced_system(layout: HostLayout, S_blocks: np.ndarray) -> np.ndarray:
    """Assemble the block-CSR S into a dense [6C,6C] matrix (tests, tiny problems only)."""
    Cn = layout.C
    S = np.zeros((6 * Cn, 6 * Cn))
    rows = np.repeat(np.arange(Cn), np.diff(layout.s_row_ptr))
    for b, (i, j) in enumerate(zip(rows, layout.s_col)):
        S[6 * i : 6 * i + 6, 6 * j : 6 * j + 6] = S_blocks[b]
    return S
