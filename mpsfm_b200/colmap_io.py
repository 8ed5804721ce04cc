"""COLMAP binary model files <-> struct-of-arrays (wire formats).

The reference persists a reconstruction with ``MpsfmReconstruction.write`` (mpsfm/sfm/scene/reconstruction/base.py:135-141:
``self.rec.write(rec_dir)``, pycolmap) and re-loads it with ``read`` (:143-145).  ``pycolmap.Reconstruction.write`` /
``read`` produce and consume COLMAP's three little-endian binary files; this module reads and writes the same bytes
directly into the arrays the engine works on, so that consecutive BA rounds (or an offline re-optimisation of a model on
disk) need no per-object pycolmap walk.  Layout of the files (COLMAP ``scene/reconstruction_io.cc``, the fork the
reference pins predates the rig / frame files):

``cameras.bin``   uint64 K; per camera: uint32 camera_id, int32 model_id, uint64 width, uint64 height,
                  double params[num_params(model_id)]
``images.bin``    uint64 C; per image: uint32 image_id, double qw qx qy qz, double tx ty tz, uint32 camera_id,
                  name bytes + NUL, uint64 M, M x (double x, double y, uint64 point3D_id)   (2^64-1: no point)
``points3D.bin``  uint64 P; per point: uint64 point3D_id, double xyz[3], uint8 rgb[3], double error, uint64 L,
                  L x (uint32 image_id, uint32 point2D_idx)

Quaternions are stored (w, x, y, z) on disk and (x, y, z, w) in the arrays (Eigen / pycolmap ``Rotation3d.quat``
order, like :class:`mpsfm_b200.problem.BundleProblem`).  ``depths.h5`` (:123-133) needs h5py, which this image does
not have: not built.
"""

from __future__ import annotations

import dataclasses
import os
import struct

import numpy as np

from .problem import CAMERA_NUM_PARAMS, FREE_ALL, MAX_CAMERA_PARAMS, BundleProblem, CameraModel, LossType

INVALID_POINT3D = np.uint64(18446744073709551615)  # kInvalidPoint3DId

_POINT2D = np.dtype([("xy", "<f8", 2), ("point3D_id", "<u8")])
_TRACK_EL = np.dtype([("image_id", "<u4"), ("point2D_idx", "<u4")])


@dataclasses.dataclass
class ColmapModel:
    """One COLMAP model as arrays; ragged parts (keypoints per image, track per point) are CSR."""

    cam_id: np.ndarray  # [K] uint32
    cam_model: np.ndarray  # [K] int32
    cam_width: np.ndarray  # [K] uint64
    cam_height: np.ndarray  # [K] uint64
    cam_params: np.ndarray  # [K,12] float64, the model's first num_params entries are used
    img_id: np.ndarray  # [C] uint32
    img_q: np.ndarray  # [C,4] float64, x y z w
    img_t: np.ndarray  # [C,3]
    img_camera_id: np.ndarray  # [C] uint32
    img_name: list  # [C] str
    kp_ptr: np.ndarray  # [C+1] int64
    kp_xy: np.ndarray  # [M,2] float64
    kp_point3D_id: np.ndarray  # [M] uint64
    pt_id: np.ndarray  # [P] uint64
    pt_xyz: np.ndarray  # [P,3]
    pt_rgb: np.ndarray  # [P,3] uint8
    pt_error: np.ndarray  # [P]
    track_ptr: np.ndarray  # [P+1] int64
    track_image_id: np.ndarray  # [T] uint32
    track_point2D_idx: np.ndarray  # [T] uint32


# ---------------------------------------------------------------------------------------------- read
def _read_cameras(path):
    buf = open(path, "rb").read()
    (K,) = struct.unpack_from("<Q", buf, 0)
    off = 8
    cid, model, w, h = np.zeros(K, np.uint32), np.zeros(K, np.int32), np.zeros(K, np.uint64), np.zeros(K, np.uint64)
    params = np.zeros((K, MAX_CAMERA_PARAMS))
    for k in range(K):
        cid[k], model[k], w[k], h[k] = struct.unpack_from("<IiQQ", buf, off)
        off += 24
        n = CAMERA_NUM_PARAMS[CameraModel(int(model[k]))]
        params[k, :n] = np.frombuffer(buf, "<f8", n, off)
        off += 8 * n
    if off != len(buf):
        raise ValueError(f"{path}: {len(buf) - off} trailing bytes")
    return cid, model, w, h, params


def _read_images(path):
    buf = open(path, "rb").read()
    (C,) = struct.unpack_from("<Q", buf, 0)
    off = 8
    iid, cam = np.zeros(C, np.uint32), np.zeros(C, np.uint32)
    q, t = np.zeros((C, 4)), np.zeros((C, 3))
    names, counts, parts = [], np.zeros(C, np.int64), []
    for k in range(C):
        iid[k], qw, qx, qy, qz, t[k, 0], t[k, 1], t[k, 2], cam[k] = struct.unpack_from("<I7dI", buf, off)
        q[k] = (qx, qy, qz, qw)
        off += 64
        end = buf.index(b"\0", off)
        names.append(buf[off:end].decode("utf-8"))
        off = end + 1
        (M,) = struct.unpack_from("<Q", buf, off)
        off += 8
        parts.append(np.frombuffer(buf, _POINT2D, M, off))
        counts[k] = M
        off += 24 * M
    if off != len(buf):
        raise ValueError(f"{path}: {len(buf) - off} trailing bytes")
    p2 = np.concatenate(parts) if parts else np.zeros(0, _POINT2D)
    ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    return iid, q, t, cam, names, ptr, np.ascontiguousarray(p2["xy"]), np.ascontiguousarray(p2["point3D_id"])


_POINT_HEAD = np.dtype([("id", "<u8"), ("xyz", "<f8", 3), ("rgb", "u1", 3), ("error", "<f8"), ("len", "<u8")])  # 51 bytes


def _segment_positions(starts, lengths):
    """Concatenation of the ranges [starts[k], starts[k] + lengths[k])."""
    total = int(lengths.sum())
    if total == 0:
        return np.zeros(0, np.int64)
    ends = np.cumsum(lengths)
    return np.arange(total, dtype=np.int64) + np.repeat(starts - (ends - lengths), lengths)


def _read_points(path):
    buf = open(path, "rb").read()
    (P,) = struct.unpack_from("<Q", buf, 0)
    # the only sequential part: where every variable-length record starts (track length at byte 43 of a record)
    offs = np.empty(P, dtype=np.int64)
    off = 8
    for k in range(P):
        offs[k] = off
        off += 51 + 8 * int.from_bytes(buf[off + 43:off + 51], "little")
    if off != len(buf):
        raise ValueError(f"{path}: {len(buf) - off} trailing bytes")
    u8 = np.frombuffer(buf, np.uint8)
    head = u8[offs[:, None] + np.arange(51)].copy().view(_POINT_HEAD).reshape(P)
    L = head["len"].astype(np.int64)
    pos = _segment_positions(offs + 51, 8 * L)
    tr = u8[pos].copy().view(_TRACK_EL)
    ptr = np.concatenate([[0], np.cumsum(L)]).astype(np.int64)
    return (np.ascontiguousarray(head["id"]), np.ascontiguousarray(head["xyz"]), np.ascontiguousarray(head["rgb"]),
            np.ascontiguousarray(head["error"]), ptr, np.ascontiguousarray(tr["image_id"]),
            np.ascontiguousarray(tr["point2D_idx"]))


def read_model(path) -> ColmapModel:
    """``pycolmap.Reconstruction.read(path)`` for the binary files, into arrays."""
    cams = _read_cameras(os.path.join(path, "cameras.bin"))
    imgs = _read_images(os.path.join(path, "images.bin"))
    pts = _read_points(os.path.join(path, "points3D.bin"))
    return ColmapModel(*cams, *imgs, *pts)


# --------------------------------------------------------------------------------------------- write
def write_model(path, m: ColmapModel) -> None:
    """``pycolmap.Reconstruction.write(path)``: the three binary files, byte for byte in COLMAP's layout."""
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "cameras.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(m.cam_id)))
        for k in range(len(m.cam_id)):
            n = CAMERA_NUM_PARAMS[CameraModel(int(m.cam_model[k]))]
            f.write(struct.pack("<IiQQ", int(m.cam_id[k]), int(m.cam_model[k]), int(m.cam_width[k]), int(m.cam_height[k])))
            f.write(np.asarray(m.cam_params[k, :n], "<f8").tobytes())
    with open(os.path.join(path, "images.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(m.img_id)))
        for k in range(len(m.img_id)):
            qx, qy, qz, qw = (float(v) for v in m.img_q[k])
            f.write(struct.pack("<I7dI", int(m.img_id[k]), qw, qx, qy, qz, *(float(v) for v in m.img_t[k]),
                                int(m.img_camera_id[k])))
            f.write(m.img_name[k].encode("utf-8") + b"\0")
            a, b = int(m.kp_ptr[k]), int(m.kp_ptr[k + 1])
            f.write(struct.pack("<Q", b - a))
            rec = np.empty(b - a, _POINT2D)
            rec["xy"], rec["point3D_id"] = m.kp_xy[a:b], m.kp_point3D_id[a:b]
            f.write(rec.tobytes())
    P = len(m.pt_id)
    L = np.diff(m.track_ptr).astype(np.int64)
    offs = 8 + np.concatenate([[0], np.cumsum(51 + 8 * L)]).astype(np.int64)
    out = np.zeros(int(offs[-1]), np.uint8)
    out[:8] = np.frombuffer(struct.pack("<Q", P), np.uint8)
    head = np.empty(P, _POINT_HEAD)
    head["id"], head["xyz"], head["rgb"], head["error"], head["len"] = m.pt_id, m.pt_xyz, m.pt_rgb, m.pt_error, L
    if P:
        out[offs[:-1, None] + np.arange(51)] = head.view(np.uint8).reshape(P, 51)
    tr = np.empty(int(m.track_ptr[-1]), _TRACK_EL)
    tr["image_id"], tr["point2D_idx"] = m.track_image_id, m.track_point2D_idx
    out[_segment_positions(offs[:-1] + 51, 8 * L)] = tr.view(np.uint8)
    with open(os.path.join(path, "points3D.bin"), "wb") as f:
        f.write(out.tobytes())


# ------------------------------------------------------------------------------------- to / from the engine's arrays
def to_bundle_problem(m: ColmapModel, kp_std: float = 1.0, reproj_loss=LossType.SOFT_L1, reproj_loss_scale: float = 1.5):
    """The reprojection part of a global bundle over the whole model (every image free except the gauge COLMAP fixes in
    ``Optimizer.__build_problem``, bundle_adjustment.py:111-122, is left to the caller's ``img_free_mask``): one
    observation per track element.  Returns (problem, obs_track_index) where ``obs_track_index[k]`` is the position
    of observation k in ``m.track_*`` -- what :func:`apply_filter` needs to edit the model."""
    C, P = len(m.img_id), len(m.pt_id)
    img_row = {int(i): k for k, i in enumerate(m.img_id)}
    cam_row = {int(c): k for k, c in enumerate(m.cam_id)}
    obs_pt = np.repeat(np.arange(P, dtype=np.int64), np.diff(m.track_ptr)).astype(np.int32)
    obs_img = np.array([img_row[int(i)] for i in m.track_image_id], dtype=np.int32).reshape(-1)
    slot = m.kp_ptr[obs_img] + m.track_point2D_idx.astype(np.int64)
    if np.any(m.track_point2D_idx.astype(np.int64) >= np.diff(m.kp_ptr)[obs_img]):
        raise ValueError("a track element points past the keypoints of its image")
    if not np.array_equal(m.kp_point3D_id[slot], m.pt_id[obs_pt]):
        raise ValueError("tracks and keypoint point3D ids disagree")
    none = np.zeros(0)
    problem = BundleProblem(
        img_q=m.img_q.copy(), img_t=m.img_t.copy(),
        img_camera=np.array([cam_row[int(c)] for c in m.img_camera_id], dtype=np.int32).reshape(-1),
        img_free_mask=np.full(C, FREE_ALL, dtype=np.uint8), cam_model=m.cam_model.astype(np.int32),
        cam_params=m.cam_params.copy(), pt_xyz=m.pt_xyz.copy(), pt_const=np.zeros(P, dtype=np.uint8), obs_img=obs_img,
        obs_pt=obs_pt, obs_xy=m.kp_xy[slot], depth_obs=none.astype(np.int32), depth_d=none, depth_w=none, depth_a=none,
        shift_scale=np.zeros((C, 2)), reproj_loss=LossType.parse(reproj_loss),
        reproj_loss_scale=float(reproj_loss_scale * kp_std), reproj_loss_weight=float(1 / kp_std**2),
        depth_loss=LossType.TRIVIAL).normalized()
    return problem, np.arange(obs_pt.shape[0], dtype=np.int64)


def update_state(m: ColmapModel, img_q, img_t, pt_xyz) -> None:
    """Write a solved state (``BundleAdjuster.state()``) back into the model's arrays, in place."""
    m.img_q[...] = img_q
    m.img_t[...] = img_t
    m.pt_xyz[...] = pt_xyz


def apply_filter(m: ColmapModel, obs_keep, pt_keep, pt_error) -> ColmapModel:
    """The edits ``ObservationManager.filter_points3D`` makes, from ``BundleAdjuster.filter_tracks`` of the problem
    :func:`to_bundle_problem` built: removed observations leave their tracks and reset their keypoint's point3D id,
    removed points disappear, surviving points get their mean reprojection error.  Returns a new model."""
    obs_keep, pt_keep = np.asarray(obs_keep, bool), np.asarray(pt_keep, bool)
    obs_pt = np.repeat(np.arange(len(m.pt_id), dtype=np.int64), np.diff(m.track_ptr))
    keep_el = obs_keep & pt_keep[obs_pt]
    img_row = {int(i): k for k, i in enumerate(m.img_id)}
    obs_img = np.array([img_row[int(i)] for i in m.track_image_id], dtype=np.int64).reshape(-1)
    slot = m.kp_ptr[obs_img] + m.track_point2D_idx.astype(np.int64)
    kp_pid = m.kp_point3D_id.copy()
    kp_pid[slot[~keep_el]] = INVALID_POINT3D
    counts = np.bincount(obs_pt[keep_el], minlength=len(m.pt_id))[pt_keep]
    err = np.where(np.asarray(pt_error) >= 0, pt_error, m.pt_error)[pt_keep]
    return dataclasses.replace(
        m, kp_point3D_id=kp_pid, pt_id=m.pt_id[pt_keep], pt_xyz=m.pt_xyz[pt_keep], pt_rgb=m.pt_rgb[pt_keep], pt_error=err,
        track_ptr=np.concatenate([[0], np.cumsum(counts)]).astype(np.int64), track_image_id=m.track_image_id[keep_el],
        track_point2D_idx=m.track_point2D_idx[keep_el])


# ----------------------------------------------------------------------------------------------
# depth cache next to the model: the reference's ``depths.h5`` (mpsfm/sfm/scene/reconstruction/base.py:122-133,
# written by ``write`` :135-141 beside ``rec/``)
# ----------------------------------------------------------------------------------------------
_DEPTH_GROUPS = (("depths", "data"), ("continuity", "continuity_mask"), ("scales", "scale"), ("pdepth", "data_prior"))


def cache_depths(registered_images, path) -> str:
    """``MpsfmReconstruction.cache_depths``: per registered image ``depths/<imid>`` (depth.data), ``continuity/<imid>``
    (continuity_mask), ``scales/<imid>`` (scale) and ``pdepth/<imid>`` (data_prior), each only where it is not None.

    With h5py importable the file is the reference's HDF5 (same dataset names, gzip on the maps).  h5py is not in this
    image ("unavailable on this box"): the twin is a compressed ``.npz`` holding the SAME keys -- ``path`` keeps its
    stem and gets the ``.npz`` suffix -- which ``read_cached_depths`` reads back either way.  Returns the path written."""
    path = str(path)
    entries = {}
    for imid, image in registered_images.items():
        depth = image.depth
        for group, attr in _DEPTH_GROUPS:
            value = getattr(depth, attr, None)
            if value is not None:
                entries[f"{group}/{imid}"] = np.asarray(value)
    try:
        import h5py
    except ImportError:
        h5py = None
    if h5py is not None:
        with h5py.File(path, "w") as f:
            for key, value in entries.items():
                if key.startswith("scales/"):
                    f.create_dataset(key, data=value)
                else:
                    f.create_dataset(key, data=value, compression="gzip")
        return path
    out = (path[: -len(".h5")] if path.endswith(".h5") else path) + ".npz"
    np.savez_compressed(out, **entries)
    return out


def read_cached_depths(path) -> dict:
    """{group: {image id: array}} of a file written by ``cache_depths`` (HDF5 or the ``.npz`` twin)."""
    path = str(path)
    out = {group: {} for group, _ in _DEPTH_GROUPS}
    if path.endswith(".npz"):
        with np.load(path) as z:
            for key in z.files:
                group, imid = key.split("/")
                out[group][int(imid)] = z[key]
        return out
    import h5py

    with h5py.File(path, "r") as f:
        for group in out:
            if group in f:
                for imid, ds in f[group].items():
                    out[group][int(imid)] = ds[()]
    return out
