"""Host-side description of one bundle-adjustment problem as struct-of-arrays.

This is the array form of what the reference hands to Ceres in
``Optimizer.__build_problem`` (/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:67-185):
one reprojection residual block per (image, point3D) observation
(``pycolmap.create_default_bundle_adjuster``, :104) and an optional log-depth prior residual
block per observation (``pycolmap.create_depth_bundle_adjuster``, :163-176).

Vocabulary follows the reference: *images* carry poses (``cam_from_world``), *cameras* carry
intrinsics (always constant, :92-94), *points3D* carry xyz, *observations* tie an image to a point.
"""

from __future__ import annotations

import dataclasses
import enum

import numpy as np


class CameraModel(enum.IntEnum):
    """COLMAP camera model ids (upstream ``colmap/sensor/models.h``)."""

    SIMPLE_PINHOLE = 0
    PINHOLE = 1
    SIMPLE_RADIAL = 2
    RADIAL = 3
    OPENCV = 4
    OPENCV_FISHEYE = 5
    FULL_OPENCV = 6
    FOV = 7
    SIMPLE_RADIAL_FISHEYE = 8
    RADIAL_FISHEYE = 9
    THIN_PRISM_FISHEYE = 10


CAMERA_NUM_PARAMS = {
    CameraModel.SIMPLE_PINHOLE: 3,
    CameraModel.PINHOLE: 4,
    CameraModel.SIMPLE_RADIAL: 4,
    CameraModel.RADIAL: 5,
    CameraModel.OPENCV: 8,
    CameraModel.OPENCV_FISHEYE: 8,
    CameraModel.FULL_OPENCV: 12,
    CameraModel.FOV: 5,
    CameraModel.SIMPLE_RADIAL_FISHEYE: 4,
    CameraModel.RADIAL_FISHEYE: 5,
    CameraModel.THIN_PRISM_FISHEYE: 12,
}
MAX_CAMERA_PARAMS = 12


class LossType(enum.IntEnum):
    """Ceres loss functions reachable from the reference (bundle_adjustment.py:44-48, :100)
    plus Huber, which BASELINE.json's north_star names."""

    TRIVIAL = 0
    SOFT_L1 = 1
    CAUCHY = 2
    HUBER = 3

    @classmethod
    def parse(cls, name) -> "LossType":
        if isinstance(name, cls):
            return name
        if isinstance(name, (int, np.integer)):
            return cls(int(name))
        key = str(name).split(".")[-1].upper().replace("SOFTL1", "SOFT_L1")
        return cls[key]


# bit j of img_free_mask <-> tangent coordinate j of an image: 0..2 rotation (half-angle
# vector of EigenQuaternionManifold), 3..5 translation.
FREE_ALL = 0b111111
FREE_NONE = 0
FREE_FIX_TX = 0b110111  # SubsetManifold(3, [0]) on the translation (bundle_adjustment.py:119-121)


@dataclasses.dataclass
class BundleProblem:
    """One BA problem, natural (caller) observation order.

    Shapes: C images, K cameras, P points3D, N observations, N_d depth-prior rows.
    All floating arrays are float64, all index arrays int32.
    """

    # images (poses). Quaternion storage order is Eigen's x,y,z,w like pycolmap's Rotation3d.quat.
    img_q: np.ndarray  # [C,4]
    img_t: np.ndarray  # [C,3]
    img_camera: np.ndarray  # [C] -> row of cam_model/cam_params
    img_free_mask: np.ndarray  # [C] uint8, see FREE_*
    # cameras (intrinsics; constant in every reference call site)
    cam_model: np.ndarray  # [K] int32 CameraModel
    cam_params: np.ndarray  # [K,12]
    # points3D
    pt_xyz: np.ndarray  # [P,3]
    pt_const: np.ndarray  # [P] uint8, 1 = constant parameter block
    # observations (reprojection residual blocks)
    obs_img: np.ndarray  # [N]
    obs_pt: np.ndarray  # [N]
    obs_xy: np.ndarray  # [N,2]
    # depth-prior residual blocks: row j refers to observation depth_obs[j]
    depth_obs: np.ndarray  # [N_d] int32 index into observations
    depth_d: np.ndarray  # [N_d] prior depth d (>0)
    depth_w: np.ndarray  # [N_d] ScaledLoss magnitude  w = d^2 / clip(v, 1e-6)   (:161)
    depth_a: np.ndarray  # [N_d] loss scale            a = m * sqrt(v) / d       (:160)
    # per-image (shift b, log-scale s) of the depth prior; constant blocks (:83, :178-182)
    shift_scale: np.ndarray  # [C,2]
    # losses
    reproj_loss: LossType = LossType.SOFT_L1
    reproj_loss_scale: float = 1.5  # reproj_loss_scale * kp_std (:101)
    reproj_loss_weight: float = 1.0  # loss_function_magnitude = 1/kp_std^2 (:99)
    depth_loss: LossType = LossType.CAUCHY
    # EXTENSION (no counterpart in the reference): point-to-plane normal-prior rows.  Row j refers to
    # observation plane_obs[j]:  r = m_j . p_cam - c_j  with  m = sqrt(w_n) n  (n: prior unit normal in the camera
    # frame) and  c = m . q  (q: prior surface point = prior depth x the keypoint's ray).  Loss: depth_loss with the
    # global scale plane_loss_scale, weight 1.  None / empty: no such rows.
    plane_obs: np.ndarray | None = None  # [N_n] int32 index into observations
    plane_m: np.ndarray | None = None  # [N_n,3]
    plane_c: np.ndarray | None = None  # [N_n]
    plane_loss_scale: float = 1.0

    # ------------------------------------------------------------------
    @property
    def num_images(self) -> int:
        return int(self.img_q.shape[0])

    @property
    def num_points(self) -> int:
        return int(self.pt_xyz.shape[0])

    @property
    def num_obs(self) -> int:
        return int(self.obs_img.shape[0])

    @property
    def num_depth(self) -> int:
        return int(self.depth_obs.shape[0])

    @property
    def num_plane(self) -> int:
        return 0 if self.plane_obs is None else int(np.asarray(self.plane_obs).shape[0])

    def normalized(self) -> "BundleProblem":
        """Return a copy with canonical dtypes/contiguity and unit quaternions.

        COLMAP normalises each quaternion in place before adding it to the Ceres problem
        (upstream ``DefaultBundleAdjuster::AddImageToProblem``)."""
        f = lambda a, s: np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(s))  # noqa: E731
        i = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.int32).reshape(-1))  # noqa: E731
        q = f(self.img_q, (-1, 4)).copy()
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        params = np.zeros((np.asarray(self.cam_params).shape[0], MAX_CAMERA_PARAMS))
        cp = np.asarray(self.cam_params, dtype=np.float64)
        params[:, : cp.shape[1]] = cp
        return dataclasses.replace(
            self,
            img_q=q,
            img_t=f(self.img_t, (-1, 3)).copy(),
            img_camera=i(self.img_camera),
            img_free_mask=np.ascontiguousarray(np.asarray(self.img_free_mask, dtype=np.uint8).reshape(-1)),
            cam_model=i(self.cam_model),
            cam_params=params,
            pt_xyz=f(self.pt_xyz, (-1, 3)).copy(),
            pt_const=np.ascontiguousarray(np.asarray(self.pt_const, dtype=np.uint8).reshape(-1)),
            obs_img=i(self.obs_img),
            obs_pt=i(self.obs_pt),
            obs_xy=f(self.obs_xy, (-1, 2)),
            depth_obs=i(self.depth_obs),
            depth_d=f(self.depth_d, (-1,)),
            depth_w=f(self.depth_w, (-1,)),
            depth_a=f(self.depth_a, (-1,)),
            shift_scale=f(self.shift_scale, (-1, 2)),
            plane_obs=None if self.plane_obs is None else i(self.plane_obs),
            plane_m=None if self.plane_m is None else f(self.plane_m, (-1, 3)),
            plane_c=None if self.plane_c is None else f(self.plane_c, (-1,)),
            reproj_loss=LossType.parse(self.reproj_loss),
            depth_loss=LossType.parse(self.depth_loss),
        )

    def validate(self) -> None:
        C, P, N = self.num_images, self.num_points, self.num_obs
        if self.img_t.shape != (C, 3) or self.img_camera.shape != (C,) or self.img_free_mask.shape != (C,):
            raise ValueError("image arrays disagree on the number of images")
        if self.shift_scale.shape != (C, 2):
            raise ValueError("shift_scale must be [C,2]")
        if self.pt_const.shape != (P,):
            raise ValueError("pt_const must be [P]")
        if self.obs_pt.shape != (N,) or self.obs_xy.shape != (N, 2):
            raise ValueError("observation arrays disagree on N")
        if N and (self.obs_img.min() < 0 or self.obs_img.max() >= C):
            raise ValueError("obs_img out of range")
        if N and (self.obs_pt.min() < 0 or self.obs_pt.max() >= P):
            raise ValueError("obs_pt out of range")
        if C and (self.img_camera.min() < 0 or self.img_camera.max() >= self.cam_model.shape[0]):
            raise ValueError("img_camera out of range")
        nd = self.num_depth
        if not (self.depth_d.shape == self.depth_w.shape == self.depth_a.shape == (nd,)):
            raise ValueError("depth prior arrays disagree on N_d")
        if nd:
            if self.depth_obs.min() < 0 or self.depth_obs.max() >= N:
                raise ValueError("depth_obs out of range")
            if int(np.bincount(self.depth_obs, minlength=N).max()) > 1:
                raise ValueError("at most one depth prior per observation")
            if not np.all(self.depth_d > 0):
                raise ValueError("depth priors must be > 0 (bundle_adjustment.py:139)")
        nn = self.num_plane
        if nn:
            if self.plane_m is None or self.plane_c is None or self.plane_m.shape != (nn, 3) or self.plane_c.shape != (nn,):
                raise ValueError("normal-prior arrays disagree on N_n")
            if self.plane_obs.min() < 0 or self.plane_obs.max() >= N:
                raise ValueError("plane_obs out of range")
            if int(np.bincount(self.plane_obs, minlength=N).max()) > 1:
                raise ValueError("at most one normal prior per observation")
            if not self.plane_loss_scale > 0:
                raise ValueError("plane_loss_scale must be > 0")
        for m in np.unique(self.cam_model):
            if int(m) not in CameraModel._value2member_map_:
                raise ValueError(f"unknown camera model id {m}")

    def copy(self) -> "BundleProblem":
        return dataclasses.replace(
            self, **{f.name: getattr(self, f.name).copy() for f in dataclasses.fields(self)
                     if isinstance(getattr(self, f.name), np.ndarray)}
        )
