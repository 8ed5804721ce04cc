"""A pycolmap-shaped scene without pycolmap.  TEST INFRASTRUCTURE.

Duck-types exactly the attributes the reference's ``Optimizer`` reads and writes
(/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py; list in SURVEY.md section 8 row b):

  mpsfm_rec.images[imid]            .camera_id .cam_from_world.{rotation.quat (xyzw), translation, matrix()}
                                    .get_observation_point2D_idxs() .keypoint_coordinates(idxs)
                                    .point3D_ids(idxs) .kp_std .depth
  image.depth                       .activated .scale .valid_at_kps .data_at_kps .data_prior_at_kps
                                    .uncertainty_update
  mpsfm_rec.rec.cameras[camid]      .model .params .calibration_matrix()
  mpsfm_rec.points3D[id]            .xyz .track.length() .track.elements[k].{image_id, point2D_idx}
  mpsfm_rec.point3D_coordinates(ids), .project_image_3d_points(imid, ids), .reg_image_ids(), .point_covs.data

Parameter blocks are numpy arrays that the optimizer must update IN PLACE (Ceres parameter blocks alias
pycolmap memory in the reference), which is what the in-place contract tests check.

The depth-map sampling (``_data_at_kps``) and the projection helper are pluggable: the golden-vector
generator (tests/golden/make_golden.py, runs only where /root/reference exists) plugs in the reference's
own ``PriorUtils`` / ``project3D_colmap``; everywhere else the numpy restatements of ``oracle/`` are used.
"""

from __future__ import annotations

import numpy as np

INVALID_POINT3D = 18446744073709551615  # pycolmap's kInvalidPoint3DId


class Rotation3d:
    def __init__(self, quat_xyzw):
        self.quat = np.array(quat_xyzw, dtype=np.float64)

    def matrix(self):
        x, y, z, w = self.quat
        return np.array([
            [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
            [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
            [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
        ])

    def normalize(self):
        self.quat /= np.linalg.norm(self.quat)


class Rigid3d:
    def __init__(self, quat_xyzw, translation):
        self.rotation = Rotation3d(quat_xyzw)
        self.translation = np.array(translation, dtype=np.float64)

    def matrix(self):
        return np.concatenate([self.rotation.matrix(), self.translation[:, None]], axis=1)

    def __mul__(self, points):
        points = np.asarray(points, dtype=np.float64)
        return points @ self.rotation.matrix().T + self.translation


class CameraModelId:
    """Stand-in for pycolmap.CameraModelId: has ``.name`` and ``.value``."""

    NAMES = ["SIMPLE_PINHOLE", "PINHOLE", "SIMPLE_RADIAL", "RADIAL", "OPENCV", "OPENCV_FISHEYE", "FULL_OPENCV",
             "FOV", "SIMPLE_RADIAL_FISHEYE", "RADIAL_FISHEYE", "THIN_PRISM_FISHEYE"]

    def __init__(self, name):
        self.name = name
        self.value = self.NAMES.index(name)


class Camera:
    """pycolmap.Camera plus the mpsfm wrapper's ``sx``/``sy`` (keypoint -> depth-map pixel scale,
    /root/reference/mpsfm/sfm/scene/camera.py)."""

    def __init__(self, camera_id, model, params, width, height, sx=1.0, sy=1.0):
        self.camera_id = camera_id
        self.model = CameraModelId(model)
        self.params = np.array(params, dtype=np.float64)
        self.width, self.height = width, height
        self.sx, self.sy = sx, sy

    def calibration_matrix(self):
        p = self.params
        if self.model.name in ("SIMPLE_PINHOLE", "SIMPLE_RADIAL", "RADIAL", "SIMPLE_RADIAL_FISHEYE", "RADIAL_FISHEYE"):
            fx = fy = p[0]
            cx, cy = p[1], p[2]
        else:
            fx, fy, cx, cy = p[0], p[1], p[2], p[3]
        return np.array([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]])


class TrackElement:
    def __init__(self, image_id, point2D_idx):
        self.image_id, self.point2D_idx = image_id, point2D_idx


class Track:
    def __init__(self):
        self.elements = []

    def length(self):
        return len(self.elements)


class Point3D:
    def __init__(self, xyz):
        self.xyz = np.array(xyz, dtype=np.float64)
        self.track = Track()


class Point2D:
    def __init__(self, xy, point3D_id=INVALID_POINT3D):
        self.xy = np.array(xy, dtype=np.float64)
        self.point3D_id = point3D_id

    def has_point3D(self):
        return self.point3D_id != INVALID_POINT3D


def numpy_data_at_kps(depth, kps, data):
    from oracle.sampling import data_at_kps

    return data_at_kps(np.asarray(kps, dtype=np.float64), np.asarray(data, dtype=np.float64), depth.camera.sx, depth.camera.sy)


class Depth:
    """The attributes of /root/reference/mpsfm/sfm/scene/image/depth.py the optimizer touches."""

    sampler = staticmethod(numpy_data_at_kps)  # (depth, kps, map) -> values; swapped by the golden generator

    def __init__(self, camera, data_prior, uncertainty, valid, kps):
        self.camera = camera
        self.data_prior = np.array(data_prior, dtype=np.float64)
        self.data = self.data_prior.copy()  # "update" depth (refined by integration in the reference)
        self.uncertainty = np.array(uncertainty, dtype=np.float64)
        self.valid = np.array(valid, dtype=bool)
        self.scale, self.shift = 1, 0
        self.activated = True
        self.uncertainty_update = self.uncertainty_at_kps(kps)

    def _data_at_kps(self, kps, data):
        return type(self).sampler(self, kps, data)

    def data_prior_at_kps(self, kps):
        return self._data_at_kps(kps, self.data_prior)

    def data_at_kps(self, kps):
        return self._data_at_kps(kps, self.data)

    def uncertainty_at_kps(self, kps):
        return self._data_at_kps(kps, self.uncertainty)

    def valid_at_kps(self, kps):
        return self._data_at_kps(kps, self.valid) == 1


class Normals:
    """The attribute of /root/reference/mpsfm/sfm/scene/image/normals.py the normal-prior extension reads: ``data``,
    the [H,W,3] unit normals of the monocular network in the camera frame, on the depth map's grid."""

    def __init__(self, data):
        self.data = np.array(data, dtype=np.float64)


class Image:
    def __init__(self, image_id, camera_id, cam_from_world, keypoints, kp_std=1.0, name=None):
        self.image_id, self.camera_id = image_id, camera_id
        self.cam_from_world = cam_from_world
        self.points2D = [Point2D(xy) for xy in np.asarray(keypoints, dtype=np.float64)]
        self.kp_std = kp_std
        self.name = name or f"im{image_id}.jpg"
        self.depth = None
        self.normals = None

    # --- fork API used by the reference (bundle_adjustment.py:52-54, :126-136)
    def get_observation_point2D_idxs(self):
        return [i for i, p in enumerate(self.points2D) if p.has_point3D()]

    def keypoint_coordinates(self, idxs):
        return np.array([self.points2D[i].xy for i in idxs], dtype=np.float64).reshape(-1, 2)

    def point3D_ids(self, idxs=None):
        if idxs is None:
            idxs = range(len(self.points2D))
        return [self.points2D[i].point3D_id for i in idxs]


class PointCovs:
    def __init__(self):
        self.data = {}


class _Rec:
    def __init__(self):
        self.cameras = {}


def numpy_project(image, camera, points3D):
    from oracle.sampling import project3D_colmap

    return project3D_colmap(image.cam_from_world.matrix(), camera.calibration_matrix(), np.asarray(points3D, dtype=np.float64))


class ObservationManager:
    """The two pycolmap ObservationManager edits the track filter makes (``rec.obs`` in MpsfmReconstruction)."""

    def __init__(self, rec):
        self._rec = rec

    def delete_observation(self, image_id, point2D_idx):
        p2d = self._rec.images[image_id].points2D[point2D_idx]
        track = self._rec.points3D[p2d.point3D_id].track
        track.elements = [e for e in track.elements if (e.image_id, e.point2D_idx) != (image_id, point2D_idx)]
        p2d.point3D_id = INVALID_POINT3D

    def delete_point3D(self, point3D_id):
        for e in self._rec.points3D[point3D_id].track.elements:
            self._rec.images[e.image_id].points2D[e.point2D_idx].point3D_id = INVALID_POINT3D
        del self._rec.points3D[point3D_id]


class Reconstruction:
    """Stand-in for MpsfmReconstruction (+ the wrapped pycolmap.Reconstruction as ``.rec``)."""

    projector = staticmethod(numpy_project)  # swapped for the reference's project3D_colmap by the golden generator

    def __init__(self):
        self.images = {}
        self.points3D = {}
        self.rec = _Rec()
        self.point_covs = PointCovs()
        self._next_point3D_id = 1
        self.obs = ObservationManager(self)

    @property
    def cameras(self):
        return self.rec.cameras

    def reg_image_ids(self):
        return list(self.images.keys())

    def add_point3D(self, xyz, observations):
        """observations: list of (image_id, point2D_idx)."""
        pid = self._next_point3D_id
        self._next_point3D_id += 1
        pt = Point3D(xyz)
        for imid, idx in observations:
            pt.track.elements.append(TrackElement(imid, idx))
            self.images[imid].points2D[idx].point3D_id = pid
        self.points3D[pid] = pt
        return pid

    def point3D_coordinates(self, ids):
        return np.array([self.points3D[i].xyz for i in ids], dtype=np.float64).reshape(-1, 3)

    def project_image_3d_points(self, imid, pts3dids=None):
        # /root/reference/mpsfm/sfm/scene/reconstruction/mixins/points3D_utils.py:9-25
        image = self.images[imid]
        if pts3dids is None:
            pts2dids = image.get_observation_point2D_idxs()
            pts3dids = image.point3D_ids(pts2dids)
            if len(pts3dids) == 0:
                return None, None, None, None, False
        else:
            pts2dids = None
        points3D = self.point3D_coordinates(pts3dids)
        camera = self.rec.cameras[image.camera_id]
        kps, depth = type(self).projector(image, camera, points3D)
        return pts2dids, pts3dids, kps, depth, True
