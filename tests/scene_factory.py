"""Seeded fake-pycolmap scenes for the drop-in ``Optimizer`` tests.  TEST INFRASTRUCTURE.

Images on a ring of radius 40 look at a sphere of radius 8; every point3D lies on the camera-facing side
of the sphere and is observed by a run of consecutive images, so that the low-resolution monocular depth
maps (z-depth of the ray / sphere intersection, with a smooth multiplicative error, a variance map and a
validity mask with holes) are consistent with the structure, like the cached priors of the reference's
``local/example`` run.  Keypoints are quantised through float16 like the reference's correspondences
(/root/reference/mpsfm/sfm/scene/correspondences/base.py:123); every image also carries keypoints without a
point3D.
"""

from __future__ import annotations

import numpy as np

import fake_scene as fs

WIDTH, HEIGHT = 1600, 1200
F = 1200.0


def _look_at(centre, target, up=np.array([0.0, 1.0, 0.0])):
    z = target - centre
    z /= np.linalg.norm(z)
    x = np.cross(up, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    return np.stack([x, y, z], axis=0)  # rows = camera axes


def _quat_from_R(R):
    from mpsfm_b200.synthetic import quat_from_rotmat

    return quat_from_rotmat(R[None])[0]


def build_scene(seed=0, num_images=8, num_points=300, track_len=4, extra_kps=25, map_hw=(48, 64), sphere=8.0,
                pose_noise=(0.01, 0.05), point_noise=0.1, model="PINHOLE", holes=0.05, depth_sigma=0.05,
                outlier_patches=4):
    rng = np.random.Generator(np.random.PCG64(seed))
    rec = fs.Reconstruction()
    H, W = map_hw
    sx, sy = (W - 1) / (WIDTH - 1), (H - 1) / (HEIGHT - 1)
    params = {"PINHOLE": [F, F, WIDTH / 2, HEIGHT / 2], "SIMPLE_RADIAL": [F, WIDTH / 2, HEIGHT / 2, 0.02],
              "SIMPLE_PINHOLE": [F, WIDTH / 2, HEIGHT / 2]}[model]
    rec.rec.cameras[1] = fs.Camera(1, model, params, WIDTH, HEIGHT, sx=sx, sy=sy)

    theta = 2 * np.pi * np.arange(num_images) / max(num_images, 12)  # an arc: neighbouring images overlap
    centres = np.stack([40 * np.cos(theta), 1.0 * np.sin(2 * theta), 40 * np.sin(theta)], axis=1)
    R_true = np.stack([_look_at(c, rng.normal(0, 0.5, 3)) for c in centres])
    t_true = -np.einsum("nij,nj->ni", R_true, centres)

    # points on the camera-facing side of the sphere; tracks = runs of consecutive images
    starts = rng.integers(0, max(num_images - track_len + 1, 1), size=num_points)
    tracks = [list(range(s, min(s + track_len, num_images))) for s in starts]
    X_true = np.empty((num_points, 3))
    for p, tr in enumerate(tracks):
        d = centres[tr].mean(axis=0)
        d = d / np.linalg.norm(d) + rng.normal(0, 0.35, 3)
        X_true[p] = sphere * d / np.linalg.norm(d)

    def project(i, X):
        pc = X @ R_true[i].T + t_true[i]
        u, v = pc[:, 0] / pc[:, 2], pc[:, 1] / pc[:, 2]
        if model == "SIMPLE_RADIAL":
            g = 1 + params[3] * (u * u + v * v)
            return np.stack([F * u * g + params[1], F * v * g + params[2]], axis=1), pc[:, 2]
        return np.stack([F * u + WIDTH / 2, F * v + HEIGHT / 2], axis=1), pc[:, 2]

    # keypoints per image: observations first (shuffled with the extras afterwards)
    per_image = [[] for _ in range(num_images)]
    for p, tr in enumerate(tracks):
        for i in tr:
            per_image[i].append(p)
    obs_of_point = [[] for _ in range(num_points)]
    for i in range(num_images):
        pts = np.array(per_image[i], dtype=np.int64)
        xy, _ = project(i, X_true[pts]) if len(pts) else (np.zeros((0, 2)), None)
        xy = xy + rng.normal(0, 1.0, xy.shape)
        extra = np.stack([rng.uniform(0, WIDTH - 1, extra_kps), rng.uniform(0, HEIGHT - 1, extra_kps)], axis=1)
        kps = np.concatenate([xy, extra], axis=0)
        order = rng.permutation(kps.shape[0])
        kps = kps[order].astype(np.float16).astype(np.float64)
        inv = np.empty_like(order)
        inv[order] = np.arange(order.shape[0])
        # start pose = perturbed truth
        rv = rng.normal(0, pose_noise[0], 3)
        ang = np.linalg.norm(rv)
        k = rv / max(ang, 1e-300)
        K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        dR = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * (K @ K)
        R0 = dR @ R_true[i]
        c0 = centres[i] + rng.normal(0, pose_noise[1], 3)
        im = fs.Image(i + 1, 1, fs.Rigid3d(_quat_from_R(R0) * (1.0 + 0.01 * (i % 3)), -R0 @ c0), kps,
                      kp_std=float(1.0 + 0.1 * (i % 2)))
        rec.images[i + 1] = im
        for local, p in enumerate(pts):
            obs_of_point[p].append((i + 1, int(inv[local])))

    for p in range(num_points):
        rec.add_point3D(X_true[p] + rng.normal(0, point_noise, 3), obs_of_point[p])

    # monocular depth maps: z-depth of the ray / sphere intersection on the low-resolution grid
    gx, gy = np.meshgrid(np.arange(W) / sx, np.arange(H) / sy)
    rays_cam = np.stack([(gx - WIDTH / 2) / F, (gy - HEIGHT / 2) / F, np.ones_like(gx)], axis=-1)
    for i in range(num_images):
        o = centres[i]
        d = rays_cam @ R_true[i]  # world directions (rows of R are the camera axes)
        b = d @ o
        a = np.sum(d * d, axis=-1)
        disc = b * b - a * (o @ o - sphere * sphere)
        hit = disc > 0
        lam = np.where(hit, (-b - np.sqrt(np.where(hit, disc, 0.0))) / a, 0.0)  # camera z = lambda (ray z = 1)
        smooth = 1.0 + 0.03 * np.sin(gx / 300.0 + i) * np.cos(gy / 250.0)
        prior = np.where(hit, lam * smooth * np.exp(rng.normal(0, depth_sigma / 3, lam.shape)), 0.0)
        for _ in range(outlier_patches):  # grossly wrong regions: exercise the scale / gross-outlier filters
            y0, x0 = rng.integers(0, H - 6), rng.integers(0, W - 8)
            prior[y0 : y0 + 6, x0 : x0 + 8] *= rng.choice([rng.uniform(0.4, 0.62), rng.uniform(1.6, 2.3)])
        var = (depth_sigma * np.maximum(prior, 0.1)) ** 2
        valid = hit & (rng.random(lam.shape) > holes)
        im = rec.images[i + 1]
        cam = rec.rec.cameras[1]
        im.depth = fs.Depth(cam, np.where(prior > 0, prior, 0.1), var, valid,
                            np.array([p.xy for p in im.points2D]).reshape(-1, 2))
        im.depth.data = im.depth.data_prior * (1.0 + 0.01 * np.cos(gx / 400.0))  # "integrated" depth differs from the prior
    return rec


def attach_normals(rec):
    """Monocular normals for every image, derived from its own prior depth map the way a geometry network's output
    relates to it: n ~ d(p)/dx x d(p)/dy of the back-projected grid p = depth * ray, oriented towards the camera."""
    cam = rec.rec.cameras[1]
    for im in rec.images.values():
        depth = im.depth.data_prior
        H, W = depth.shape
        gx, gy = np.meshgrid(np.arange(W) / cam.sx, np.arange(H) / cam.sy)
        p = depth[..., None] * np.stack([(gx - WIDTH / 2) / F, (gy - HEIGHT / 2) / F, np.ones_like(gx)], axis=-1)
        n = np.cross(np.gradient(p, axis=1), np.gradient(p, axis=0))
        n /= np.maximum(np.linalg.norm(n, axis=-1, keepdims=True), 1e-300)
        n[n[..., 2] > 0] *= -1.0
        im.normals = fs.Normals(n)
    return rec
