"""Frozen synthetic inputs (numpy, fp32) for the parity tests and bench.py.

Camera conventions are the reference's: `getWorld2View2` / `getProjectionMatrix(znear=0.01,
zfar=100)` (utils/graphics_utils.py:38-71) and `Camera` (scene/cameras.py:86-89): the matrices
handed to the rasterizer are the TRANSPOSED world->view and (proj @ world->view) matrices, i.e.
flat memory index m[4*col+row] of the column-vector matrices.

Scene families follow SURVEY.md section 8(d):
  * `random_init_scene`  -- the reference's Blender random init (scene/dataset_readers.py:290-298,
    scene/gaussian_model.py:149-176): uniform points in [-1.3,1.3]^3, isotropic 3-NN scale,
    opacity 0.1, rot (1,0,0,0), DC-only SH.
  * `trained_like_scene` -- a trained-scene-shaped cloud: 60% central blob + 40% background shell,
    log-normal anisotropic scales, random rotations, bimodal opacity, full SH-3.
"""
import math

import numpy as np

C0 = 0.28209479177387814


def projection_matrix(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:51-71 (column-vector form, before the reference's transpose)."""
    t_y, t_x = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = t_y * znear, t_x * znear
    P = np.zeros((4, 4), np.float32)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def look_at_w2c(eye, target, up=(0.0, 0.0, 1.0)):
    """World->camera 4x4 (x right, y down, z forward: the COLMAP convention 3DGS uses)."""
    eye = np.asarray(eye, np.float64)
    f = np.asarray(target, np.float64) - eye
    f /= np.linalg.norm(f)
    r = np.cross(f, np.asarray(up, np.float64))
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    R = np.stack([r, d, f])            # rows = camera axes in world coordinates
    w2c = np.eye(4)
    w2c[:3, :3] = R
    w2c[:3, 3] = -R @ eye
    return w2c.astype(np.float32)


class Camera:
    """The fields gaussian_renderer/__init__.py:36-50 reads from a reference Camera."""

    def __init__(self, w2c, width, height, fovx, fovy=None, znear=0.01, zfar=100.0):
        self.image_width, self.image_height = int(width), int(height)
        self.FoVx = float(fovx)
        if fovy is None:                      # focal2fov(fov2focal(fovx, W), H)
            focal = width / (2 * math.tan(fovx / 2))
            fovy = 2 * math.atan(height / (2 * focal))
        self.FoVy = float(fovy)
        w2c = np.asarray(w2c, np.float32)
        proj = projection_matrix(znear, zfar, self.FoVx, self.FoVy)
        self.world_view_transform = np.ascontiguousarray(w2c.T)                         # cameras.py:86
        self.projection_matrix = np.ascontiguousarray(proj.T)                           # cameras.py:87
        self.full_proj_transform = (self.world_view_transform @ self.projection_matrix).astype(np.float32)  # :88
        self.camera_center = np.linalg.inv(self.world_view_transform.astype(np.float64))[3, :3].astype(np.float32)  # :89
        self.tanfovx = math.tan(self.FoVx * 0.5)
        self.tanfovy = math.tan(self.FoVy * 0.5)


def orbit_cameras(n, width, height, seed, fovx=0.9, radius=(4.0, 5.0), jitter=0.3):
    rng = np.random.default_rng(seed)
    cams = []
    for k in range(n):
        ang = 2 * math.pi * k / max(n, 1) * 3.0 + rng.uniform(0, 0.2)
        rad = rng.uniform(*radius)
        h = 0.6 * math.sin(2 * math.pi * k / max(n, 1)) + rng.normal(0, 0.1)
        eye = np.array([rad * math.cos(ang), rad * math.sin(ang), h])
        target = rng.normal(0, jitter, 3)
        cams.append(Camera(look_at_w2c(eye, target), width, height, fovx))
    return cams


def blender_camera(transform_matrix, width, height, fovx):
    """readCamerasFromTransforms (scene/dataset_readers.py:239-247): flip y/z of the NeRF c2w."""
    c2w = np.array(transform_matrix, np.float64)
    c2w[:3, 1:3] *= -1
    w2c = np.linalg.inv(c2w)
    return Camera(w2c.astype(np.float32), width, height, fovx)


# The hard-coded chair camera of the fork's host renderer (GS-IRON/npu-1/base.cpp:453-457 uses a
# look-at of this shape; fx=fy=1111.1 at 800x800 -> FoVx = 2 atan(400/1111.1)).
def chair_camera(width=800, height=800):
    fovx = 2 * math.atan(width / (2 * 1111.1))
    return Camera(look_at_w2c([2.5, -2.8, 1.6], [0, 0, 0.2]), width, height, fovx)


def _knn3_mean_sq_dist(xyz):
    """distCUDA2 (submodules/simple-knn/simple_knn.cu:154-183): mean squared distance to the 3 NN."""
    from scipy.spatial import cKDTree
    d, _ = cKDTree(xyz).query(xyz, k=4)
    return (d[:, 1:] ** 2).mean(axis=1).astype(np.float32)


def random_init_scene(P=100_000, seed=1001, sh_degree=3):
    rng = np.random.default_rng(seed)
    xyz = (rng.random((P, 3)) * 2.6 - 1.3).astype(np.float32)
    dist2 = np.maximum(_knn3_mean_sq_dist(xyz), 1e-7)
    scales = np.repeat(np.sqrt(dist2)[:, None], 3, axis=1).astype(np.float32)   # exp(log(sqrt(dist2)))
    rot = np.zeros((P, 4), np.float32); rot[:, 0] = 1
    opac = np.full((P, 1), 0.1, np.float32)
    M = 16
    shs = np.zeros((P, M, 3), np.float32)
    rgb = 0.5 + (rng.random((P, 3)).astype(np.float32) - 0.5) * 0.2
    shs[:, 0, :] = (rgb - 0.5) / C0
    return dict(means3D=xyz, scales=scales, rotations=rot, opacities=opac, shs=shs, sh_degree=sh_degree)


def trained_like_scene(P, seed, s0=0.012, extent=1.0, sh_degree=3):
    rng = np.random.default_rng(seed)
    n_blob = int(0.6 * P)
    blob = rng.normal(0, 1, (n_blob, 3)) * np.array([2.0, 2.0, 0.8])
    blob = np.clip(blob, -6, 6)
    n_sh = P - n_blob
    d = rng.normal(0, 1, (n_sh, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = rng.uniform(4.0, 12.0, (n_sh, 1))
    xyz = (np.concatenate([blob, d * r]) * extent).astype(np.float32)
    perm = rng.permutation(P)
    xyz = xyz[perm]
    scales = np.exp(rng.normal(math.log(s0), 0.8, (P, 3))).astype(np.float32) * extent
    q = rng.normal(0, 1, (P, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = (1.0 / (1.0 + np.exp(-rng.normal(0, 2.5, (P, 1))))).astype(np.float32)
    shs = np.empty((P, 16, 3), np.float32)
    shs[:, 0] = rng.normal(0, 1, (P, 3))
    k = 1
    for l in (1, 2, 3):
        n = 2 * l + 1
        shs[:, k:k + n] = rng.normal(0, 0.25 / l, (P, n, 3))
        k += n
    return dict(means3D=xyz, scales=scales, rotations=q.astype(np.float32), opacities=opac, shs=shs,
                sh_degree=sh_degree)


def upstream_grads(H, W, seed, dense=True):
    """dL/dcolor [3,H,W] and dL/dinvdepth [1,H,W] for the backward tests (dense U(-1,1))."""
    rng = np.random.default_rng(seed)
    if dense:
        return (rng.uniform(-1, 1, (3, H, W)).astype(np.float32),
                rng.uniform(-1, 1, (1, H, W)).astype(np.float32))
    t = rng.random((3, H, W)).astype(np.float32)
    return t, rng.random((1, H, W)).astype(np.float32)
