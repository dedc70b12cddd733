"""Camera loading for real scenes (the caller side of the rasterizer's camera arguments).

Host mirror, in Python like the reference, of
  /root/reference/scene/colmap_loader.py:43-53,156-271   COLMAP cameras/images .bin and .txt, qvec2rotmat
  /root/reference/scene/dataset_readers.py:48-118        getNerfppNorm, readColmapCameras (PINHOLE / SIMPLE_PINHOLE)
  /root/reference/scene/dataset_readers.py:186-198       LLFF hold-out split, cameras sorted by image name
  /root/reference/scene/dataset_readers.py:228-271       readCamerasFromTransforms (Blender / NeRF-synthetic JSON)
  /root/reference/utils/graphics_utils.py:38-76          getWorld2View2, getProjectionMatrix, fov2focal, focal2fov
  /root/reference/scene/cameras.py:80-89                 the four per-camera tensors the renderer reads
Only the geometry is loaded: images are the training targets, not inputs of the path (the image SIZE of a Blender
frame is read from the PNG header, so PIL is not needed).  Files are parsed from one buffer with struct.unpack_from;
the 2D observations of images.bin are skipped by their byte count, not materialised.
"""
import json
import math
import os
import struct
from typing import NamedTuple

import numpy as np

from .synth import Camera

# COLMAP camera models: id -> (name, number of parameters)   (colmap_loader.py:24-36)
COLMAP_MODELS = {0: ("SIMPLE_PINHOLE", 3), 1: ("PINHOLE", 4), 2: ("SIMPLE_RADIAL", 4), 3: ("RADIAL", 5), 4: ("OPENCV", 8),
                 5: ("OPENCV_FISHEYE", 8), 6: ("FULL_OPENCV", 12), 7: ("FOV", 5), 8: ("SIMPLE_RADIAL_FISHEYE", 4),
                 9: ("RADIAL_FISHEYE", 5), 10: ("THIN_PRISM_FISHEYE", 12)}


class Intrinsics(NamedTuple):
    id: int
    model: str
    width: int
    height: int
    params: np.ndarray


class Extrinsics(NamedTuple):
    id: int
    qvec: np.ndarray
    tvec: np.ndarray
    camera_id: int
    name: str


class CameraInfo(NamedTuple):          # dataset_readers.py:26-38 without the image / depth paths
    uid: int
    R: np.ndarray                      # camera-to-world rotation (the transpose of the w2c block: "glm" convention)
    T: np.ndarray                      # world-to-camera translation
    FovY: float
    FovX: float
    image_name: str
    width: int
    height: int
    is_test: bool


def fov2focal(fov, pixels):
    return pixels / (2 * math.tan(fov / 2))


def focal2fov(focal, pixels):
    return 2 * math.atan(pixels / (2 * focal))


def qvec2rotmat(q):
    w, x, y, z = (float(v) for v in q)
    return np.array([[1 - 2 * y * y - 2 * z * z, 2 * x * y - 2 * w * z, 2 * z * x + 2 * w * y],
                     [2 * x * y + 2 * w * z, 1 - 2 * x * x - 2 * z * z, 2 * y * z - 2 * w * x],
                     [2 * z * x - 2 * w * y, 2 * y * z + 2 * w * x, 1 - 2 * x * x - 2 * y * y]])


# ---- COLMAP sparse model -------------------------------------------------------------------------------
def read_intrinsics_binary(path):
    buf = open(path, "rb").read()
    (n,), off, out = struct.unpack_from("<Q", buf, 0), 8, {}
    for _ in range(n):
        cid, model_id, w, h = struct.unpack_from("<iiQQ", buf, off)
        off += 24
        if model_id not in COLMAP_MODELS:
            raise ValueError("%s: unknown COLMAP camera model id %d" % (path, model_id))
        name, npar = COLMAP_MODELS[model_id]
        out[cid] = Intrinsics(cid, name, w, h, np.array(struct.unpack_from("<%dd" % npar, buf, off)))
        off += 8 * npar
    if len(out) != n:
        raise ValueError("%s: duplicate camera ids" % path)
    return out


def read_extrinsics_binary(path):
    buf = open(path, "rb").read()
    (n,), off, out = struct.unpack_from("<Q", buf, 0), 8, {}
    for _ in range(n):
        rec = struct.unpack_from("<i7di", buf, off)
        off += 64
        end = buf.index(b"\x00", off)
        name = buf[off:end].decode("utf-8")
        (npts,) = struct.unpack_from("<Q", buf, end + 1)
        off = end + 9 + 24 * npts                          # (x, y, point3D_id) observations: not needed by the path
        if off > len(buf):
            raise ValueError("%s: truncated" % path)
        out[rec[0]] = Extrinsics(rec[0], np.array(rec[1:5]), np.array(rec[5:8]), rec[8], name)
    return out


def _data_lines(path):
    with open(path, "r") as f:
        for line in f:
            line = line.strip()
            if line and line[0] != "#":
                yield line
            elif not line:
                yield ""                                   # an image without observations has an empty second line


def read_intrinsics_text(path):
    out = {}
    for line in _data_lines(path):
        if not line:
            continue
        e = line.split()
        if e[1] != "PINHOLE":                              # colmap_loader.py:171: the text reader asserts PINHOLE
            raise AssertionError("While the loader support other types, the rest of the code assumes PINHOLE")
        out[int(e[0])] = Intrinsics(int(e[0]), e[1], int(e[2]), int(e[3]), np.array([float(v) for v in e[4:]]))
    return out


def read_extrinsics_text(path):
    out, it = {}, _data_lines(path)
    for line in it:
        if not line:
            continue
        e = line.split()
        iid = int(e[0])
        out[iid] = Extrinsics(iid, np.array([float(v) for v in e[1:5]]), np.array([float(v) for v in e[5:8]]), int(e[8]), e[9])
        next(it, None)                                     # the observations line
    return out


def read_colmap_model(path):
    """sparse/0 of a COLMAP scene: binary files first, text as the fallback (dataset_readers.py:146-155)."""
    d = os.path.join(path, "sparse/0")
    try:
        return read_extrinsics_binary(os.path.join(d, "images.bin")), read_intrinsics_binary(os.path.join(d, "cameras.bin"))
    except (OSError, ValueError, struct.error):
        return read_extrinsics_text(os.path.join(d, "images.txt")), read_intrinsics_text(os.path.join(d, "cameras.txt"))


def colmap_camera_infos(extrinsics, intrinsics, test_names=()):
    """readColmapCameras (dataset_readers.py:71-118), in the order of `extrinsics`."""
    test_names, infos = set(test_names), []
    for ex in extrinsics.values():
        intr = intrinsics[ex.camera_id]
        if intr.model == "SIMPLE_PINHOLE":
            fx = fy = intr.params[0]
        elif intr.model == "PINHOLE":
            fx, fy = intr.params[0], intr.params[1]
        else:
            raise AssertionError("Colmap camera model not handled: only undistorted datasets (PINHOLE or SIMPLE_PINHOLE "
                                 "cameras) supported!")
        infos.append(CameraInfo(intr.id, qvec2rotmat(ex.qvec).T, np.array(ex.tvec), focal2fov(fy, intr.height),
                                focal2fov(fx, intr.width), ex.name, intr.width, intr.height, ex.name in test_names))
    return infos


def read_colmap_scene(path, eval=False, llffhold=8, train_test_exp=False):
    """readColmapSceneInfo's camera part (dataset_readers.py:145-207): (train, test, nerf_normalization)."""
    extr, intr = read_colmap_model(path)
    test_names = []
    if eval:
        if "360" in path:
            llffhold = 8
        if llffhold:
            test_names = [n for i, n in enumerate(sorted(e.name for e in extr.values())) if i % llffhold == 0]
        else:
            with open(os.path.join(path, "sparse/0", "test.txt")) as f:
                test_names = [line.strip() for line in f]
    infos = sorted(colmap_camera_infos(extr, intr, test_names), key=lambda c: c.image_name)
    train = [c for c in infos if train_test_exp or not c.is_test]
    return train, [c for c in infos if c.is_test], nerfpp_norm(train)


# ---- Blender / NeRF-synthetic ---------------------------------------------------------------------------
def png_size(path):
    """(width, height) from the IHDR chunk."""
    with open(path, "rb") as f:
        head = f.read(24)
    if len(head) < 24 or head[:8] != b"\x89PNG\r\n\x1a\n" or head[12:16] != b"IHDR":
        raise ValueError("%s: not a PNG" % path)
    return struct.unpack(">II", head[16:24])


def read_blender_transforms(path, transformsfile, is_test=False, extension=".png", size=None):
    """readCamerasFromTransforms (dataset_readers.py:228-271).  size=(W, H) overrides reading the PNG headers."""
    with open(os.path.join(path, transformsfile)) as f:
        contents = json.load(f)
    fovx, infos = contents["camera_angle_x"], []
    for idx, frame in enumerate(contents["frames"]):
        cam_name = os.path.join(path, frame["file_path"] + extension)
        c2w = np.array(frame["transform_matrix"], dtype=np.float64)
        c2w[:3, 1:3] *= -1                                   # Blender (y up, z back) -> COLMAP (y down, z forward)
        w2c = np.linalg.inv(c2w)
        W, H = size if size is not None else png_size(cam_name)
        infos.append(CameraInfo(idx, w2c[:3, :3].T.copy(), w2c[:3, 3].copy(), focal2fov(fov2focal(fovx, W), H), fovx,
                                os.path.splitext(os.path.basename(cam_name))[0], W, H, is_test))
    return infos


def read_blender_scene(path, eval=False, extension=".png", size=None):
    """readNerfSyntheticInfo's camera part (dataset_readers.py:273-285)."""
    train = read_blender_transforms(path, "transforms_train.json", False, extension, size)
    test = read_blender_transforms(path, "transforms_test.json", True, extension, size)
    if not eval:
        train, test = train + test, []
    return train, test, nerfpp_norm(train)


# ---- camera tensors -------------------------------------------------------------------------------------
def world_to_view(R, T, translate=(0.0, 0.0, 0.0), scale=1.0):
    """getWorld2View2 (graphics_utils.py:38-49)."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = np.asarray(R).T
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    c2w = np.linalg.inv(Rt)
    c2w[:3, 3] = (c2w[:3, 3] + np.asarray(translate)) * scale
    return np.float32(np.linalg.inv(c2w))


def nerfpp_norm(infos):
    """getNerfppNorm (dataset_readers.py:48-69): {"translate", "radius"}; radius = cameras_extent = spatial_lr_scale."""
    centers = np.hstack([np.linalg.inv(world_to_view(c.R, c.T))[:3, 3:4] for c in infos])
    center = np.mean(centers, axis=1, keepdims=True)
    diagonal = np.max(np.linalg.norm(centers - center, axis=0, keepdims=True))
    return {"translate": -center.flatten(), "radius": diagonal * 1.1}


def to_camera(info, trans=(0.0, 0.0, 0.0), scale=1.0, resolution_scale=1.0):
    """CameraInfo -> the Camera the rasterizer settings are built from (scene/cameras.py:80-89).  resolution_scale
    divides the image size as loadCam does (utils/camera_utils.py: round(orig / scale))."""
    W, H = round(info.width / resolution_scale), round(info.height / resolution_scale)
    return Camera(world_to_view(info.R, info.T, np.asarray(trans), scale), W, H, info.FovX, info.FovY)


# ---- COLMAP points3D and the initial point cloud (scene/colmap_loader.py:83-154, dataset_readers.py:120-143) ----
def read_points3D_binary(path):
    """-> (xyz [N,3] f64, rgb [N,3] f64, error [N,1] f64); the 2D tracks are skipped by their byte count."""
    buf = open(path, "rb").read()
    (n,), off = struct.unpack_from("<Q", buf, 0), 8
    xyz, rgb, err = np.empty((n, 3)), np.empty((n, 3)), np.empty((n, 1))
    for k in range(n):
        rec = struct.unpack_from("<QdddBBBdQ", buf, off)
        xyz[k], rgb[k], err[k] = rec[1:4], rec[4:7], rec[7]
        off += 51 + 8 * rec[8]
    if off > len(buf):
        raise ValueError("%s: truncated" % path)
    return xyz, rgb, err


def read_points3D_text(path):
    rows = [line.split() for line in _data_lines(path) if line]
    xyz = np.array([[float(v) for v in e[1:4]] for e in rows]).reshape(-1, 3)
    rgb = np.array([[int(v) for v in e[4:7]] for e in rows], dtype=np.float64).reshape(-1, 3)
    err = np.array([[float(e[7])] for e in rows]).reshape(-1, 1)
    return xyz, rgb, err


_PCD_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("nx", "<f4"), ("ny", "<f4"), ("nz", "<f4"),
                       ("red", "u1"), ("green", "u1"), ("blue", "u1")])


def store_ply(path, xyz, rgb):
    """storePly (dataset_readers.py:128-143): points3D.ply with float xyz + zero normals + uchar colours."""
    el = np.zeros(len(xyz), _PCD_DTYPE)
    for k, name in enumerate("xyz"):
        el[name] = np.asarray(xyz)[:, k]
    for k, name in enumerate(("red", "green", "blue")):
        el[name] = np.asarray(rgb)[:, k]
    head = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % len(el)
    head += "".join("property %s %s\n" % ("float" if _PCD_DTYPE[n].kind == "f" else "uchar", n) for n in _PCD_DTYPE.names)
    with open(path, "wb") as f:
        f.write((head + "end_header\n").encode("ascii") + el.tobytes())


def fetch_ply(path):
    """fetchPly (dataset_readers.py:120-126): (points [N,3] f32, colors [N,3] in 0..1, normals [N,3] f32).  Properties
    are looked up by name; float / uchar / double / int properties of any order are accepted."""
    buf = open(path, "rb").read()
    end = buf.index(b"end_header\n") + len(b"end_header\n")
    lines = buf[:end].decode("ascii").split("\n")
    if lines[0].strip() != "ply" or not any(l.startswith("format binary_little_endian") for l in lines):
        raise ValueError("%s: not a binary little-endian PLY" % path)
    np_of = dict(float="<f4", float32="<f4", double="<f8", float64="<f8", uchar="u1", uint8="u1", char="i1", int8="i1",
                 short="<i2", int16="<i2", ushort="<u2", uint16="<u2", int="<i4", int32="<i4", uint="<u4", uint32="<u4")
    n, fields, in_vertex = 0, [], False
    for l in lines:
        t = l.split()
        if t[:1] == ["element"]:
            in_vertex = t[1] == "vertex"
            n = int(t[2]) if in_vertex else n
        elif t[:1] == ["property"] and in_vertex:
            if t[1] == "list":
                raise ValueError("%s: list property inside the vertex element" % path)
            fields.append((t[2], np_of[t[1]]))
    v = np.frombuffer(buf, np.dtype(fields), count=n, offset=end)
    cols = lambda names: np.vstack([v[k] for k in names]).T
    return cols("xyz"), cols(("red", "green", "blue")) / 255.0, cols(("nx", "ny", "nz"))
