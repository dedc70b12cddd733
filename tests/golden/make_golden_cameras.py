"""Golden vectors of the camera loaders (gsr_b200/cameras.py), produced by the REFERENCE's own Python code.

Run in the build container (needs /root/reference):   python tests/golden/make_golden_cameras.py
Writes
  tests/golden/colmap_tiny/sparse/0/{cameras,images}.bin      a 7-image COLMAP model, PINHOLE + SIMPLE_PINHOLE
  tests/golden/colmap_tiny_txt/sparse/0/{cameras,images}.txt  the text form (PINHOLE only: the reference asserts it)
  tests/golden/lego_transforms/transforms_{train,test}.json   the camera part of the NeRF-synthetic lego set the
                                                              reference ships (file_path + transform_matrix)
  tests/golden/cameras_golden.npz                             what the reference returns for them
The reference modules are loaded by file path (scene/__init__.py pulls in the CUDA extensions); `plyfile` (absent
here) and scene.gaussian_model are stubbed: neither is touched by the camera functions.
"""
import importlib.util
import json
import os
import struct
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def reference_modules():
    for pkg in ("scene", "utils"):
        sys.modules[pkg] = types.ModuleType(pkg)
        sys.modules[pkg].__path__ = []
    cl = load("scene.colmap_loader", "scene/colmap_loader.py")
    gu = load("utils.graphics_utils", "utils/graphics_utils.py")
    load("utils.sh_utils", "utils/sh_utils.py")
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object
    sys.modules["plyfile"] = ply
    gm = types.ModuleType("scene.gaussian_model")
    gm.BasicPointCloud = gu.BasicPointCloud
    sys.modules["scene.gaussian_model"] = gm
    dr = load("scene.dataset_readers", "scene/dataset_readers.py")
    # dataset_readers.py:257 hands PIL an int8 array, which the PIL of this image rejects; the pixels are not part of
    # these vectors, so the call is given the same bytes as uint8 (only image.size is read afterwards)
    pil = dr.Image
    shim = types.SimpleNamespace(open=pil.open, fromarray=lambda a, mode=None: pil.fromarray(a.view(np.uint8), mode))
    dr.Image = shim
    return cl, gu, dr


def write_colmap(d, cams, images, text):
    os.makedirs(d, exist_ok=True)
    if text:
        with open(os.path.join(d, "cameras.txt"), "w") as f:
            f.write("# Camera list with one line of data per camera:\n#   CAMERA_ID, MODEL, WIDTH, HEIGHT, PARAMS[]\n")
            for cid, (model, w, h, params) in cams.items():
                f.write("%d %s %d %d %s\n" % (cid, model, w, h, " ".join(repr(float(p)) for p in params)))
        with open(os.path.join(d, "images.txt"), "w") as f:
            f.write("# Image list with two lines of data per image:\n\n")
            for iid, (q, t, cid, name, pts) in images.items():
                f.write("%d %s %s %d %s\n" % (iid, " ".join(repr(float(v)) for v in q), " ".join(repr(float(v)) for v in t), cid, name))
                f.write(" ".join("%r %r %d" % (float(x), float(y), int(p)) for x, y, p in pts) + "\n")
        return
    ids = {"SIMPLE_PINHOLE": 0, "PINHOLE": 1}
    with open(os.path.join(d, "cameras.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(cams)))
        for cid, (model, w, h, params) in cams.items():
            f.write(struct.pack("<iiQQ", cid, ids[model], w, h) + struct.pack("<%dd" % len(params), *params))
    with open(os.path.join(d, "images.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(images)))
        for iid, (q, t, cid, name, pts) in images.items():
            f.write(struct.pack("<i7di", iid, *q, *t, cid) + name.encode() + b"\x00" + struct.pack("<Q", len(pts)))
            for x, y, p in pts:
                f.write(struct.pack("<ddq", x, y, p))


def write_points3d(d, pts, text):
    if text:
        with open(os.path.join(d, "points3D.txt"), "w") as f:
            f.write("# 3D point list with one line of data per point:\n#   POINT3D_ID, X, Y, Z, R, G, B, ERROR, TRACK[] as (IMAGE_ID, POINT2D_IDX)\n")
            for pid, (xyz, rgb, err, track) in pts.items():
                f.write("%d %s %d %d %d %r %s\n" % (pid, " ".join(repr(float(v)) for v in xyz), *rgb, float(err),
                                                    " ".join("%d %d" % t for t in track)))
        return
    with open(os.path.join(d, "points3D.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(pts)))
        for pid, (xyz, rgb, err, track) in pts.items():
            f.write(struct.pack("<QdddBBBdQ", pid, *xyz, *rgb, err, len(track)))
            for t in track:
                f.write(struct.pack("<ii", *t))


def tiny_points(rng):
    pts = {}
    for pid in (3, 1, 20, 7, 8):
        track = [(int(rng.integers(1, 13)), int(rng.integers(0, 5))) for _ in range(int(rng.integers(0, 4)))]
        pts[pid] = (rng.standard_normal(3) * 2, tuple(int(v) for v in rng.integers(0, 256, 3)), float(rng.random()), track)
    return pts


def tiny_model(rng, models):
    cams = {}
    for k, model in enumerate(models):
        w, h = 640 + 32 * k, 480 - 16 * k
        cams[k + 1] = (model, w, h, [500.25 + 7 * k, w / 2, h / 2] if model == "SIMPLE_PINHOLE" else [510.5 + k, 498.75 - k, w / 2, h / 2])
    images = {}
    for i, iid in enumerate([5, 2, 9, 1, 12, 7, 3]):                     # file order != name order != id order
        q = rng.standard_normal(4)
        q /= np.linalg.norm(q)
        npts = [0, 3, 1, 5, 0, 2, 4][i]
        pts = [(rng.uniform(0, 600), rng.uniform(0, 400), int(rng.integers(-1, 1000))) for _ in range(npts)]
        images[iid] = (q, rng.standard_normal(3) * 3, 1 + i % len(models), "frame_%02d.JPG" % ((i * 5) % 7), pts)
    return cams, images


def camera_tensors(gu, torch, info, trans, scale):
    """scene/cameras.py:86-89 on the CPU (the Camera class itself calls .cuda())."""
    wv = torch.tensor(gu.getWorld2View2(info.R, info.T, trans, scale)).transpose(0, 1)
    pr = gu.getProjectionMatrix(znear=0.01, zfar=100.0, fovX=info.FovX, fovY=info.FovY).transpose(0, 1)
    full = (wv.unsqueeze(0).bmm(pr.unsqueeze(0))).squeeze(0)
    return wv.numpy(), pr.numpy(), full.numpy(), wv.inverse()[3, :3].numpy()


def pack_infos(prefix, infos, out):
    out[prefix + "_R"] = np.stack([c.R for c in infos])
    out[prefix + "_T"] = np.stack([c.T for c in infos])
    out[prefix + "_fov"] = np.array([[c.FovX, c.FovY] for c in infos])
    out[prefix + "_size"] = np.array([[c.width, c.height] for c in infos])
    out[prefix + "_uid"] = np.array([c.uid for c in infos])
    out[prefix + "_name"] = np.array([c.image_name for c in infos])
    out[prefix + "_is_test"] = np.array([c.is_test for c in infos])


if __name__ == "__main__":
    import torch
    cl, gu, dr = reference_modules()
    rng = np.random.default_rng(20)
    out = {}
    for sub, models, text in (("colmap_tiny", ["PINHOLE", "SIMPLE_PINHOLE"], False), ("colmap_tiny_txt", ["PINHOLE"], True)):
        cams, images = tiny_model(rng, models)
        d = os.path.join(HERE, sub, "sparse/0")
        write_colmap(d, cams, images, text)
        write_points3d(d, tiny_points(np.random.default_rng(77)), text)
        pr = cl.read_points3D_text(os.path.join(d, "points3D.txt")) if text else cl.read_points3D_binary(os.path.join(d, "points3D.bin"))
        out[sub + "_p3d_xyz"], out[sub + "_p3d_rgb"], out[sub + "_p3d_err"] = pr
        if text:
            extr, intr = cl.read_extrinsics_text(os.path.join(d, "images.txt")), cl.read_intrinsics_text(os.path.join(d, "cameras.txt"))
        else:
            extr, intr = cl.read_extrinsics_binary(os.path.join(d, "images.bin")), cl.read_intrinsics_binary(os.path.join(d, "cameras.bin"))
        names = sorted(e.name for e in extr.values())
        test_names = [n for i, n in enumerate(names) if i % 3 == 0]      # readColmapSceneInfo with eval, llffhold 3
        infos = sorted(dr.readColmapCameras(extr, intr, None, "images", "", test_names), key=lambda c: c.image_name)
        pack_infos(sub, infos, out)
        norm = dr.getNerfppNorm([c for c in infos if not c.is_test])
        out[sub + "_translate"], out[sub + "_radius"] = norm["translate"], np.float64(norm["radius"])
        tens = [camera_tensors(gu, torch, c, norm["translate"], 0.5) for c in infos]
        for k, nm in enumerate(("world_view", "projection", "full_proj", "center")):
            out["%s_%s" % (sub, nm)] = np.stack([t[k] for t in tens])
    # Blender: the lego set shipped with the reference (the PNGs are read there; only the cameras are kept here)
    lego = os.path.join(REF, "nerf_synthetic/lego")
    os.makedirs(os.path.join(HERE, "lego_transforms"), exist_ok=True)
    for split in ("train", "test"):
        with open(os.path.join(lego, "transforms_%s.json" % split)) as f:
            t = json.load(f)
        keep = t["frames"] if split == "train" else t["frames"][:8]
        slim = {"camera_angle_x": t["camera_angle_x"],
                "frames": [{"file_path": fr["file_path"], "transform_matrix": fr["transform_matrix"]} for fr in keep]}
        with open(os.path.join(HERE, "lego_transforms", "transforms_%s.json" % split), "w") as f:
            json.dump(slim, f)
    print()
    infos = dr.readCamerasFromTransforms(lego, "transforms_train.json", "", False, False)
    pack_infos("lego", infos, out)
    norm = dr.getNerfppNorm(infos)
    out["lego_translate"], out["lego_radius"] = norm["translate"], np.float64(norm["radius"])
    tens = [camera_tensors(gu, torch, c, np.array([0.0, 0.0, 0.0]), 1.0) for c in infos[:10]]
    for k, nm in enumerate(("world_view", "projection", "full_proj", "center")):
        out["lego_%s" % nm] = np.stack([t[k] for t in tens])
    np.savez_compressed(os.path.join(HERE, "cameras_golden.npz"), **out)
    print("wrote cameras_golden.npz:", {k: v.shape for k, v in out.items() if k.endswith("_R")})
