"""Camera loaders (gsr_b200/cameras.py) against golden vectors produced by the reference's own Python code
(tests/golden/make_golden_cameras.py: colmap_loader.py, dataset_readers.py, graphics_utils.py, cameras.py:86-89).
Parsed values (R, T, FoV, sizes, names, split) must be IDENTICAL; matrices formed in fp32 within 1e-6."""
import os
import struct
import zlib

import numpy as np
import pytest

from gsr_b200 import cameras as cm

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "cameras_golden.npz"))


def check_infos(gold, prefix, infos):
    assert [c.image_name for c in infos] == list(gold[prefix + "_name"])
    assert np.array_equal(np.stack([c.R for c in infos]), gold[prefix + "_R"])
    assert np.array_equal(np.stack([c.T for c in infos]), gold[prefix + "_T"])
    assert np.array_equal(np.array([[c.FovX, c.FovY] for c in infos]), gold[prefix + "_fov"])
    assert np.array_equal(np.array([[c.width, c.height] for c in infos]), gold[prefix + "_size"])
    assert np.array_equal(np.array([c.uid for c in infos]), gold[prefix + "_uid"])
    assert np.array_equal(np.array([c.is_test for c in infos]), gold[prefix + "_is_test"])


def check_tensors(gold, prefix, cams):
    for k, c in enumerate(cams):
        assert np.abs(c.world_view_transform - gold[prefix + "_world_view"][k]).max() <= 1e-6
        assert np.abs(c.projection_matrix - gold[prefix + "_projection"][k]).max() <= 1e-6
        assert np.abs(c.full_proj_transform - gold[prefix + "_full_proj"][k]).max() <= 2e-6 * np.abs(gold[prefix + "_full_proj"][k]).max()
        assert np.abs(c.camera_center - gold[prefix + "_center"][k]).max() <= 2e-6 * max(1.0, np.abs(gold[prefix + "_center"][k]).max())


@pytest.mark.parametrize("sub", ["colmap_tiny", "colmap_tiny_txt"])
def test_colmap_model_vs_reference(gold, sub):
    train, test, norm = cm.read_colmap_scene(os.path.join(GOLD, sub), eval=True, llffhold=3)
    infos = sorted(train + test, key=lambda c: c.image_name)
    check_infos(gold, sub, infos)
    assert [c.image_name for c in test] == [n for i, n in enumerate(sorted(c.image_name for c in infos)) if i % 3 == 0]
    assert np.allclose(norm["translate"], gold[sub + "_translate"], rtol=0, atol=1e-12)
    assert abs(norm["radius"] - float(gold[sub + "_radius"])) <= 1e-12
    check_tensors(gold, sub, [cm.to_camera(c, norm["translate"], 0.5) for c in infos])


def test_binary_is_preferred_and_text_is_the_fallback(tmp_path):
    extr, intr = cm.read_colmap_model(os.path.join(GOLD, "colmap_tiny"))
    assert {i.model for i in intr.values()} == {"PINHOLE", "SIMPLE_PINHOLE"} and len(extr) == 7
    extr_t, intr_t = cm.read_colmap_model(os.path.join(GOLD, "colmap_tiny_txt"))
    assert len(extr_t) == 7 and all(i.model == "PINHOLE" for i in intr_t.values())
    with pytest.raises(OSError):
        cm.read_colmap_model(str(tmp_path))


def test_colmap_error_paths(tmp_path):
    d = tmp_path / "sparse" / "0"
    d.mkdir(parents=True)
    (d / "cameras.bin").write_bytes(struct.pack("<Q", 1) + struct.pack("<iiQQ", 1, 4, 64, 48) + struct.pack("<8d", *range(8)))
    (d / "images.bin").write_bytes(struct.pack("<Q", 1) + struct.pack("<i7di", 1, 1, 0, 0, 0, 0, 0, 0, 1) + b"a.png\x00" + struct.pack("<Q", 0))
    extr, intr = cm.read_colmap_model(str(tmp_path))
    with pytest.raises(AssertionError, match="only undistorted datasets"):       # OPENCV model: dataset_readers.py:96
        cm.colmap_camera_infos(extr, intr)
    (d / "cameras.txt").write_text("1 OPENCV 64 48 1 1 1 1 0 0 0 0\n")
    with pytest.raises(AssertionError, match="assumes PINHOLE"):                 # colmap_loader.py:171
        cm.read_intrinsics_text(str(d / "cameras.txt"))
    (d / "images.bin").write_bytes(struct.pack("<Q", 1) + struct.pack("<i7di", 1, 1, 0, 0, 0, 0, 0, 0, 1) + b"a.png\x00" + struct.pack("<Q", 3))
    with pytest.raises(ValueError, match="truncated"):
        cm.read_extrinsics_binary(str(d / "images.bin"))


def test_blender_transforms_vs_reference(gold):
    path = os.path.join(GOLD, "lego_transforms")
    infos = cm.read_blender_transforms(path, "transforms_train.json", size=(800, 800))
    assert len(infos) == 100
    check_infos(gold, "lego", infos)
    norm = cm.nerfpp_norm(infos)
    assert np.allclose(norm["translate"], gold["lego_translate"], rtol=0, atol=1e-12)
    assert abs(norm["radius"] - float(gold["lego_radius"])) <= 1e-12
    check_tensors(gold, "lego", [cm.to_camera(c) for c in infos[:10]])
    train, test, _ = cm.read_blender_scene(path, eval=False, size=(800, 800))
    assert len(train) == 108 and test == [] and train[100].is_test
    train, test, _ = cm.read_blender_scene(path, eval=True, size=(800, 800))
    assert len(train) == 100 and len(test) == 8


def test_image_size_from_png_header(tmp_path):
    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data))
    W, H = 37, 21
    raw = b"".join(b"\x00" + bytes(3 * W) for _ in range(H))
    png = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", W, H, 8, 2, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b"")
    (tmp_path / "train").mkdir()
    (tmp_path / "train" / "r_0.png").write_bytes(png)
    (tmp_path / "t.json").write_text('{"camera_angle_x": 0.7, "frames": [{"file_path": "./train/r_0", "transform_matrix": '
                                     '[[1,0,0,0],[0,1,0,0],[0,0,1,4],[0,0,0,1]]}]}')
    (info,) = cm.read_blender_transforms(str(tmp_path), "t.json")
    assert (info.width, info.height, info.image_name) == (W, H, "r_0")
    assert info.FovY == cm.focal2fov(cm.fov2focal(0.7, W), H)
    (tmp_path / "bad.png").write_bytes(b"not a png at all, really not")
    with pytest.raises(ValueError, match="not a PNG"):
        cm.png_size(str(tmp_path / "bad.png"))


@pytest.mark.parametrize("sub", ["colmap_tiny", "colmap_tiny_txt"])
def test_points3d_vs_reference(gold, sub):
    d = os.path.join(GOLD, sub, "sparse/0")
    xyz, rgb, err = cm.read_points3D_text(os.path.join(d, "points3D.txt")) if sub.endswith("txt") else \
        cm.read_points3D_binary(os.path.join(d, "points3D.bin"))
    assert np.array_equal(xyz, gold[sub + "_p3d_xyz"]) and np.array_equal(rgb, gold[sub + "_p3d_rgb"])
    assert np.array_equal(err, gold[sub + "_p3d_err"])


def test_point_cloud_ply_round_trip_and_foreign_layout(tmp_path):
    rng = np.random.default_rng(6)
    xyz, rgb = rng.standard_normal((11, 3)), rng.integers(0, 256, (11, 3))
    p = str(tmp_path / "points3D.ply")
    cm.store_ply(p, xyz, rgb)
    raw = open(p, "rb").read()
    head = raw[:raw.index(b"end_header\n")].decode()
    # the header storePly's structured dtype produces (dataset_readers.py:130-132): 6 floats then 3 uchars, 27 B/record
    assert head.split("\n")[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 11"]
    assert [l.split()[1:] for l in head.split("\n")[3:] if l] == [["float", n] for n in ("x", "y", "z", "nx", "ny", "nz")] + \
        [["uchar", n] for n in ("red", "green", "blue")]
    assert len(raw) == len(head) + len("end_header\n") + 11 * 27
    pts, col, nrm = cm.fetch_ply(p)
    assert np.array_equal(pts, xyz.astype(np.float32)) and np.array_equal(col, rgb / 255.0) and not nrm.any()
    # another tool's layout: colours first, doubles for the position, an extra property
    dt = np.dtype([("red", "u1"), ("green", "u1"), ("blue", "u1"), ("x", "<f8"), ("y", "<f8"), ("z", "<f8"),
                   ("nx", "<f4"), ("ny", "<f4"), ("nz", "<f4"), ("confidence", "<i4")])
    el = np.zeros(4, dt)
    el["x"], el["green"], el["nz"] = [1, 2, 3, 4], [10, 20, 30, 40], 1.0
    hdr = "ply\nformat binary_little_endian 1.0\nelement vertex 4\n" + "".join(
        "property %s %s\n" % ({"u1": "uchar", "<f8": "double", "<f4": "float", "<i4": "int"}[dt[n].str.replace("|", "")], n) for n in dt.names)
    q = str(tmp_path / "other.ply")
    open(q, "wb").write((hdr + "element face 0\nproperty list uchar int vertex_indices\nend_header\n").encode() + el.tobytes())
    pts, col, nrm = cm.fetch_ply(q)
    assert np.array_equal(pts[:, 0], [1, 2, 3, 4]) and np.allclose(col[:, 1], np.array([10, 20, 30, 40]) / 255.0) and (nrm[:, 2] == 1).all()
    open(q, "wb").write(b"ply\nformat ascii 1.0\nend_header\n")
    with pytest.raises(ValueError, match="binary little-endian"):
        cm.fetch_ply(q)
