"""GPU parity of the PLY transposition kernels (csrc/ply_io.cu) through the C ABI against the numpy oracle
(oracle/ply_oracle.py): bit-exact for the raw values (pure data movement), <= 2 ulp-level tolerance for the fused
activations; reference fixtures, ragged sizes, property orders other than save_ply's, unaligned records, and a
render of the reference's single-Gaussian fixture through the rasterizer."""
import os

import numpy as np
import pytest

from oracle import ply_oracle as po
from test_ply_cpu import GOLD, _raw

pytestmark = pytest.mark.gpu


def _rand_params(P, M, seed):
    rng = np.random.default_rng(seed)
    p = dict(xyz=rng.standard_normal((P, 3)), features_dc=rng.standard_normal((P, 1, 3)), features_rest=rng.standard_normal((P, M - 1, 3)),
             opacity=rng.standard_normal((P, 1)) * 3, scaling=rng.standard_normal((P, 3)) - 3, rotation=rng.standard_normal((P, 4)))
    return {k: np.ascontiguousarray(v, np.float32) for k, v in p.items()}


@pytest.mark.parametrize("name", ["single_gaussian.ply", "two_gaussians.ply"])
def test_load_reference_fixture(cuda, name):
    from gsr_b200 import ply
    got = ply.load_ply(os.path.join(GOLD, name), cuda)
    want = po.load_ply(_raw(name))
    for k in want:
        assert np.array_equal(got[k].cpu().numpy(), want[k]), k


@pytest.mark.parametrize("P,deg", [(1, 3), (127, 3), (128, 3), (1000, 3), (300, 1), (129, 0), (70001, 3)])
def test_save_then_load_round_trip_and_oracle(cuda, tmp_path, P, deg):
    import torch
    from gsr_b200 import ply
    M = (deg + 1) ** 2
    p = _rand_params(P, M, P + deg)
    t = {k: torch.from_numpy(v).to(cuda) for k, v in p.items()}
    path = str(tmp_path / "m.ply")
    ply.save_ply(path, t["xyz"], torch.cat([t["features_dc"], t["features_rest"]], dim=1), t["opacity"], t["scaling"], t["rotation"])
    with open(path, "rb") as f:
        assert f.read() == po.save_ply(p)                       # the file is byte-identical to the reference's save_ply
    got = ply.load_ply(path, cuda, max_sh_degree=deg)
    for k in p:
        assert np.array_equal(got[k].cpu().numpy(), p[k]), k
    act = ply.load_ply_for_render(path, cuda, sh_degree=deg)
    want = po.activate(p)
    assert np.array_equal(act["xyz"].cpu().numpy(), want["xyz"]) and np.array_equal(act["shs"].cpu().numpy(), want["shs"])
    for k in ("opacity", "scales", "rotations"):
        assert np.allclose(act[k].cpu().numpy(), want[k], rtol=2e-6, atol=1e-7), k


def test_foreign_property_order_unaligned_records_and_extra_types(cuda):
    """Files written by other tools: attributes in any order, extra uchar/double properties (odd record size ->
    byte-wise path), a lower SH degree than the output tensor (missing coefficients are zero)."""
    import torch
    from gsr_b200 import ply
    rng = np.random.default_rng(5)
    P = 333
    names = ["rot_2", "x", "opacity", "f_dc_1", "scale_0", "y", "rot_0", "f_dc_0", "z", "scale_2", "rot_3", "f_dc_2", "scale_1", "rot_1"] + \
            ["f_rest_%d" % i for i in range(9)]
    fields = [("red", "u1")] + [(n, "<f4") for n in names] + [("w", "<f8")]
    v = np.zeros(P, dtype=np.dtype(fields))
    for n in names:
        v[n] = rng.standard_normal(P).astype(np.float32)
    tmap = {"u1": "uchar", "<f4": "float", "<f8": "double"}
    head = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P + "".join("property %s %s\n" % (tmap[t], n) for n, t in fields) +
            "end_header\n").encode()
    raw = head + v.tobytes()
    lay = ply.parse_header(raw)
    assert lay.record_bytes == 1 + 4 * len(names) + 8 and lay.record_bytes % 4 != 0
    dev = torch.from_numpy(np.frombuffer(raw, np.uint8, offset=lay.header_bytes).copy()).to(cuda)
    got = ply.unpack(lay, dev, sh_degree=3, activate=False)
    want = po.load_ply(raw, max_sh_degree=1)
    assert np.array_equal(got["xyz"].cpu().numpy(), want["xyz"])
    assert np.array_equal(got["shs"][:, :1].cpu().numpy(), want["features_dc"])
    assert np.array_equal(got["shs"][:, 1:4].cpu().numpy(), want["features_rest"])
    assert float(got["shs"][:, 4:].abs().max()) == 0.0
    assert np.array_equal(got["rotations"].cpu().numpy(), want["rotation"]) and np.array_equal(got["scales"].cpu().numpy(), want["scaling"])


def test_fixture_renders_through_the_rasterizer(cuda):
    """The fork's single-Gaussian fixture, loaded on the GPU, rendered with the chair camera of tests/cases.py and
    compared with the oracle rasterizer fed by the oracle loader."""
    import torch
    import cases
    from gsr_b200 import capi, ply
    a = ply.load_ply_for_render(os.path.join(GOLD, "two_gaussians.ply"), cuda)
    w = po.activate(po.load_ply(_raw("two_gaussians.ply")))
    _, cam, opts = cases.get_case("two")
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(cuda)
    st = capi.forward(means3D=a["xyz"], opacities=a["opacity"], shs=a["shs"], scales=a["scales"], rotations=a["rotations"],
                      viewmatrix=t(cam.world_view_transform), projmatrix=t(cam.full_proj_transform), cam_pos=t(cam.camera_center),
                      background=t(np.array(opts["bg"], np.float32)), width=cam.image_width, height=cam.image_height,
                      tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy, sh_degree=3)
    scene = dict(means3D=w["xyz"], shs=w["shs"], opacities=w["opacity"], scales=w["scales"], rotations=w["rotations"], sh_degree=3)
    o = cases.run_oracle(scene, cam, dict(opts, mode="sh"), None)
    assert np.array_equal(st["radii"].cpu().numpy(), o["radii"])
    assert np.abs(st["out_color"].cpu().numpy() - o["out_color"])[:, ~o["borderline"]].max() <= 1e-5


def test_real_scene_flow_ply_plus_colmap_camera(cuda, tmp_path):
    """What a caller with a trained scene does: point_cloud.ply from disk (device unpack), cameras from the COLMAP
    sparse model (gsr_b200/cameras.py), one render per camera through the C ABI, checked against the oracle rasterizer
    fed by the oracle loader."""
    import torch
    import cases
    from gsr_b200 import capi, ply, synth, cameras as cm
    train, test, norm = cm.read_colmap_scene(os.path.join(GOLD, "colmap_tiny"), eval=True, llffhold=3)
    infos = (train + test)[:3]
    rng = np.random.default_rng(8)
    sc = synth.trained_like_scene(3000, 8, s0=0.05)
    path = str(tmp_path / "point_cloud.ply")
    for k, info in enumerate(infos):
        cam = cm.to_camera(info, resolution_scale=4.0)
        assert (cam.image_width, cam.image_height) == (round(info.width / 4), round(info.height / 4))
        # put the cloud 4 units in front of this camera: forward axis = third row of the w2c rotation
        w2c = cam.world_view_transform.T
        centre = cam.camera_center + 4.0 * w2c[2, :3]
        xyz = (centre + 0.6 * rng.standard_normal((3000, 3))).astype(np.float32)
        raw = {k2: torch.from_numpy(np.ascontiguousarray(v)) for k2, v in
               dict(xyz=xyz, shs=sc["shs"], opacity=np.log(sc["opacities"] / (1 - sc["opacities"])).reshape(-1, 1),
                    scaling=np.log(sc["scales"]), rotation=sc["rotations"]).items()}
        ply.save_ply(path, *[raw[n].to(cuda) for n in ("xyz", "shs", "opacity", "scaling", "rotation")])
        a = ply.load_ply_for_render(path, cuda)
        w = po.activate(po.load_ply(open(path, "rb").read()))
        t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(cuda)
        st = capi.forward(means3D=a["xyz"], opacities=a["opacity"], shs=a["shs"], scales=a["scales"], rotations=a["rotations"],
                          viewmatrix=t(cam.world_view_transform), projmatrix=t(cam.full_proj_transform), cam_pos=t(cam.camera_center),
                          background=t(np.zeros(3, np.float32)), width=cam.image_width, height=cam.image_height,
                          tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy, sh_degree=3)
        scene = dict(means3D=w["xyz"], shs=w["shs"], opacities=w["opacity"], scales=w["scales"], rotations=w["rotations"], sh_degree=3)
        o = cases.run_oracle(scene, cam, dict(bg=(0.0, 0.0, 0.0), antialiasing=False, mode="sh"), None)
        assert int((o["radii"] > 0).sum()) > 1000, "camera %d does not see the cloud" % k
        assert np.array_equal(st["radii"].cpu().numpy(), o["radii"])
        assert np.abs(st["out_color"].cpu().numpy() - o["out_color"])[:, ~o["borderline"]].max() <= 1e-5
