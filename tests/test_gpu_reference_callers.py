"""Acceptance run of the reference's OWN callers, unchanged, on this package (north_star: "gaussian_renderer/__init__.py
and train.py use it unchanged").

The caller files are imported / executed AS THEY ARE: from /root/reference when it exists (build container), else from
oracle/_ref/callers/ where oracle/build_ref.py staged them byte for byte (git-ignored; travels with the gpurun snapshot).
Only the two absent third-party imports are stubbed (tests/stubs: `plyfile`, `simple_knn._C`).

  (i)  gaussian_renderer.render() (gaussian_renderer/__init__.py:18-128) on a reference GaussianModel, every branch
       (separate_sh on/off, use_trained_exp on/off, scaling_modifier, the python SH / covariance paths), compared
       with the CPU oracle on the very tensors the model hands to the rasterizer;
  (ii) `python train.py ...` (the stock script: Scene loading, render, loss with fused_ssim, backward, densification
       statistics, densify_and_prune, opacity reset, Adam, evaluation, save_ply) for a few iterations on a tiny
       COLMAP-format dataset written by the test.
"""
import json
import math
import os
import subprocess
import sys
import types

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gaussian-splatting-npu_b200")
STUBS = os.path.join(ROOT, "tests", "stubs")


def _callers_dir():
    from oracle import build_ref
    d = build_ref.callers_dir()
    if d is None:
        pytest.skip("reference callers not staged (oracle/build_ref.py stage_callers() needs /root/reference at build time)")
    return d


@pytest.fixture(scope="module")
def ref_modules(cuda):
    """Imports the reference's gaussian_renderer / scene packages (unmodified) with this repo's rasterizer on the path."""
    d = _callers_dir()
    for p in (STUBS, d):
        if p not in sys.path:
            sys.path.insert(0, p)
    import diff_gaussian_rasterization as dgr
    assert os.path.realpath(dgr.__file__).startswith(os.path.realpath(PKG)), "the rasterizer under test must be this repo's"
    import gaussian_renderer
    from gaussian_renderer import network_gui     # a submodule: train.py imports it the same way
    from scene.cameras import MiniCam
    from scene.gaussian_model import GaussianModel
    assert os.path.realpath(gaussian_renderer.__file__).startswith(os.path.realpath(d))
    return types.SimpleNamespace(render=gaussian_renderer.render, MiniCam=MiniCam, GaussianModel=GaussianModel,
                                 network_gui=network_gui)


def _model_from_scene(ref, scene):
    """A reference GaussianModel holding the synthetic scene in its raw (pre-activation) parameterisation."""
    import torch
    from torch import nn
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    g = ref.GaussianModel(3)
    g.active_sh_degree = 3
    g._xyz = nn.Parameter(t(scene["means3D"]))
    g._features_dc = nn.Parameter(t(scene["shs"][:, :1]))
    g._features_rest = nn.Parameter(t(scene["shs"][:, 1:]))
    g._scaling = nn.Parameter(torch.log(t(scene["scales"])))
    g._rotation = nn.Parameter(t(scene["rotations"]))
    op = t(scene["opacities"]).clamp(1e-4, 1 - 1e-4)
    g._opacity = nn.Parameter(torch.log(op / (1 - op)))
    g.exposure_mapping = {"view0": 0}
    g.pretrained_exposures = None
    e = torch.eye(3, 4, device="cuda")[None].repeat(1, 1, 1)
    e[0] += torch.tensor([[0.05, -0.02, 0.01, 0.02], [0.0, -0.04, 0.03, -0.01], [0.02, 0.01, 0.06, 0.0]], device="cuda")
    g._exposure = nn.Parameter(e)
    return g


def _minicam(ref, cam):
    import torch
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    mc = ref.MiniCam(cam.image_width, cam.image_height, cam.FoVy, cam.FoVx, 0.01, 100.0, t(cam.world_view_transform),
                     t(cam.full_proj_transform))
    mc.image_name = "view0"
    return mc


RENDER_MODES = [
    dict(id="plain"), dict(id="separate_sh", separate_sh=True), dict(id="trained_exp", use_trained_exp=True),
    dict(id="separate_sh+exp+modifier", separate_sh=True, use_trained_exp=True, scaling_modifier=0.7),
    dict(id="python_sh_and_cov", convert_SHs_python=True, compute_cov3D_python=True, scaling_modifier=1.3),
    dict(id="antialiasing", antialiasing=True),
]


@pytest.mark.parametrize("mode", RENDER_MODES, ids=[m["id"] for m in RENDER_MODES])
def test_reference_render_unchanged(ref_modules, mode):
    import torch
    ref = ref_modules
    scene, cam, opts = cases.get_case("mid640x360")
    scene = {k: (v[:60_000] if isinstance(v, np.ndarray) else v) for k, v in scene.items()}
    pc = _model_from_scene(ref, scene)
    mc = _minicam(ref, cam)
    pipe = types.SimpleNamespace(debug=False, antialiasing=bool(mode.get("antialiasing", False)),
                                 convert_SHs_python=bool(mode.get("convert_SHs_python", False)),
                                 compute_cov3D_python=bool(mode.get("compute_cov3D_python", False)))
    bg = torch.tensor([0.2, 0.5, 0.3], device="cuda")
    smod = float(mode.get("scaling_modifier", 1.0))
    pkg = ref.render(mc, pc, pipe, bg, scaling_modifier=smod, separate_sh=bool(mode.get("separate_sh", False)),
                     use_trained_exp=bool(mode.get("use_trained_exp", False)))
    H, W = cam.image_height, cam.image_width
    dcol, ddep = cases.grads_for(cam)
    loss = (pkg["render"] * torch.from_numpy(dcol).cuda()).sum() + (pkg["depth"] * torch.from_numpy(ddep).cuda()).sum()
    loss.backward()
    torch.cuda.synchronize()

    # the oracle on exactly the tensors the reference model handed to the rasterizer
    n = lambda x: x.detach().cpu().numpy()
    sc = dict(means3D=n(pc.get_xyz), opacities=n(pc.get_opacity), sh_degree=3)
    o = dict(bg=tuple(n(bg).tolist()), antialiasing=pipe.antialiasing, mode="sh", scale_modifier=smod)
    if pipe.convert_SHs_python:
        from utils.sh_utils import eval_sh
        with torch.no_grad():
            shs_view = pc.get_features.transpose(1, 2).view(-1, 3, 16)
            d = pc.get_xyz - mc.camera_center.repeat(pc.get_features.shape[0], 1)
            sc["colors_precomp"] = n(torch.clamp_min(eval_sh(3, shs_view, d / d.norm(dim=1, keepdim=True)) + 0.5, 0.0))
        o["mode"] = "colors"
    else:
        sc["shs"] = n(pc.get_features)
    kw_scene = dict(sc)
    if pipe.compute_cov3D_python:
        kw = cases.inputs_of(dict(kw_scene, scales=None, rotations=None), cam, dict(o, mode=o["mode"]))
        kw.pop("scales"), kw.pop("rotations"), kw.pop("scale_modifier", None)
        kw["cov3D_precomp"] = n(pc.get_covariance(smod))
    else:
        kw_scene["scales"], kw_scene["rotations"] = n(pc.get_scaling), n(pc.get_rotation)
        kw = cases.inputs_of(kw_scene, cam, o)
    from oracle import gs_oracle as go
    S = go.Scene(W=W, H=H, **kw)
    f = go.forward(S)
    raw = f["out_color"].astype(np.float64)
    if mode.get("use_trained_exp"):
        E = n(pc._exposure[0]).astype(np.float64)
        raw = np.einsum("hwk,kc->chw", raw.transpose(1, 2, 0), E[:3, :3]) + E[:3, 3, None, None]
    want = np.clip(raw, 0.0, 1.0)
    ok = ~f["borderline"].astype(bool)
    assert np.array_equal(n(pkg["radii"]), f["radii"])
    assert np.array_equal(n(pkg["visibility_filter"]).ravel(), np.nonzero(f["radii"] > 0)[0])
    assert np.abs(n(pkg["render"]).astype(np.float64) - want)[:, ok].max() <= 2e-5
    assert np.abs(n(pkg["depth"])[0] - f["invdepth"][0])[ok].max() <= 1e-5
    # screen-space gradient (what add_densification_stats reads): oracle backward with the clamp / exposure chain applied
    passes = ((raw >= 0.0) & (raw <= 1.0)).astype(np.float64)
    up = dcol.astype(np.float64) * passes
    if mode.get("use_trained_exp"):
        up = np.einsum("chw,kc->khw", up, E[:3, :3])
    g = go.backward(S, f, up.astype(np.float32), ddep)
    got = n(pkg["viewspace_points"].grad)
    scale = np.abs(g["dL_dmean2D"]).max()
    assert np.abs(got[:, :2] - g["dL_dmean2D"][:, :2]).max() / scale <= 2e-4
    assert pc._xyz.grad is not None and torch.isfinite(pc._xyz.grad).all() and float(pc._xyz.grad.abs().max()) > 0
    if not pipe.compute_cov3D_python:
        assert float(pc._scaling.grad.abs().max()) > 0 and float(pc._rotation.grad.abs().max()) > 0
    if mode.get("use_trained_exp"):
        assert float(pc._exposure.grad.abs().max()) > 0


def _rotmat2qvec(R):
    """COLMAP's quaternion (w, x, y, z) of a rotation matrix (the inverse of colmap_loader.qvec2rotmat)."""
    K = np.array([[R[0, 0] - R[1, 1] - R[2, 2], 0, 0, 0],
                  [R[1, 0] + R[0, 1], R[1, 1] - R[0, 0] - R[2, 2], 0, 0],
                  [R[2, 0] + R[0, 2], R[2, 1] + R[1, 2], R[2, 2] - R[0, 0] - R[1, 1], 0],
                  [R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1], R[0, 0] + R[1, 1] + R[2, 2]]]) / 3.0
    vals, vecs = np.linalg.eigh(K)
    q = vecs[[3, 0, 1, 2], np.argmax(vals)]
    return q if q[0] >= 0 else -q


def _write_colmap_dataset(root, n_images=10, size=64, n_points=20_000):
    """A tiny COLMAP-format dataset in the TEXT flavour the reference falls back to (dataset_readers.py:150-154,
    :212-215): sparse/0/{cameras,images,points3D}.txt + images/*.png, coloured blobs seen from an orbit.
    (The reference's Blender reader cannot be used in this image: dataset_readers.py:259 hands PIL an int8 array,
    which Pillow 12 rejects -- independent of the rasterizer.)"""
    from PIL import Image
    from gsr_b200 import synth
    rng = np.random.default_rng(7)
    os.makedirs(os.path.join(root, "sparse", "0"), exist_ok=True)
    os.makedirs(os.path.join(root, "images"), exist_ok=True)
    focal = size / (2 * math.tan(0.35))
    with open(os.path.join(root, "sparse", "0", "cameras.txt"), "w") as f:
        f.write("# Camera list\n1 PINHOLE %d %d %.6f %.6f %.1f %.1f\n" % (size, size, focal, focal, size / 2, size / 2))
    with open(os.path.join(root, "sparse", "0", "images.txt"), "w") as f:
        f.write("# Image list\n")
        for k in range(n_images):
            ang = 2 * math.pi * k / n_images
            w2c = synth.look_at_w2c([4.0 * math.cos(ang), 4.0 * math.sin(ang), 1.0], [0, 0, 0]).astype(np.float64)
            q, t = _rotmat2qvec(w2c[:3, :3]), w2c[:3, 3]
            f.write("%d %.9f %.9f %.9f %.9f %.9f %.9f %.9f 1 img_%02d.png\n\n" % (k + 1, q[0], q[1], q[2], q[3], t[0], t[1], t[2], k))
            yy, xx = np.mgrid[0:size, 0:size] / size
            img = np.zeros((size, size, 3), np.float64)
            for _ in range(4):
                cx, cy, r = rng.random(), rng.random(), 0.1 + 0.2 * rng.random()
                img += np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * r * r))[..., None] * rng.random(3)
            Image.fromarray((np.clip(img, 0, 1) * 255).astype(np.uint8)).save(os.path.join(root, "images", "img_%02d.png" % k))
    with open(os.path.join(root, "sparse", "0", "points3D.txt"), "w") as f:
        f.write("# 3D point list\n")
        xyz = rng.normal(0, 0.6, (n_points, 3))
        rgb = rng.integers(0, 256, (n_points, 3))
        for i in range(n_points):
            f.write("%d %.6f %.6f %.6f %d %d %d 0.5\n" % (i + 1, xyz[i, 0], xyz[i, 1], xyz[i, 2], rgb[i, 0], rgb[i, 1], rgb[i, 2]))


@pytest.mark.parametrize("sparse_adam", [False, True], ids=["torch_adam", "sparse_adam_separate_sh"])
def test_stock_train_py_runs_unchanged(cuda, tmp_path, sparse_adam):
    """`python train.py` of the reference, unmodified, for a few iterations: densification statistics + densify_and_prune
    + opacity reset + optimizer + evaluation + save_ply all run on this repo's rasterizer and fused_ssim.  With
    GSR_EXPORT_SPARSE_ADAM=1 the script takes its separate_sh / SparseGaussianAdam path (train.py:37-41,111,177-181)."""
    d = _callers_dir()
    data, out = str(tmp_path / "data"), str(tmp_path / "out")
    _write_colmap_dataset(data)
    iters = 12
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([PKG, STUBS, d] + ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
    env.pop("GSR_EXPORT_SPARSE_ADAM", None)
    cmd = [sys.executable, os.path.join(d, "train.py"), "-s", data, "-m", out, "--eval", "--iterations", str(iters),
           "--test_iterations", str(iters), "--save_iterations", str(iters), "--densify_from_iter", "3",
           "--densification_interval", "4", "--opacity_reset_interval", "9", "--disable_viewer"]
    if sparse_adam:
        env["GSR_EXPORT_SPARSE_ADAM"] = "1"
        cmd += ["--optimizer_type", "sparse_adam"]
    r = subprocess.run(cmd, env=env, cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    log = r.stdout[-4000:] + "\n" + r.stderr[-4000:]
    assert r.returncode == 0, log
    assert "Training complete." in r.stdout, log
    assert ("[ITER %d] Evaluating test" % iters) in r.stdout, log
    ply = os.path.join(out, "point_cloud", "iteration_%d" % iters, "point_cloud.ply")
    assert os.path.exists(ply), log
    # the file the reference wrote is read back by this repo's loader and renders
    import torch
    from gsr_b200 import ply as gply, capi
    m = gply.load_ply_for_render(ply, device="cuda")
    P = m["xyz"].shape[0]
    assert P >= 1000 and all(bool(torch.isfinite(v).all()) for v in m.values())
    cam = cases._lego_cam(2, 64, 64)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    st = capi.forward(means3D=m["xyz"], opacities=m["opacity"], shs=m["shs"], scales=m["scales"], rotations=m["rotations"],
                      viewmatrix=t(cam.world_view_transform), projmatrix=t(cam.full_proj_transform), cam_pos=t(cam.camera_center),
                      background=torch.zeros(3, device="cuda"), width=64, height=64, tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy)
    assert bool(torch.isfinite(st["out_color"]).all()) and st["num_rendered"] > 0
