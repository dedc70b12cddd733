"""Seeded parity cases shared by the CPU tests, the GPU tests and tests/golden/make_golden.py.

A case is (scene dict of numpy arrays, synth.Camera, options).  Sizes are chosen so that the CPU
oracle finishes in seconds.  `run_oracle`, `run_cuda` and `run_ref` evaluate the same case through
the three implementations and return dicts of numpy arrays with identical keys.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gaussian-splatting-npu_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

from gsr_b200 import synth  # noqa: E402


def _lego_cam(k, W=800, H=800):
    with open(os.path.join(ROOT, "tests", "golden", "lego_cams.json")) as f:
        t = json.load(f)
    return synth.blender_camera(t["transform_matrix"][k % len(t["transform_matrix"])], W, H, t["camera_angle_x"])


def _one_gaussian():
    # a single isotropic Gaussian in front of the chair camera (values of GS-IRON/base/gaussian.py's
    # single_gaussian.ply: scale exp(-3), opacity sigmoid(2), DC colour only)
    sc = dict(means3D=np.array([[0.0, 0.0, 0.2]], np.float32),
              scales=np.full((1, 3), np.exp(-3.0), np.float32),
              rotations=np.array([[1.0, 0, 0, 0]], np.float32),
              opacities=np.array([[1 / (1 + np.exp(-2.0))]], np.float32),
              shs=np.zeros((1, 16, 3), np.float32), sh_degree=3)
    sc["shs"][0, 0] = [1.0, 0.2, -0.5]
    return sc


def _two_gaussians():
    sc = dict(means3D=np.array([[0.0, 0.0, 0.2], [0.05, 0.02, 0.5]], np.float32),
              scales=np.array([[0.05, 0.02, 0.03], [0.08, 0.08, 0.01]], np.float32),
              rotations=np.array([[1.0, 0, 0, 0], [0.8, 0.1, -0.5, 0.3]], np.float32),
              opacities=np.array([[0.9], [0.6]], np.float32),
              shs=np.random.default_rng(11).normal(0, 0.5, (2, 16, 3)).astype(np.float32), sh_degree=3)
    return sc


def get_case(name):
    """-> (scene, camera, opts) ; opts: bg, antialiasing, mode ('sh'|'colors'|'cov3d'), sh_degree."""
    o = dict(bg=(0.0, 0.0, 0.0), antialiasing=False, mode="sh")
    if name == "one":
        return _one_gaussian(), synth.chair_camera(128, 128), o
    if name == "two":
        return _two_gaussians(), synth.chair_camera(96, 80), dict(o, bg=(1.0, 1.0, 1.0))
    if name == "tiny64":
        return synth.trained_like_scene(1000, 64, s0=0.08), synth.orbit_cameras(3, 64, 64, 64)[1], o
    if name == "odd167x101":
        return synth.trained_like_scene(50_000, 167, s0=0.03), synth.orbit_cameras(3, 167, 101, 167)[2], dict(o, bg=(0.3, 0.6, 0.9))
    if name == "aa160x96":
        return synth.trained_like_scene(30_000, 160, s0=0.02), synth.orbit_cameras(3, 160, 96, 160)[0], dict(o, antialiasing=True, bg=(1.0, 1.0, 1.0))
    if name == "deg1":
        sc = synth.trained_like_scene(20_000, 21, s0=0.03, sh_degree=1)
        return sc, synth.orbit_cameras(2, 128, 72, 21)[1], o
    if name == "deg0":
        sc = synth.trained_like_scene(20_000, 20, s0=0.03, sh_degree=0)
        return sc, synth.orbit_cameras(2, 128, 72, 20)[0], o
    if name == "colors":
        sc = synth.trained_like_scene(20_000, 31, s0=0.03)
        sc["colors_precomp"] = np.random.default_rng(31).random((20_000, 3)).astype(np.float32)
        return sc, synth.orbit_cameras(2, 144, 80, 31)[1], dict(o, mode="colors")
    if name == "cov3d":
        sc = synth.trained_like_scene(20_000, 32, s0=0.03)
        return sc, synth.orbit_cameras(2, 144, 80, 32)[0], dict(o, mode="cov3d")
    if name == "lego800":
        return synth.random_init_scene(100_000, 1001), _lego_cam(1), o
    if name == "lego800_white":
        return synth.random_init_scene(100_000, 1001), _lego_cam(3), dict(o, bg=(1.0, 1.0, 1.0))
    if name == "mid640x360":
        return synth.trained_like_scene(400_000, 640, s0=0.008), synth.orbit_cameras(4, 640, 360, 640)[2], o
    # ---- arguments of the boundary that round 1 left untested ----
    if name in ("scale_half", "scale_two"):      # the viewer path passes scaling_modifier (train.py:79-80)
        sc = synth.trained_like_scene(20_000, 41, s0=0.03)
        return sc, synth.orbit_cameras(2, 144, 80, 41)[1], dict(o, scale_modifier=0.5 if name == "scale_half" else 2.0)
    if name == "deg2":                           # active degree 2, M = 16 rows (training between iterations 2000-3000)
        sc = synth.trained_like_scene(20_000, 22, s0=0.03, sh_degree=2)
        return sc, synth.orbit_cameras(2, 128, 72, 22)[0], o
    if name in ("m1", "m4", "m9"):               # narrower SH tensors: the non-SH16 branches of preprocess fwd / bwd
        M = int(name[1:])
        D = {1: 0, 4: 1, 9: 2}[M]
        sc = synth.trained_like_scene(20_000, 50 + M, s0=0.03, sh_degree=D)
        sc["shs"] = np.ascontiguousarray(sc["shs"][:, :M])
        return sc, synth.orbit_cameras(2, 136, 88, 50 + M)[1], dict(o, bg=(0.2, 0.4, 0.1))
    if name in ("quat_small", "quat_big"):       # forward.cu:123 uses the quaternion un-normalised
        sc = synth.trained_like_scene(20_000, 61, s0=0.03)
        sc["rotations"] = (sc["rotations"] * (0.3 if name == "quat_small" else 3.0)).astype(np.float32)
        if name == "quat_big":
            sc["scales"] = (sc["scales"] * 0.2).astype(np.float32)      # |q|^2 = 9 stretches the covariance ~ 81x
        return sc, synth.orbit_cameras(2, 144, 80, 61)[0], o
    if name == "deep_list":
        # >= 2000 blended contributors per pixel: large, faint Gaussians (alpha just above 1/255, so T stays above
        # 1e-4 for thousands of layers) stacked along the view axis -- the longest T-recovery chain of the backward
        # blend (backward.cu:573 divides, render_bwd.cu multiplies by an approximate reciprocal)
        rng = np.random.default_rng(77)
        P = 3000
        sc = dict(means3D=np.stack([rng.normal(0, 0.02, P), rng.normal(0, 0.02, P), rng.uniform(-1.0, 1.0, P)], 1).astype(np.float32),
                  scales=np.full((P, 3), 8.0, np.float32) * np.exp(rng.normal(0, 0.05, (P, 3))).astype(np.float32),
                  rotations=np.tile(np.array([[1.0, 0, 0, 0]], np.float32), (P, 1)),
                  opacities=rng.uniform(0.00420, 0.00445, (P, 1)).astype(np.float32),
                  shs=rng.normal(0, 0.5, (P, 16, 3)).astype(np.float32), sh_degree=3)
        cam = synth.Camera(synth.look_at_w2c([0.0, -0.001, -4.0], [0.0, 0.0, 0.0], up=(0.0, -1.0, 0.0)), 48, 32, 0.5)
        return sc, cam, dict(o, bg=(0.5, 0.5, 0.5))
    if name == "empty_view":     # camera looks away: every Gaussian culled, num_rendered = 0
        sc = synth.trained_like_scene(2_000, 99, s0=0.05)
        sc["means3D"] = (sc["means3D"] * 0.05).astype(np.float32)
        cam = synth.Camera(synth.look_at_w2c([5.0, 0, 0], [10.0, 0, 0]), 64, 48, 0.9)
        return sc, cam, o
    raise KeyError(name)


SMALL_CASES = ["one", "two", "tiny64", "odd167x101", "aa160x96", "deg1", "deg0", "colors", "cov3d", "empty_view"]
ARG_CASES = ["scale_half", "scale_two", "deg2", "m1", "m4", "m9", "quat_small", "quat_big", "deep_list"]
MEDIUM_CASES = ["lego800", "lego800_white", "mid640x360"]
GOLDEN_CASES = ["two", "tiny64", "odd167x101", "aa160x96", "colors", "cov3d"]


def cov3d_of(scene):
    """Python-side covariance (scene/gaussian_model.py:56-61 build_covariance_from_scaling_rotation)."""
    s, q = scene["scales"].astype(np.float64), scene["rotations"].astype(np.float64)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                  2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                  2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)
    L = R * s[:, None, :]
    S = L @ L.transpose(0, 2, 1)
    return np.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).astype(np.float32)


def inputs_of(scene, cam, opts):
    """Common keyword set (numpy) for the three implementations."""
    kw = dict(means3D=scene["means3D"], opacities=scene["opacities"], viewmatrix=cam.world_view_transform,
              projmatrix=cam.full_proj_transform, cam_pos=cam.camera_center,
              background=np.array(opts["bg"], np.float32), tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy,
              sh_degree=scene["sh_degree"], antialiasing=opts["antialiasing"])
    if opts.get("scale_modifier", 1.0) != 1.0:
        kw["scale_modifier"] = float(opts["scale_modifier"])
    if opts["mode"] == "colors":
        kw["colors_precomp"] = scene["colors_precomp"]
    else:
        kw["shs"] = scene["shs"]
    if opts["mode"] == "cov3d":
        kw["cov3D_precomp"] = cov3d_of(scene)
    else:
        kw["scales"], kw["rotations"] = scene["scales"], scene["rotations"]
    return kw


FWD_KEYS = ["radii", "tiles_touched", "depths", "means2D", "conic_opacity", "rgb", "cov3D", "clamped",
            "point_offsets", "keys", "point_list", "ranges", "out_color", "final_T", "n_contrib", "invdepth"]
BWD_KEYS = ["dL_dmeans2D", "dL_dconic", "dL_dopacity", "dL_dcolors", "dL_dinvdepths", "dL_dmeans3D",
            "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations"]


def run_oracle(scene, cam, opts, grads=None):
    from oracle import gs_oracle as go
    kw = inputs_of(scene, cam, opts)
    S = go.Scene(W=cam.image_width, H=cam.image_height, **kw)
    f = go.forward(S)
    out = {k: f[k] for k in FWD_KEYS}
    out["cov3D"] = np.where((f["radii"] > 0)[:, None], f["cov3D"], 0.0).astype(np.float32)  # defined for visible rows only
    out["num_rendered"] = f["num_rendered"]
    out["borderline"] = f["borderline"].astype(bool)
    out["pairs_evaluated"], out["pairs_blended"] = f["pairs_evaluated"], f["pairs_blended"]
    if grads is not None:
        g = go.backward(S, f, grads[0], grads[1])
        out.update({"dL_dmeans2D": g["dL_dmean2D"], "dL_dconic": g["dL_dconic"].reshape(-1, 2, 2),
                    "dL_dopacity": g["dL_dopacity"], "dL_dcolors": g["dL_dcolors"], "dL_dinvdepths": g["dL_dinvdepths"],
                    "dL_dmeans3D": g["dL_dmeans3D"], "dL_dcov3D": g["dL_dcov3D"], "dL_dsh": g["dL_dsh"],
                    "dL_dscales": g["dL_dscales"], "dL_drotations": g["dL_drotations"]})
    return out


def _to_dev(kw):
    import torch
    out = {}
    for k, v in kw.items():
        out[k] = torch.from_numpy(np.ascontiguousarray(v)).cuda() if isinstance(v, np.ndarray) else v
    return out


def _np(t):
    return None if t is None else t.detach().cpu().numpy()


def run_cuda(scene, cam, opts, grads=None):
    """This repo's kernels through the C ABI (gsr_b200.capi)."""
    import torch
    from gsr_b200 import capi
    kw = _to_dev(inputs_of(scene, cam, opts))
    st = capi.forward(width=cam.image_width, height=cam.image_height, **kw)
    torch.cuda.synchronize()
    it = capi.introspect(st)
    vis = _np(st["radii"]) > 0
    out = dict(num_rendered=st["num_rendered"], radii=_np(st["radii"]), out_color=_np(st["out_color"]),
               invdepth=_np(st["out_invdepth"]))
    for k in ("tiles_touched", "point_offsets", "point_list", "ranges", "n_contrib"):
        out[k] = _np(it[k]).view(np.uint32)
    out["keys"] = _np(it["keys"]).view(np.uint64)
    out["final_T"] = _np(it["final_T"])
    # per-Gaussian arrays are only defined for visible Gaussians: zero the rest like the oracle
    for k in ("depths", "means2D", "conic_opacity", "rgb", "cov3D", "clamped"):
        a = _np(it[k]).copy()
        a[~vis] = 0
        out[k] = a
    if opts["mode"] == "colors":
        out["rgb"] = np.zeros_like(out["rgb"]); out["clamped"] = np.zeros_like(out["clamped"])
    if opts["mode"] == "cov3d":
        out["cov3D"] = np.zeros_like(out["cov3D"])
    if grads is not None:
        gkw = {k: v for k, v in kw.items()}
        g = capi.backward(st, dL_dcolor=torch.from_numpy(grads[0]).cuda(),
                          dL_dinvdepth=None if grads[1] is None else torch.from_numpy(grads[1]).cuda(), **gkw)
        torch.cuda.synchronize()
        for k in BWD_KEYS:
            out[k] = _np(g[k])
    return out


def run_ref(scene, cam, opts, grads=None):
    """The unmodified reference CUDA rasterizer (oracle/_ref)."""
    import torch
    from oracle import ref_cuda
    kw = _to_dev(inputs_of(scene, cam, opts))
    R = ref_cuda.RefRasterizer()
    f = R.forward(width=cam.image_width, height=cam.image_height, **kw)
    it = R.intermediates()
    vis = _np(f["radii"]) > 0
    out = dict(num_rendered=f["num_rendered"], radii=_np(f["radii"]), out_color=_np(f["out_color"]),
               invdepth=_np(f["out_invdepth"]))
    for k in ("tiles_touched", "point_offsets", "point_list", "ranges", "n_contrib"):
        out[k] = _np(it[k]).view(np.uint32)
    out["keys"] = _np(it["keys"]).view(np.uint64)
    out["final_T"] = _np(it["final_T"])
    for k in ("depths", "means2D", "conic_opacity", "rgb", "cov3D", "clamped"):
        a = _np(it[k]).copy()
        a[~vis] = 0          # the reference leaves culled rows uninitialised
        out[k] = a
    if opts["mode"] == "colors":
        out["rgb"] = np.zeros_like(out["rgb"]); out["clamped"] = np.zeros_like(out["clamped"])
    if opts["mode"] == "cov3d":
        out["cov3D"] = np.zeros_like(out["cov3D"])
    T = ((cam.image_width + 15) // 16) * ((cam.image_height + 15) // 16)
    out["ranges"] = out["ranges"][:T]
    if grads is not None:
        g = R.backward(f["radii"], dL_dcolor=torch.from_numpy(grads[0]).cuda(),
                       dL_dinvdepth=None if grads[1] is None else torch.from_numpy(grads[1]).cuda(), **kw)
        for k in BWD_KEYS:
            out[k] = _np(g[k])
    R.close()
    return out


def grads_for(cam, seed=5):
    return synth.upstream_grads(cam.image_height, cam.image_width, seed)


INT_KEYS = ["radii", "tiles_touched", "point_offsets", "keys", "point_list", "ranges", "n_contrib"]
BITEXACT_FLOAT_KEYS = ["depths", "means2D", "conic_opacity"]


def compare(a, b, borderline=None, grad_rtol=1e-4):
    """Mismatch report between two result dicts (a = under test, b = reference).  grad_rtol: one tolerance, or a
    dict per gradient key (the elementwise `_viol` count uses it)."""
    rep = {}
    rep["num_rendered"] = (a["num_rendered"], b["num_rendered"])
    for k in INT_KEYS:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        if x.shape != y.shape:
            rep[k] = "shape %s vs %s" % (x.shape, y.shape)
            continue
        bad = x != y
        if k == "n_contrib" and borderline is not None:
            rep[k + "_borderline_mismatch"] = int((bad & borderline).sum())
            bad = bad & ~borderline
        rep[k] = int(bad.sum())
    for k in BITEXACT_FLOAT_KEYS:
        x, y = np.asarray(a[k]), np.asarray(b[k])
        rep[k + "_bits"] = int((x.view(np.uint32) != y.view(np.uint32)).sum())
    for k in ("rgb", "cov3D", "out_color", "final_T", "invdepth"):
        x, y = np.asarray(a[k], np.float64), np.asarray(b[k], np.float64)
        d = np.abs(x - y)
        if borderline is not None and k in ("out_color", "final_T", "invdepth"):
            m = np.broadcast_to(borderline.reshape((1,) * (d.ndim - 2) + borderline.shape), d.shape)
            rep[k + "_maxabs_borderline"] = float(d[m].max()) if m.any() else 0.0
            d = np.where(m, 0.0, d)
        rep[k + "_maxabs"] = float(d.max()) if d.size else 0.0
    rep["clamped"] = int((np.asarray(a["clamped"]) != np.asarray(b["clamped"])).sum())
    for k in BWD_KEYS:
        if k in a and k in b and a[k] is not None and b[k] is not None:
            x, y = np.asarray(a[k], np.float64), np.asarray(b[k], np.float64)
            if x.size == 0:
                continue
            scale = np.abs(y).max()
            if scale < 1e-12:
                # analytically-zero gradient (e.g. rotation of an isotropic Gaussian): only noise to compare
                rep[k + "_absmax"] = float(np.abs(x - y).max())
                continue
            rep[k + "_relmax"] = float(np.abs(x - y).max() / scale)
            # elementwise: |dx| <= rtol*|y| + 0.1*rtol*max|y|   (GRAD_RTOL = 1e-4, north_star)
            rt = grad_rtol.get(k, 1e-4) if isinstance(grad_rtol, dict) else grad_rtol
            tol = rt * np.abs(y) + 0.1 * rt * scale
            rep[k + "_viol"] = int((np.abs(x - y) > tol).sum())
            rep[k + "_size"] = int(x.size)
    return rep
