"""Drop-in behaviour of the public Python package on a GPU: autograd path, markVisible, empty and
degenerate inputs, debug flag, stream/device handling."""
import numpy as np
import pytest

import cases
from test_oracle_golden import GRAD_TOL

pytestmark = pytest.mark.gpu


def _settings(cam, opts, scene, dev, debug=False):
    import torch
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return GaussianRasterizationSettings(
        image_height=cam.image_height, image_width=cam.image_width, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=t(np.array(opts["bg"], np.float32)), scale_modifier=1.0, viewmatrix=t(cam.world_view_transform),
        projmatrix=t(cam.full_proj_transform), sh_degree=scene["sh_degree"], campos=t(cam.camera_center),
        prefiltered=False, debug=debug, antialiasing=opts["antialiasing"])


@pytest.mark.parametrize("name", ["tiny64", "aa160x96", "colors", "cov3d"])
def test_autograd_path_matches_oracle(cuda, name):
    import torch
    from diff_gaussian_rasterization import GaussianRasterizer
    scene, cam, opts = cases.get_case(name)
    grads = cases.grads_for(cam)
    o = cases.run_oracle(scene, cam, opts, grads)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda).requires_grad_(True)
    means3D, opac = t(scene["means3D"]), t(scene["opacities"])
    means2D = torch.zeros_like(means3D, requires_grad=True)
    kw, leaves = {}, {"dL_dmeans3D": means3D, "dL_dopacity": opac, "dL_dmeans2D": means2D}
    if opts["mode"] == "colors":
        kw["colors_precomp"] = leaves["dL_dcolors"] = t(scene["colors_precomp"])
    else:
        kw["shs"] = leaves["dL_dsh"] = t(scene["shs"])
    if opts["mode"] == "cov3d":
        kw["cov3D_precomp"] = leaves["dL_dcov3D"] = t(cases.cov3d_of(scene))
    else:
        kw["scales"] = leaves["dL_dscales"] = t(scene["scales"])
        kw["rotations"] = leaves["dL_drotations"] = t(scene["rotations"])
    color, radii, invd = GaussianRasterizer(_settings(cam, opts, scene, cuda))(means3D=means3D, means2D=means2D,
                                                                                opacities=opac, **kw)
    assert color.shape == (3, cam.image_height, cam.image_width) and invd.shape == (1, cam.image_height, cam.image_width)
    assert radii.dtype == torch.int32 and np.array_equal(radii.cpu().numpy(), o["radii"])
    ok = ~o["borderline"]
    assert np.abs(color.detach().cpu().numpy() - o["out_color"])[:, ok].max() <= 1e-5
    assert np.abs(invd.detach().cpu().numpy() - o["invdepth"])[:, ok].max() <= 1e-5
    (color * torch.from_numpy(grads[0]).to(cuda)).sum().add((invd * torch.from_numpy(grads[1]).to(cuda)).sum()).backward()
    for k, leaf in leaves.items():
        want = o[k].astype(np.float64)
        got = leaf.grad.cpu().numpy().astype(np.float64).reshape(want.shape)
        if np.abs(want).max() < 1e-12:
            continue
        assert np.abs(got - want).max() / np.abs(want).max() <= GRAD_TOL[k], k
    assert means2D.grad.shape == (scene["means3D"].shape[0], 3) and float(means2D.grad[:, 2].abs().max()) == 0.0


def test_mark_visible_and_no_grad_render(cuda):
    import torch
    from diff_gaussian_rasterization import GaussianRasterizer
    from oracle import gs_oracle as go
    scene, cam, opts = cases.get_case("odd167x101")
    r = GaussianRasterizer(_settings(cam, opts, scene, cuda))
    vis = r.markVisible(torch.from_numpy(scene["means3D"]).to(cuda))
    assert vis.dtype == torch.bool
    assert np.array_equal(vis.cpu().numpy(), go.mark_visible(scene["means3D"], cam.world_view_transform))
    with torch.no_grad():   # render.py path: forward only, scratch never handed back
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda)
        color, radii, invd = r(means3D=t(scene["means3D"]), means2D=torch.zeros(len(scene["means3D"]), 3, device=cuda),
                               opacities=t(scene["opacities"]), shs=t(scene["shs"]), scales=t(scene["scales"]),
                               rotations=t(scene["rotations"]))
    assert bool(torch.isfinite(color).all())


def test_empty_inputs_and_fully_culled_view(cuda):
    import torch
    from diff_gaussian_rasterization import GaussianRasterizer
    scene, cam, opts = cases.get_case("empty_view")
    rs = _settings(cam, dict(opts, bg=(0.25, 0.5, 0.75)), scene, cuda)
    z = lambda *s: torch.zeros(*s, device=cuda)
    # P = 0: the reference skips every kernel and returns the zero-initialised image (rasterize_points.cu:88)
    color, radii, invd = GaussianRasterizer(rs)(means3D=z(0, 3), means2D=z(0, 3), opacities=z(0, 1), shs=z(0, 16, 3),
                                                scales=z(0, 3), rotations=z(0, 4))
    assert radii.numel() == 0 and float(color.abs().max()) == 0.0
    # every Gaussian culled: num_rendered = 0, image = background, all gradients zero
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda).requires_grad_(True)
    m3, sh, op, sc, ro = t(scene["means3D"]), t(scene["shs"]), t(scene["opacities"]), t(scene["scales"]), t(scene["rotations"])
    color, radii, invd = GaussianRasterizer(rs)(means3D=m3, means2D=torch.zeros_like(m3, requires_grad=True), opacities=op,
                                                shs=sh, scales=sc, rotations=ro)
    assert int(radii.abs().max()) == 0
    for ch, v in enumerate((0.25, 0.5, 0.75)):
        assert float((color[ch] - v).abs().max()) == 0.0
    color.sum().backward()
    for p in (m3, sh, op, sc, ro):
        assert float(p.grad.abs().max()) == 0.0


def test_debug_flag_and_side_stream(cuda):
    import torch
    from diff_gaussian_rasterization import GaussianRasterizer
    scene, cam, opts = cases.get_case("tiny64")
    o = cases.run_oracle(scene, cam, opts)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda)
    args = dict(means3D=t(scene["means3D"]), means2D=torch.zeros(1000, 3, device=cuda), opacities=t(scene["opacities"]),
                shs=t(scene["shs"]), scales=t(scene["scales"]), rotations=t(scene["rotations"]))
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    with torch.cuda.stream(s):      # launches follow torch's current stream (reference: legacy default stream)
        color, radii, _ = GaussianRasterizer(_settings(cam, opts, scene, cuda, debug=True))(**args)
    s.synchronize()
    assert np.array_equal(radii.cpu().numpy(), o["radii"])
    assert np.abs(color.cpu().numpy() - o["out_color"])[:, ~o["borderline"]].max() <= 1e-5


def test_prefiltered_trap_becomes_an_error(cuda):
    import torch
    from diff_gaussian_rasterization import GaussianRasterizer
    scene, cam, opts = cases.get_case("tiny64")      # contains Gaussians behind the camera
    rs = _settings(cam, opts, scene, cuda)._replace(prefiltered=True)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda)
    with pytest.raises(RuntimeError, match="prefiltered"):
        GaussianRasterizer(rs)(means3D=t(scene["means3D"]), means2D=torch.zeros(1000, 3, device=cuda),
                               opacities=t(scene["opacities"]), shs=t(scene["shs"]), scales=t(scene["scales"]),
                               rotations=t(scene["rotations"]))


@pytest.mark.parametrize("name", ["tiny64", "aa160x96"])
def test_accumulate_mode_sums_views(cuda, name):
    """gsr_backward_args.accumulate: view 0 written + view 1 accumulated into the same buffers equals the
    sum of two separate write-mode backwards; rows invisible in view 1 keep view 0's values bit-exactly
    (what autograd does for the reference when one step renders several cameras)."""
    import torch
    from gsr_b200 import capi
    scene, cam, opts = cases.get_case(name)
    grads = cases.grads_for(cam)
    kw0 = cases._to_dev(cases.inputs_of(scene, cam, opts))
    # second view: same camera looking at a scene shifted sideways -> a different visible set
    scene1 = dict(scene); scene1["means3D"] = scene["means3D"] + np.array([0.35, 0.0, 0.0], np.float32)
    kw1 = cases._to_dev(cases.inputs_of(scene1, cam, opts))
    dcol = torch.from_numpy(grads[0]).cuda(); ddep = torch.from_numpy(grads[1]).cuda()
    keys = ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations")
    sep = []
    for kw in (kw0, kw1):
        st = capi.forward(width=cam.image_width, height=cam.image_height, **kw)
        g = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
        sep.append(({k: g[k].clone() for k in keys}, st["radii"].clone()))
    P = scene["means3D"].shape[0]
    M = scene["shs"].shape[1]
    bucket = dict(dL_dmeans3D=torch.full((P, 3), 7.0, device=cuda), dL_dsh=torch.full((P, M, 3), 7.0, device=cuda),
                  dL_dopacity=torch.full((P, 1), 7.0, device=cuda), dL_dscales=torch.full((P, 3), 7.0, device=cuda),
                  dL_drotations=torch.full((P, 4), 7.0, device=cuda))
    st = capi.forward(width=cam.image_width, height=cam.image_height, **kw0)
    capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=bucket, **kw0)                      # overwrites the 7s
    st = capi.forward(width=cam.image_width, height=cam.image_height, **kw1)
    capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=bucket, accumulate=True, **kw1)
    torch.cuda.synchronize()
    vis1 = (sep[1][1] > 0).cpu().numpy()
    assert vis1.any() and (~vis1).any()
    for k in keys:
        want = (sep[0][0][k].double() + sep[1][0][k].double()).cpu().numpy().reshape(P, -1)
        got = bucket[k].double().cpu().numpy().reshape(P, -1)
        scale = np.abs(want).max() + 1e-30
        assert np.abs(got - want).max() / scale <= 1e-4, k
        first = sep[0][0][k].cpu().numpy().reshape(P, -1)
        # atomic-order noise only enters through rows view 1 touched
        assert np.abs(got[~vis1] - first[~vis1]).max() / scale <= 1e-4, k


def test_separate_sh_arguments_dc_plus_rest(cuda):
    """gaussian_renderer/__init__.py:90-100 (separate_sh): rasterizer(dc=[P,1,3], shs=[P,15,3]) renders the same bits as
    one [P,16,3] tensor and autograd splits dL_dsh between the two leaves."""
    import torch
    from diff_gaussian_rasterization import GaussianRasterizer
    scene, cam, opts = cases.get_case("tiny64")
    g0, g1 = cases.grads_for(cam)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda).requires_grad_(True)
    R = GaussianRasterizer(_settings(cam, opts, scene, cuda))
    res = {}
    for split in (False, True):
        means3D, opac, sc, rot = t(scene["means3D"]), t(scene["opacities"]), t(scene["scales"]), t(scene["rotations"])
        means2D = torch.zeros_like(means3D, requires_grad=True)
        if split:
            dc, rest = t(scene["shs"][:, :1]), t(scene["shs"][:, 1:])
            color, radii, invd = R(means3D=means3D, means2D=means2D, dc=dc, shs=rest, colors_precomp=None, opacities=opac,
                                   scales=sc, rotations=rot, cov3D_precomp=None)
        else:
            shs = t(scene["shs"])
            color, radii, invd = R(means3D=means3D, means2D=means2D, shs=shs, opacities=opac, scales=sc, rotations=rot)
        (color * torch.from_numpy(g0).to(cuda)).sum().backward()
        res[split] = (color.detach(), radii, torch.cat([dc.grad, rest.grad], 1) if split else shs.grad, means3D.grad)
    assert torch.equal(res[False][0], res[True][0]) and torch.equal(res[False][1], res[True][1])
    for a, b in zip(res[False][2:], res[True][2:]):           # gradients: fp32 RED order differs from run to run
        assert (a - b).abs().max().item() <= 1e-4 * a.abs().max().item()


@pytest.mark.parametrize("n_streams", [2, 3])
def test_view_streams_match_single_stream(cuda, n_streams):
    """gsr_b200.dist.ViewStreams: the views of a step spread over several CUDA streams give the same bucket as the
    sequential loop (view 0 overwrites, the others add; only the order of the fp32 adds differs)."""
    import torch
    from gsr_b200 import capi
    from gsr_b200.dist import ViewStreams
    scene, cam, opts = cases.get_case("odd167x101")
    grads = cases.grads_for(cam)
    dcol = torch.from_numpy(grads[0]).cuda(); ddep = torch.from_numpy(grads[1]).cuda()
    views = []
    for k in range(5):
        sc = dict(scene); sc["means3D"] = scene["means3D"] + np.array([0.2 * k - 0.4, 0.05 * k, 0.0], np.float32)
        views.append(cases._to_dev(cases.inputs_of(sc, cam, opts)))
    P, M = scene["means3D"].shape[0], scene["shs"].shape[1]
    W, H = cam.image_width, cam.image_height

    def run(S):
        bucket = dict(dL_dmeans3D=torch.full((P, 3), 7.0, device=cuda), dL_dsh=torch.full((P, M, 3), 7.0, device=cuda),
                      dL_dopacity=torch.full((P, 1), 7.0, device=cuda), dL_dscales=torch.full((P, 3), 7.0, device=cuda),
                      dL_drotations=torch.full((P, 4), 7.0, device=cuda))
        vs = ViewStreams(cuda, S)
        for _ in range(3):                                   # repeated steps reuse the streams
            vs.run(len(views), lambda v: capi.forward(width=W, height=H, **views[v]),
                   lambda v, st, acc: capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=bucket, accumulate=acc, **views[v]))
        torch.cuda.synchronize()
        return {k: b.double().cpu().numpy() for k, b in bucket.items()}
    want, got = run(1), run(n_streams)
    # dL_dscales / dL_drotations come out of the covariance backward with heavy cancellation: the order of the fp32 atomics
    # of the blend (different from run to run on ONE stream as well) moves them by ~1e-4 of the tensor maximum -- the same
    # three tensors tests/test_oracle_golden.py holds to 1e-3 for the reference's own noise
    tol = dict(dL_dscales=1e-3, dL_drotations=1e-3)
    for k in want:
        assert np.abs(got[k] - want[k]).max() <= tol.get(k, 1e-4) * (np.abs(want[k]).max() + 1e-30), k


def test_workspace_reuse_gives_the_same_render(cuda):
    """capi.Workspace: growth-only scratch reused across forwards (big view, then a small one, then the big one again)
    renders the same bits as per-call scratch, and backward works from the aliased state."""
    import torch
    from gsr_b200 import capi
    ws = capi.Workspace(cuda)
    grads = None
    for name in ("odd167x101", "tiny64", "odd167x101"):
        scene, cam, opts = cases.get_case(name)
        kw = cases._to_dev(cases.inputs_of(scene, cam, opts))
        a = capi.forward(width=cam.image_width, height=cam.image_height, **kw)
        b = capi.forward(width=cam.image_width, height=cam.image_height, workspace=ws, **kw)
        assert a["num_rendered"] == b["num_rendered"]
        assert torch.equal(a["out_color"], b["out_color"]) and torch.equal(a["radii"], b["radii"])
        g = cases.grads_for(cam)
        dcol, ddep = torch.from_numpy(g[0]).cuda(), torch.from_numpy(g[1]).cuda()
        ga = capi.backward(a, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
        gb = capi.backward(b, dL_dcolor=dcol, dL_dinvdepth=ddep, **kw)
        for k in ("dL_dmeans3D", "dL_dsh", "dL_dopacity"):
            assert (ga[k] - gb[k]).abs().max().item() <= 1e-4 * (ga[k].abs().max().item() + 1e-30), (name, k)
    assert ws.binning.t.numel() > 0


@pytest.mark.parametrize("n_views", [3, 10])
def test_grad_sink_immediate_and_deferred_equal_autograd_sum(cuda, n_views):
    """diff_gaussian_rasterization.grad_sink: the parameter gradients of the views of a step written into caller-owned
    tensors -- view by view (accumulate) or, with defer=True, by ONE batched per-Gaussian launch per 8 views when the
    context closes -- equal what plain autograd accumulates in .grad; means2D gradients keep flowing through autograd."""
    import torch
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, grad_sink
    from gsr_b200 import synth
    P, W, H = 20_000, 167, 101
    scene = synth.trained_like_scene(P, 31, s0=0.03)
    cams = synth.orbit_cameras(max(8, n_views), W, H, 31)[:n_views]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda)
    names = ("means3D", "shs", "opacities", "scales", "rotations")
    rng = np.random.default_rng(7)
    weights = [t(rng.standard_normal((3, H, W)).astype(np.float32)) for _ in cams]
    bg = torch.tensor([0.2, 0.1, 0.3], device=cuda)

    def render_all(leaves, ctx_for):
        m2d = []
        for v, c in enumerate(cams):
            rs = GaussianRasterizationSettings(
                image_height=H, image_width=W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, bg=bg, scale_modifier=1.0,
                viewmatrix=t(c.world_view_transform), projmatrix=t(c.full_proj_transform), sh_degree=3, campos=t(c.camera_center),
                prefiltered=False, debug=False, antialiasing=False)
            means2D = torch.zeros(P, 3, device=cuda, requires_grad=True)
            color, radii, invd = GaussianRasterizer(rs)(means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
                                                        shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
            loss = (color * weights[v]).sum() + 0.3 * invd.sum()
            with ctx_for(v):
                loss.backward()
            m2d.append(means2D)
        return m2d

    import contextlib
    # plain autograd: .grad accumulates over the views
    ref = {k: t(scene[k]).requires_grad_(True) for k in names}
    m2d_ref = render_all(ref, lambda v: contextlib.nullcontext())
    want = {k: ref[k].grad.double() for k in names}
    tol = dict(scales=1e-3, rotations=1e-3)

    def check(sinks, m2d):
        for k in names:
            err = (sinks[k].double() - want[k]).abs().max().item() / (want[k].abs().max().item() + 1e-30)
            assert err <= tol.get(k, 1e-5), (k, err)
        for a, b in zip(m2d, m2d_ref):
            assert (a.grad - b.grad).abs().max().item() <= 1e-5 * (b.grad.abs().max().item() + 1e-30)

    # immediate sink, one context per view
    leaves = {k: t(scene[k]).requires_grad_(True) for k in names}
    sinks = {k: torch.full_like(leaves[k], 9.0) for k in names}
    m2d = render_all(leaves, lambda v: grad_sink(*(sinks[k] for k in names), accumulate=(v > 0)))
    assert all(leaves[k].grad is None for k in names)
    check(sinks, m2d)

    # deferred sink, one context around the whole step (n_views = 10: two batched launches, the second one adds)
    leaves = {k: t(scene[k]).requires_grad_(True) for k in names}
    sinks = {k: torch.full_like(leaves[k], 9.0) for k in names}
    with grad_sink(*(sinks[k] for k in names), accumulate=False, defer=True):
        m2d = render_all(leaves, lambda v: contextlib.nullcontext())
    assert all(leaves[k].grad is None for k in names)
    check(sinks, m2d)
    from diff_gaussian_rasterization import _C
    assert _C.flush_grad_sink() == 0                      # nothing left parked


def test_deferred_grad_sink_with_fresh_parameter_tensors_and_a_side_stream(cuda):
    """grad_sink(defer=True) when the views do NOT share their parameter tensors (a caller that recomputes the activations
    per render, as the reference's GaussianModel properties do) and when backward() runs under another stream than the one
    the context closes on: views that cannot share a batched launch are written out one by one instead of failing at the
    flush, the write-out waits for the blends, and the sink still holds the autograd sum."""
    import contextlib
    import torch
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, grad_sink, _C
    from gsr_b200 import synth
    P, W, H, n_views = 20_000, 167, 101, 4
    scene = synth.trained_like_scene(P, 33, s0=0.03)
    cams = synth.orbit_cameras(8, W, H, 33)[:n_views]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda)
    names = ("means3D", "shs", "opacities", "scales", "rotations")
    rng = np.random.default_rng(9)
    weights = [t(rng.standard_normal((3, H, W)).astype(np.float32)) for _ in cams]
    bg = torch.tensor([0.2, 0.1, 0.3], device=cuda)
    alive = []                                            # keeps the per-view copies alive so their addresses differ

    def render_all(leaves, fresh):
        for v, c in enumerate(cams):
            rs = GaussianRasterizationSettings(
                image_height=H, image_width=W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, bg=bg, scale_modifier=1.0,
                viewmatrix=t(c.world_view_transform), projmatrix=t(c.full_proj_transform), sh_degree=3, campos=t(c.camera_center),
                prefiltered=False, debug=False, antialiasing=False)
            args = dict(leaves)
            for k in fresh(v):
                args[k] = leaves[k].detach().clone().requires_grad_(True)
                alive.append(args[k])
            means2D = torch.zeros(P, 3, device=cuda, requires_grad=True)
            color, radii, invd = GaussianRasterizer(rs)(means3D=args["means3D"], means2D=means2D, opacities=args["opacities"],
                                                        shs=args["shs"], scales=args["scales"], rotations=args["rotations"])
            ((color * weights[v]).sum() + 0.3 * invd.sum()).backward()

    ref = {k: t(scene[k]).requires_grad_(True) for k in names}
    render_all(ref, lambda v: ())
    want = {k: ref[k].grad.double() for k in names}
    tol = dict(scales=1e-3, rotations=1e-3)

    def check(sinks):
        torch.cuda.synchronize()
        for k in names:
            err = (sinks[k].double() - want[k]).abs().max().item() / (want[k].abs().max().item() + 1e-30)
            assert err <= tol.get(k, 1e-5), (k, err)

    # (1) opacities / scales / rotations are new tensors in views 1 and 3, shs in view 2: four write-outs, no error
    leaves = {k: t(scene[k]).requires_grad_(True) for k in names}
    sinks = {k: torch.full_like(leaves[k], 9.0) for k in names}
    pick = {0: (), 1: ("opacities", "scales", "rotations"), 2: ("shs",), 3: ("opacities",)}
    with grad_sink(*(sinks[k] for k in names), accumulate=False, defer=True):
        render_all(leaves, lambda v: pick[v])
    assert _C.flush_grad_sink() == 0
    check(sinks)

    # (2) shared tensors, forward + backward on a side stream, the context closes (and flushes) on the default stream
    leaves = {k: t(scene[k]).requires_grad_(True) for k in names}
    sinks = {k: torch.full_like(leaves[k], 9.0) for k in names}
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with grad_sink(*(sinks[k] for k in names), accumulate=False, defer=True):
        with torch.cuda.stream(side):
            render_all(leaves, lambda v: ())
    check(sinks)
    del alive[:]
