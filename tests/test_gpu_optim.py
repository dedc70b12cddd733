"""GPU parity of the optimizer stage (csrc/adam.cu) through the C ABI (gsr_b200.capi -> libgsr_b200.so) against
  (i) the numpy oracle (oracle/optim_oracle.py), (ii) torch.optim.Adam itself -- the optimizer the reference
  constructs at /root/reference/scene/gaussian_model.py:178-201 -- on this GPU.
Tolerances (fp32): moments <= 1e-6 of their input scale, parameters <= 1e-4 of one step (lr) + fp32 rounding."""
import numpy as np
import pytest

from oracle import optim_oracle as oo
from test_optim_oracle_cpu import GROUPS

pytestmark = pytest.mark.gpu


def _mk(P, seed):
    rng = np.random.default_rng(seed)
    params = {n: rng.standard_normal((P, c)).astype(np.float32) for n, c, _ in GROUPS}
    return rng, params


def _segs(t, cols=None):
    return [dict(param=t["p"][n], grad=t["g"][n], exp_avg=t["m"][n], exp_avg_sq=t["v"][n], lr=lr) for n, _, lr in GROUPS]


@pytest.mark.parametrize("P", [1, 257, 4099])
def test_adam_vs_oracle_and_torch(cuda, P):
    import torch
    from gsr_b200 import capi
    rng, params = _mk(P, P)
    dev = {k: {n: torch.from_numpy(params[n].copy()).to(cuda) if k == "p" else torch.zeros(P, c, device=cuda)
               for n, c, _ in GROUPS} for k in ("p", "g", "m", "v")}
    tp = {n: torch.nn.Parameter(torch.from_numpy(params[n].copy()).to(cuda)) for n, _, _ in GROUPS}
    opt = torch.optim.Adam([{"params": [tp[n]], "lr": lr, "name": n} for n, _, lr in GROUPS], lr=0.0, eps=1e-15)
    m = {n: np.zeros_like(params[n]) for n, _, _ in GROUPS}
    v = {n: np.zeros_like(params[n]) for n, _, _ in GROUPS}
    for step in range(1, 6):
        grads = {n: (rng.standard_normal((P, c)) * 10.0 ** rng.integers(-6, 1)).astype(np.float32) for n, c, _ in GROUPS}
        for n, _, lr in GROUPS:
            dev["g"][n].copy_(torch.from_numpy(grads[n]))
            tp[n].grad = torch.from_numpy(grads[n].copy()).to(cuda)
            oo.adam_step(params[n], grads[n], m[n], v[n], step, lr)
        capi.adam_step(_segs(dev), step)
        opt.step()
        for n, _, lr in GROUPS:
            gmax = np.abs(grads[n]).max()
            for want_p, want_m, want_v in ((params[n], m[n], v[n]),
                                           (tp[n].detach().cpu().numpy(), opt.state[tp[n]]["exp_avg"].cpu().numpy(),
                                            opt.state[tp[n]]["exp_avg_sq"].cpu().numpy())):
                assert np.abs(dev["m"][n].cpu().numpy() - want_m).max() <= 1e-6 * max(gmax, np.abs(want_m).max()), (n, step)
                assert np.abs(dev["v"][n].cpu().numpy() - want_v).max() <= 1e-6 * max(gmax ** 2, np.abs(want_v).max()), (n, step)
                assert np.abs(dev["p"][n].cpu().numpy() - want_p).max() <= 1e-4 * lr + 2e-7 * np.abs(want_p).max(), (n, step)


def test_gaussian_adam_on_flat_bucket_matches_six_group_torch_adam(cuda):
    """The host mirror (gsr_b200.optim.GaussianAdam over FlatGaussians): SH row = f_dc + f_rest with their own rates,
    zero_grad fused, visible mask = sparse step."""
    import torch
    import cases
    from gsr_b200.dist import FlatGaussians
    from gsr_b200.optim import GaussianAdam, DEFAULT_LRS
    scene, cam, opts = cases.get_case("tiny64")
    G = FlatGaussians.from_numpy(scene, cuda)
    P = G.P
    opt = GaussianAdam(G, spatial_lr_scale=5.0)
    ref = {"xyz": G.xyz.clone(), "f_dc": G.shs[:, :1].clone(), "f_rest": G.shs[:, 1:].clone(), "opacity": G.opacity.clone(),
           "scaling": G.scales.clone(), "rotation": G.rotations.clone()}
    tp = {k: torch.nn.Parameter(t.contiguous()) for k, t in ref.items()}
    lrs = dict(DEFAULT_LRS)
    lrs["xyz"] *= 5.0
    topt = torch.optim.Adam([{"params": [tp[k]], "lr": lrs[k], "name": k} for k in tp], lr=0.0, eps=1e-15)
    gen = torch.Generator(device="cuda").manual_seed(1)
    for step in range(3):
        G.bucket.normal_(generator=gen)
        gv = G.grad_views()
        tp["xyz"].grad = gv["dL_dmeans3D"].clone()
        tp["f_dc"].grad = gv["dL_dsh"][:, :1].clone()
        tp["f_rest"].grad = gv["dL_dsh"][:, 1:].clone()
        tp["opacity"].grad = gv["dL_dopacity"].clone()
        tp["scaling"].grad = gv["dL_dscales"].clone()
        tp["rotation"].grad = gv["dL_drotations"].clone()
        opt.step(zero_grad=True)
        topt.step()
        for gview in G.grad_views().values():               # optimizer.zero_grad() fused into the step
            assert float(gview.abs().max()) == 0.0           # (the 64-float padding between segments is not a gradient)
    got = {"xyz": G.xyz, "f_dc": G.shs[:, :1], "f_rest": G.shs[:, 1:], "opacity": G.opacity, "scaling": G.scales,
           "rotation": G.rotations}
    for k in tp:
        d = (got[k] - tp[k].detach()).abs().max().item()
        assert d <= 1e-4 * lrs[k] + 2e-7 * tp[k].abs().max().item(), (k, d)
    # sparse step: rows outside the mask keep parameters and moments bit for bit
    vis = torch.rand(P, device=cuda, generator=gen) < 0.3
    before = (G.params.clone(), opt.exp_avg.clone(), opt.exp_avg_sq.clone())
    G.bucket.normal_(generator=gen)
    opt.step(visible=vis)
    for name, n in (("xyz", 3), ("shs", 48), ("opacity", 1), ("scales", 3), ("rotations", 4)):
        for now, was in zip((G.params, opt.exp_avg, opt.exp_avg_sq), before):
            a, b = G._view(now, name).reshape(P, -1), G._view(was, name).reshape(P, -1)
            assert torch.equal(a[~vis], b[~vis])
            assert not torch.equal(a[vis], b[vis])


def test_untouched_and_decayed_rows(cuda):
    """Zero and denormal moments are the common case (Gaussians no view has touched; moments decaying after a Gaussian
    left every frustum).  The kernel's branch-free sqrt / division must agree with the oracle there and everywhere else:
    compared elementwise on a state that mixes all the ranges."""
    import torch
    from gsr_b200 import capi
    rng = np.random.default_rng(11)
    n = 64 * 1024
    p = rng.standard_normal((n, 1)).astype(np.float32)
    g = (rng.standard_normal((n, 1)) * 10.0 ** rng.integers(-12, 3, (n, 1))).astype(np.float32)
    m = (rng.standard_normal((n, 1)) * 10.0 ** rng.integers(-44, 1, (n, 1))).astype(np.float32)
    v = (rng.random((n, 1)) * 10.0 ** rng.integers(-44, 4, (n, 1))).astype(np.float32)
    kind = rng.integers(0, 4, (n, 1))
    g[kind == 0] = 0.0                                      # not visible this step
    m[kind == 1] = 0.0
    v[kind == 1] = 0.0                                      # never touched
    g[kind == 1] = 0.0
    t = {k: torch.from_numpy(a.copy()).to(cuda) for k, a in dict(p=p, g=g, m=m, v=v).items()}
    for step in (1, 2, 1000):
        p0, m0, v0 = p.copy(), m.copy(), v.copy()
        oo.adam_step(p, g, m, v, step, 0.01)
        capi.adam_step([dict(param=t["p"], grad=t["g"], exp_avg=t["m"], exp_avg_sq=t["v"], lr=0.01)], step)
        got = {k: t[k].cpu().numpy() for k in "pmv"}
        assert all(np.isfinite(a).all() for a in got.values())
        # elementwise: one fp32 rounding of the operands (the kernel fuses multiply-adds the oracle rounds twice)
        g64 = g.astype(np.float64)
        assert (np.abs(got["m"] - m) <= 1.2e-7 * (np.abs(m0) + np.abs(g)) + 2e-45).all(), step
        assert (np.abs(got["v"] - v) <= 1.2e-7 * (v0 + g64 * g64) + 2e-45).all(), step
        assert (np.abs(got["p"] - p) <= 2.4e-7 * np.abs(p0) + 1e-5 * np.abs(p - p0) + 1e-12).all(), step
        untouched = (kind == 1)
        assert np.array_equal(got["p"][untouched], p0[untouched])          # 0 / eps = 0: the parameter does not move
        for k, a in (("p", p), ("m", m), ("v", v)):
            t[k].copy_(torch.from_numpy(a))


def test_sparse_gaussian_adam_interface(cuda):
    """SparseGaussianAdam(l, lr=0.0, eps=1e-15).step(visible, N) as gaussian_model.py:194-198 / train.py:179-181 use it,
    against torch.optim.Adam with the rows outside the mask restored after every step."""
    import torch
    from gsr_b200.optim import SparseGaussianAdam
    N = 5003
    gen = torch.Generator(device="cuda").manual_seed(5)
    shapes = dict(xyz=(N, 3), f_dc=(N, 1, 3), f_rest=(N, 15, 3), opacity=(N, 1), scaling=(N, 3), rotation=(N, 4))
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.025, scaling=0.005, rotation=0.001)
    ours = {k: torch.nn.Parameter(torch.randn(s, device=cuda, generator=gen)) for k, s in shapes.items()}
    # create_from_pcd hands the optimizer a column-major _xyz (torch.tensor keeps the strides of np.vstack([x,y,z]).T)
    ours["xyz"] = torch.nn.Parameter(torch.randn(3, N, device=cuda, generator=gen).t())
    assert not ours["xyz"].is_contiguous()
    ref = {k: torch.nn.Parameter(v.detach().clone()) for k, v in ours.items()}
    opt = SparseGaussianAdam([{"params": [ours[k]], "lr": lrs[k], "name": k} for k in ours], lr=0.0, eps=1e-15)
    topt = torch.optim.Adam([{"params": [ref[k]], "lr": lrs[k], "name": k} for k in ref], lr=0.0, eps=1e-15)
    assert isinstance(opt, torch.optim.Adam)
    for step in range(4):
        vis = torch.rand(N, device=cuda, generator=gen) < 0.4
        for k in ours:
            g = torch.randn(shapes[k], device=cuda, generator=gen)
            ours[k].grad, ref[k].grad = g.clone(), g.clone()
        before = {k: (ref[k].detach().clone(), topt.state[ref[k]].get("exp_avg", torch.zeros_like(ref[k])).clone(),
                      topt.state[ref[k]].get("exp_avg_sq", torch.zeros_like(ref[k])).clone()) for k in ref}
        mine = {k: (ours[k].detach().clone(), opt.state[ours[k]].get("exp_avg", torch.zeros_like(ours[k])).clone()) for k in ours}
        opt.step(vis, N)
        topt.step()
        for k in ref:                                          # rows outside the mask: as if the step had not happened
            st = topt.state[ref[k]]
            ref[k].data[~vis] = before[k][0][~vis]
            st["exp_avg"][~vis] = before[k][1][~vis]
            st["exp_avg_sq"][~vis] = before[k][2][~vis]
        for k in ours:
            assert torch.equal(ours[k].detach()[~vis], mine[k][0][~vis]), (k, step)       # untouched bit for bit
            assert torch.equal(opt.state[ours[k]]["exp_avg"][~vis], mine[k][1][~vis]), (k, step)
            assert (ours[k].detach() - ref[k].detach()).abs().max().item() <= 1e-4 * lrs[k] + 2e-7 * ref[k].abs().max().item(), (k, step)
            assert (opt.state[ours[k]]["exp_avg"] - topt.state[ref[k]]["exp_avg"]).abs().max().item() <= 1e-6 * 5, (k, step)
            assert int(opt.state[ours[k]]["step"]) == step + 1
    # the densification code replaces state tensors and parameters in place (gaussian_model.py:316-392): keep working
    k = "opacity"
    st = opt.state.pop(ours[k])
    grown = torch.nn.Parameter(torch.cat([ours[k].detach(), torch.zeros(7, 1, device=cuda)]))
    st["exp_avg"] = torch.cat([st["exp_avg"], torch.zeros(7, 1, device=cuda)])
    st["exp_avg_sq"] = torch.cat([st["exp_avg_sq"], torch.zeros(7, 1, device=cuda)])
    opt.param_groups[3]["params"][0] = grown
    opt.state[grown] = st
    grown.grad = torch.ones_like(grown)
    for kk in ours:
        if kk != k:
            ours[kk].grad = None
    opt.step(torch.ones(N + 7, dtype=torch.bool, device=cuda), N + 7)
    assert int(opt.state[grown]["step"]) == 5 and bool((grown.detach()[-7:] < 0).all())


def test_unaligned_segments_take_the_scalar_path(cuda):
    import torch
    from gsr_b200 import capi
    rng = np.random.default_rng(4)
    n = 1001
    host = {k: rng.standard_normal(n + 1).astype(np.float32) for k in "pg"}
    big = {k: torch.from_numpy(host[k]).to(cuda) if k in host else torch.zeros(n + 1, device=cuda) for k in "pgmv"}
    t = {k: big[k][1:].view(n, 1) for k in big}             # 4-byte aligned only
    p, m, v = host["p"][1:].reshape(n, 1).copy(), np.zeros((n, 1), np.float32), np.zeros((n, 1), np.float32)
    oo.adam_step(p, host["g"][1:].reshape(n, 1), m, v, 1, 0.01)
    capi.adam_step([dict(param=t["p"], grad=t["g"], exp_avg=t["m"], exp_avg_sq=t["v"], lr=0.01)], 1)
    assert np.abs(t["p"].cpu().numpy() - p).max() <= 1e-6


def test_densification_stats_vs_oracle_and_reference_expressions(cuda):
    import torch
    from gsr_b200 import capi
    rng = np.random.default_rng(3)
    P = 100_003
    g2 = rng.standard_normal((P, 3)).astype(np.float32)
    radii = rng.integers(0, 40, P).astype(np.int32)
    radii[rng.random(P) < 0.5] = 0
    accum = rng.random((P, 1)).astype(np.float32)
    denom = rng.integers(0, 5, (P, 1)).astype(np.float32)
    maxr = rng.integers(0, 30, P).astype(np.float32)
    t = [torch.from_numpy(a.copy()).to(cuda) for a in (g2, radii, accum, denom, maxr)]
    capi.densification_stats(*t)
    oo.densification_stats(g2, radii, accum, denom, maxr)
    assert np.allclose(t[2].cpu().numpy(), accum, rtol=1e-6, atol=0)
    assert np.array_equal(t[3].cpu().numpy(), denom) and np.array_equal(t[4].cpu().numpy(), maxr)


def test_error_paths(cuda):
    import torch
    from gsr_b200 import capi
    z = torch.zeros(4, 3, device=cuda)
    with pytest.raises(RuntimeError, match="step counts from 1"):
        capi.adam_step([dict(param=z, grad=z, exp_avg=z, exp_avg_sq=z, lr=0.1)], 0)
