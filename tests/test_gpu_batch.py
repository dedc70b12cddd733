"""gsr_forward_batch / gsr_backward_batch (include/gsr.h): the views of a step in one call.

The per-Gaussian stages run once for all cameras (csrc/preprocess_batch.cu); every per-view output must be what the
single-view entry points produce for that camera -- bit for bit for every geometric and integer artefact of the forward
(same project_gaussian code; the SH polynomial is ordinary fp32 code and agrees to 2e-6) -- and the parameter gradients must be the sum over the views of the single-view backward's (up to the order of the fp32
adds).  The single-view path itself is pinned to the oracle and the reference CUDA build in tests/test_gpu_parity.py, and
one view of the batch is compared with the oracle here as well."""
import numpy as np
import pytest
import torch

import cases
from gsr_b200 import capi, synth

pytestmark = pytest.mark.gpu

SHARED = ("means3D", "opacities", "background", "shs", "scales", "rotations", "sh_degree", "antialiasing")


def _setup(P=30_000, V=5, W=167, H=101, seed=21, aa=False, degree=3):
    scene = synth.trained_like_scene(P, seed, s0=0.03)
    scene["sh_degree"] = degree
    cams = synth.orbit_cameras(8, W, H, seed)[:V]
    opts = dict(bg=(0.1, 0.2, 0.3), antialiasing=aa, mode="sh")
    kws = [cases._to_dev(cases.inputs_of(scene, c, opts)) for c in cams]
    shared = {k: kws[0][k] for k in SHARED}
    cam_kw = [{k: kw[k] for k in ("viewmatrix", "projmatrix", "cam_pos", "tan_fovx", "tan_fovy")} for kw in kws]
    return scene, cams, opts, kws, shared, cam_kw, W, H


@pytest.mark.parametrize("aa,degree", [(False, 3), (True, 2)])
def test_forward_batch_is_bitwise_the_single_view_forward(cuda, aa, degree):
    scene, cams, opts, kws, shared, cam_kw, W, H = _setup(aa=aa, degree=degree)
    single = [capi.forward(width=W, height=H, **kw) for kw in kws]
    batch = capi.forward_batch(cam_kw, width=W, height=H, **shared)
    assert len(batch) == len(single)
    for a, b in zip(single, batch):
        assert a["num_rendered"] == b["num_rendered"] and a["num_rendered"] > 0
        for k in ("out_invdepth", "radii"):
            assert torch.equal(a[k], b[k]), k
        ia, ib = capi.introspect(a), capi.introspect(b)
        vis = a["radii"] > 0
        for k in ("tiles_touched", "point_offsets", "keys", "point_list", "ranges", "n_contrib", "final_T"):
            assert torch.equal(ia[k], ib[k]), k
        for k in ("depths", "means2D", "conic_opacity", "cov3D", "clamped"):              # defined for visible Gaussians
            assert torch.equal(ia[k][vis], ib[k][vis]), k
        # the SH polynomial is ordinary fp32 code (tolerance 1e-5 in the contract, not a bit-exact artefact): the
        # compiler may contract it differently in the two kernels
        assert (ia["rgb"][vis] - ib["rgb"][vis]).abs().max().item() <= 2e-6
        assert (a["out_color"] - b["out_color"]).abs().max().item() <= 2e-6
    # the views' chains on their own streams: same results
    side = [torch.cuda.Stream(), torch.cuda.Stream()]
    fanned = capi.forward_batch(cam_kw, width=W, height=H, streams=[torch.cuda.current_stream()] + side, **shared)
    torch.cuda.synchronize()
    for b, c in zip(batch, fanned):
        assert b["num_rendered"] == c["num_rendered"]
        assert torch.equal(b["out_color"], c["out_color"]) and torch.equal(b["radii"], c["radii"])
        assert torch.equal(capi.introspect(b)["n_contrib"], capi.introspect(c)["n_contrib"])
    # and one view against the oracle (radii, image)
    ref = cases.run_oracle(scene, cams[2], opts)
    assert np.array_equal(batch[2]["radii"].cpu().numpy(), ref["radii"])
    ok = ~ref["borderline"]
    assert np.abs(batch[2]["out_color"].cpu().numpy() - ref["out_color"])[:, ok].max() <= 1e-5


@pytest.mark.parametrize("aa", [False, True])
def test_backward_batch_sums_the_single_view_backwards(cuda, aa):
    scene, cams, opts, kws, shared, cam_kw, W, H = _setup(aa=aa)
    P = scene["means3D"].shape[0]
    rng = np.random.default_rng(5)
    dcol = [torch.from_numpy(rng.standard_normal((3, H, W)).astype(np.float32)).cuda() for _ in cams]
    ddep = [torch.from_numpy(rng.standard_normal((1, H, W)).astype(np.float32)).cuda() for _ in cams]
    states = capi.forward_batch(cam_kw, width=W, height=H, **shared)
    keys = ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations")
    want = {k: 0 for k in keys}
    per_view_want = []
    for v, (st, kw) in enumerate(zip(states, kws)):
        g = capi.backward(st, dL_dcolor=dcol[v], dL_dinvdepth=ddep[v], **kw)
        for k in keys:
            want[k] = want[k] + g[k].double()
        per_view_want.append({k: g[k].clone() for k in ("dL_dmeans2D", "dL_dconic", "dL_dcolors", "dL_dinvdepths")})
    got, per_view = capi.backward_batch(states, cam_kw, dL_dcolor=dcol, dL_dinvdepth=ddep,
                                        streams=[torch.cuda.current_stream(), torch.cuda.Stream()] if aa else None, **shared)
    tol = dict(dL_dscales=1e-3, dL_drotations=1e-3)        # the two cancellation-heavy tensors
    for k in keys:
        w = want[k].cpu().numpy()
        err = np.abs(got[k].double().cpu().numpy().reshape(w.shape) - w).max() / (np.abs(w).max() + 1e-30)
        assert err <= tol.get(k, 1e-5), (k, err)
    for v in range(len(cams)):
        for k, wv in per_view_want[v].items():
            w = wv.double().cpu().numpy()
            err = np.abs(per_view[v][k].double().cpu().numpy().reshape(w.shape) - w).max() / (np.abs(w).max() + 1e-30)
            assert err <= 1e-5, (v, k, err)
    # accumulate mode adds the same sum into what the outputs hold, visible rows only
    bucket = {k: torch.full_like(got[k], 3.0) for k in keys}
    capi.backward_batch(states, cam_kw, dL_dcolor=dcol, dL_dinvdepth=ddep, out=bucket, accumulate=True, **shared)
    seen = torch.zeros(P, dtype=torch.bool, device=cuda)
    for st in states:
        seen |= st["radii"] > 0
    assert 0 < int(seen.sum()) < P
    for k in keys:
        delta = (bucket[k].double() - 3.0).cpu().numpy()
        w = want[k].cpu().numpy().reshape(delta.shape)
        assert np.abs(delta - w).max() <= max(tol.get(k, 1e-5) * np.abs(w).max(), 1e-5), k
        assert torch.equal(bucket[k][~seen], torch.full_like(bucket[k][~seen], 3.0)), k


def test_batch_argument_checks(cuda):
    scene, cams, opts, kws, shared, cam_kw, W, H = _setup(P=2000, V=2)
    with pytest.raises(RuntimeError, match="n_views"):
        capi.forward_batch(cam_kw * 5, width=W, height=H, **shared)                      # 10 views > GSR_MAX_BATCH_VIEWS
    bad = dict(shared)
    bad["shs"] = shared["shs"][:, :4].contiguous()                                       # M = 4: single-view path only
    bad["sh_degree"] = 1
    with pytest.raises(RuntimeError, match="M = 16"):
        capi.forward_batch(cam_kw, width=W, height=H, **bad)
    # per-view image sizes may differ
    cam2 = [dict(cam_kw[0]), dict(cam_kw[1], width=96, height=64)]
    st = capi.forward_batch(cam2, width=W, height=H, **shared)
    assert st[0]["out_color"].shape == (3, H, W) and st[1]["out_color"].shape == (3, 64, 96)
