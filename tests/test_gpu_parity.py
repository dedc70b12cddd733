"""GPU parity tests proper: this repo's sm_100a kernels, called through the C ABI
(gsr_b200.capi -> libgsr_b200.so), against the CPU oracle, the committed golden vectors and -- when
oracle/_ref was built -- the unmodified reference CUDA rasterizer on the same GPU.

Bars: bit-exact integer artefacts (+ depth / pixel-centre / conic bits); colour, final_T, inverse
depth <= 1e-5 max-abs; gradients <= 1e-4 relative to the tensor's max-abs (1e-3 for the
conic->covariance chain, where the reference's own fp32 result is 3e-4 from the fp64 oracle)."""
import os

import numpy as np
import pytest

import cases
from test_oracle_golden import GRAD_TOL, ORACLE_GRAD_TOL, GOLD

pytestmark = pytest.mark.gpu

# worst measured gradient error per (comparison, case, tensor), written by test_zz_worst_errors_report
WORST = {}


def _record(kind, name, rep):
    WORST.setdefault(kind, {})[name] = {k: rep[k] for k in rep if k.endswith("_relmax") or k.endswith("_viol") or k.endswith("_absmax")}


def _assert_parity(rep, against_ref, name, tol=None):
    tol = GRAD_TOL if tol is None else tol
    assert rep["num_rendered"][0] == rep["num_rendered"][1]
    for k in cases.INT_KEYS:
        assert rep[k] == 0, (name, k, rep[k])
    for k in cases.BITEXACT_FLOAT_KEYS:
        assert rep[k + "_bits"] == 0, (name, k, rep[k + "_bits"])
    assert rep["clamped"] == 0
    for k in ("rgb", "out_color", "final_T", "invdepth"):
        assert rep[k + "_maxabs"] <= 1e-5, (name, k, rep[k + "_maxabs"])
    assert rep["cov3D_maxabs"] == 0.0
    if against_ref:   # same expf on both sides: images must agree to the last bit
        assert rep["out_color_maxabs"] <= 1e-6 and rep["final_T_maxabs"] == 0.0 and rep["invdepth_maxabs"] <= 1e-6
    for k in cases.BWD_KEYS:
        if k + "_relmax" in rep:
            assert rep[k + "_relmax"] <= tol[k], (name, k, rep[k + "_relmax"])
        # elementwise bar: |dx| <= tol |y| + 0.1 tol max|y| for every entry, not just the worst relative to the max
        if k + "_viol" in rep and k + "_relmax" in rep:
            assert rep[k + "_viol"] <= VIOL_FRAC * rep[k + "_size"], (name, k, rep[k + "_viol"], rep[k + "_size"])


# Fraction of a gradient tensor's entries allowed outside the elementwise bound.  Measured on B200 (round 2,
# gpurun_out/parity_worst.json): 0 entries in every small / medium case; at full size 3 of 12 M (c2 dL_drotations) and
# at most 6 of 36 M (c4 dL_dcov3D) against the reference CUDA build, whose own fp32 atomics move as many entries
# from run to run (tools/c4_grad_noise.py).  2e-6 allows none below 500 k entries.
VIOL_FRAC = 2e-6


@pytest.mark.parametrize("name", cases.SMALL_CASES + cases.ARG_CASES + cases.MEDIUM_CASES)
def test_cuda_vs_oracle(cuda, name):
    scene, cam, opts = cases.get_case(name)
    grads = cases.grads_for(cam)
    o = cases.run_oracle(scene, cam, opts, grads)
    c = cases.run_cuda(scene, cam, opts, grads)
    rep = cases.compare(c, o, o["borderline"], GRAD_TOL)     # the elementwise bar keeps its table; the max-norm bar is 1e-4
    _record("cuda_vs_oracle", name, rep)
    _assert_parity(rep, False, name, ORACLE_GRAD_TOL)
    assert rep.get("n_contrib_borderline_mismatch", 0) <= max(3, int(o["borderline"].sum()))
    # deep_list: every pixel ends on the T < 1e-4 termination test after ~2100 layers, a few land on the threshold
    assert o["borderline"].mean() < (5e-3 if name == "deep_list" else 1e-3)
    if name == "deep_list":
        assert int(o["n_contrib"].min()) >= 2000, "the case must keep >= 2000 layers per pixel"


@pytest.mark.parametrize("name", cases.SMALL_CASES + cases.ARG_CASES + cases.MEDIUM_CASES)
def test_cuda_vs_reference_cuda(cuda, name):
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/libref_rasterizer.so not built (needs /root/reference at build time)")
    scene, cam, opts = cases.get_case(name)
    grads = cases.grads_for(cam)
    r = cases.run_ref(scene, cam, opts, grads)
    c = cases.run_cuda(scene, cam, opts, grads)
    rep = cases.compare(c, r, None, GRAD_TOL)
    # gradients: both sides are fp32 with different summation orders; judge each against the oracle's scale
    for k in cases.BWD_KEYS:
        if k + "_relmax" in rep and np.abs(r[k]).max() < 1e-6 * max(1.0, float(np.abs(r["dL_dmeans3D"]).max())):
            rep.pop(k + "_relmax")      # analytically-zero gradient (identity rotations): noise vs noise
    _record("cuda_vs_reference_cuda", name, rep)
    _assert_parity(rep, True, name)


@pytest.mark.parametrize("name", cases.GOLDEN_CASES)
def test_cuda_vs_golden(cuda, name):
    gold = np.load(os.path.join(GOLD, name + ".npz"))
    scene, cam, opts = cases.get_case(name)
    c = cases.run_cuda(scene, cam, opts, cases.grads_for(cam))
    rows = np.arange(0, scene["means3D"].shape[0], int(gold["stride"]))
    for k in ("radii", "tiles_touched", "point_offsets", "keys", "point_list", "ranges", "n_contrib"):
        assert np.array_equal(c[k], gold[k]), k
    for k in ("depths", "means2D", "conic_opacity"):
        assert np.array_equal(c[k][rows].view(np.uint32), gold[k].view(np.uint32)), k
    for k in ("out_color", "final_T", "invdepth"):
        assert np.abs(c[k].astype(np.float64) - gold[k]).max() <= 1e-5, k
    for k in cases.BWD_KEYS:
        if k in gold.files and np.abs(gold[k]).max() > 1e-12:
            err = np.abs(c[k][rows].astype(np.float64) - gold[k]).max() / np.abs(gold[k]).max()
            assert err <= GRAD_TOL[k], (k, err)


def test_full_size_properties(cuda):
    """BASELINE config 2 at full size (3M Gaussians, 1920x1080): size-independent properties."""
    import torch
    from gsr_b200 import capi
    import bench
    w, scene, cam = bench.make_workload("c2")
    kw = cases._to_dev(cases.inputs_of(scene, cam, dict(bg=(0, 0, 0), antialiasing=False, mode="sh")))
    st = capi.forward(width=w["W"], height=w["H"], **kw)
    it = capi.introspect(st)
    N = st["num_rendered"]
    tiles = it["tiles_touched"].to(torch.int64)
    assert N == int(tiles.sum()) and N > 1_000_000
    assert torch.equal(it["point_offsets"].to(torch.int64) & 0xFFFFFFFF, torch.cumsum(tiles, 0) & 0xFFFFFFFF)
    keys = it["keys"]
    assert bool((keys[1:] >= keys[:-1]).all()), "sortedness"
    same = keys[1:] == keys[:-1]
    pl = it["point_list"].to(torch.int64)
    assert bool((pl[1:][same] > pl[:-1][same]).all()), "stability"
    depth_bits_all = it["depths"].view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    depth_bits = depth_bits_all[pl]
    assert torch.equal(keys & 0xFFFFFFFF, depth_bits), "key low word = depth bits of its Gaussian"
    # checksum: the multiset of (tile, gaussian) instances is what the rects say
    T = ((w["W"] + 15) // 16) * ((w["H"] + 15) // 16)
    # the binning split: visible Gaussians in (depth bits, index) order, instances emitted in that order
    vis_idx = torch.nonzero(tiles > 0).flatten()
    order = it["depth_order"].to(torch.int64)
    assert order.numel() == vis_idx.numel() and torch.equal(torch.sort(order).values, vis_idx), "depth_order is a permutation of the visible set"
    dkey = (depth_bits_all[order] << 32) | order
    assert bool((dkey[1:] > dkey[:-1]).all()), "depth_order sorted by (depth bits, Gaussian index)"
    assert torch.equal(it["sorted_offsets"].to(torch.int64), torch.cumsum(tiles[order], 0)), "sorted_offsets"
    assert torch.equal(it["tile_keys"].to(torch.int64), keys >> 32)
    cnt = torch.bincount(keys >> 32, minlength=T)
    r = it["ranges"].to(torch.int64)
    assert torch.equal(r[:, 1] - r[:, 0], cnt)
    assert torch.equal(torch.bincount(pl, minlength=w["P"]), tiles)
    nc = it["n_contrib"].to(torch.int64)
    assert int(nc.max()) <= int(cnt.max())
    fT = it["final_T"]
    assert float(fT.max()) <= 1.0 and float(fT.min()) >= 0.0
    assert bool(torch.isfinite(st["out_color"]).all())
    # idempotence: the same call again gives bit-identical forward outputs
    st2 = capi.forward(width=w["W"], height=w["H"], **kw)
    assert torch.equal(st2["out_color"], st["out_color"]) and torch.equal(st2["radii"], st["radii"])
    assert torch.equal(capi.introspect(st2)["point_list"], it["point_list"])
    # linearity of the backward in the upstream gradient: bwd(2g) == 2 bwd(g) up to atomics noise
    dcol = torch.rand(3, w["H"], w["W"], device="cuda") - 0.5
    g1 = capi.backward(st, dL_dcolor=dcol, **kw)
    g2 = capi.backward(st, dL_dcolor=2 * dcol, **kw)
    for k in ("dL_dmeans3D", "dL_dsh", "dL_dopacity"):
        a, b = g1[k].double(), g2[k].double()
        assert float((2 * a - b).abs().max() / b.abs().max()) < 1e-4, k
    vis = st["radii"] > 0
    assert float(g1["dL_dsh"][~vis].abs().max()) == 0.0 and float(g1["dL_dmeans3D"][~vis].abs().max()) == 0.0


@pytest.mark.parametrize("workload", ["c2", "c3", "c4"])
def test_full_size_vs_reference_cuda(cuda, workload):
    """BASELINE configs at FULL size (c2: 3M Gaussians 1920x1080; c3: 1.8M 979x546; c4: 6M 3840x2160 with
    antialiasing + inverse depth) against the unmodified reference CUDA rasterizer on this GPU: every integer
    artefact, depth / pixel-centre / conic bits and n_contrib bit-exact, images to the last bits, gradients within
    north_star's 1e-4."""
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/libref_rasterizer.so not built (needs /root/reference at build time)")
    import bench
    w, scene, cam = bench.make_workload(workload)
    opts = dict(bg=(0.0, 0.0, 0.0), antialiasing=w["aa"], mode="sh")
    grads = cases.grads_for(cam)
    r = cases.run_ref(scene, cam, opts, grads)
    c = cases.run_cuda(scene, cam, opts, grads)
    rep = cases.compare(c, r, None, GRAD_TOL)
    _record("full_size_vs_reference_cuda", workload, rep)
    assert rep["num_rendered"][0] > 1_000_000
    # north_star: gradients <= 1e-4 "given atomic-order nondeterminism".  At 6M Gaussians a handful of ill-conditioned
    # rows (dL_dconic ~ 1e6 cancelling in the conic -> covariance chain) make the REFERENCE's fp32 atomics differ from
    # run to run by 1e-3 and from the fp64-accumulating oracle by 1.6e-3 (tools/c4_grad_noise.py).
    # A gradient that misses the tolerance against the reference is therefore arbitrated by the oracle: this path must
    # be within the tolerance of the oracle, or at least as close to it as the reference is.
    failing = [k for k in cases.BWD_KEYS if rep.get(k + "_relmax", 0.0) > GRAD_TOL[k]]
    if failing:
        o = cases.run_oracle(scene, cam, opts, grads)
        for k in failing:
            tr = np.asarray(o[k], np.float64)
            s_ = np.abs(tr).max()
            e_ours = np.abs(np.asarray(c[k], np.float64).reshape(tr.shape) - tr).max() / s_
            e_ref = np.abs(np.asarray(r[k], np.float64).reshape(tr.shape) - tr).max() / s_
            assert e_ours <= max(GRAD_TOL[k], e_ref), (workload, k, e_ours, e_ref)
            WORST["full_size_vs_reference_cuda"][workload][k + "_arbitrated"] = {"ours_vs_oracle": e_ours, "reference_vs_oracle": e_ref}
            rep.pop(k + "_relmax")
    _assert_parity(rep, True, workload)


def test_c5_10m_forward_vs_reference_cuda(cuda):
    """BASELINE configs[4] at its largest size: 10M Gaussians, 1920x1080, forward only, against the unmodified reference
    CUDA rasterizer: every integer artefact bit-exact, the image to the last bits."""
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/libref_rasterizer.so not built (needs /root/reference at build time)")
    import bench
    w, scene, cam = bench.make_workload("c5")
    assert scene["means3D"].shape[0] == 10_000_000
    opts = dict(bg=(0.0, 0.0, 0.0), antialiasing=False, mode="sh")
    r = cases.run_ref(scene, cam, opts, None)
    c = cases.run_cuda(scene, cam, opts, None)
    rep = cases.compare(c, r, None)
    assert rep["num_rendered"][0] > 5_000_000
    _assert_parity(rep, True, "c5")


def test_zz_worst_errors_report():
    """Writes the worst measured gradient errors of this session (no assertion of its own): the 1e-3 rows of GRAD_TOL
    carry a measured number in the log and in gpurun_out/parity_worst.json."""
    import json
    print(json.dumps(WORST, indent=1, default=str))
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out_dir) and WORST:
        with open(os.path.join(out_dir, "parity_worst.json"), "w") as f:
            json.dump(WORST, f, indent=1, default=str)
