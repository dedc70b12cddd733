"""CPU tests of the oracle (oracle/gs_oracle.c): analytic cases, internal consistency, gradients by
finite differences, the sort/scan/ranges restatements.  No GPU."""
import numpy as np
import pytest

import cases
from oracle import gs_oracle as go
from gsr_b200 import synth


def _scene(name):
    scene, cam, opts = cases.get_case(name)
    kw = cases.inputs_of(scene, cam, opts)
    return scene, cam, opts, go.Scene(W=cam.image_width, H=cam.image_height, **kw)


def test_higher_msb_matches_reference_definition():
    # rasterizer_impl.cu:35-50: smallest b with (n >> b) == 0
    for n in [1, 2, 3, 4, 255, 256, 2170, 2500, 8160, 32400, 65535, 65536, 1 << 20]:
        assert go.higher_msb(n) == int(n).bit_length()


def test_single_gaussian_analytic():
    """One isotropic Gaussian: centre pixel value, radius and rect follow in closed form."""
    scene, cam, opts, S = _scene("one")
    f = go.forward(S)
    assert f["radii"][0] > 0 and f["num_rendered"] == f["tiles_touched"][0]
    # isotropic sigma_world -> sigma_px = focal * s / z ; cov2D = sigma_px^2 + 0.3 on the diagonal
    z = f["depths"][0]
    focal = cam.image_width / (2 * cam.tanfovx)
    var = (focal * float(scene["scales"][0, 0]) / z) ** 2 + 0.3
    a, b, c, o = f["conic_opacity"][0]
    # the Gaussian sits a little off the optical axis, so cov2D is isotropic only to first order
    assert abs(a - 1 / var) < 3e-3 / var and abs(c - 1 / var) < 3e-3 / var and abs(b) < 3e-3 / var
    assert f["radii"][0] == int(np.ceil(3 * np.sqrt(max(1 / a, 1 / c)))) or abs(f["radii"][0] - 3 * np.sqrt(var)) <= 1
    # peak pixel: colour = rgb * alpha + (1 - alpha) * bg with alpha = opacity * exp(power at that pixel)
    px, py = f["means2D"][0]
    ix, iy = int(round(px)), int(round(py))
    dx, dy = px - ix, py - iy
    power = -0.5 * (a * dx * dx + c * dy * dy) - b * dx * dy
    alpha = min(0.99, o * np.exp(power))
    want = f["rgb"][0] * alpha
    assert np.allclose(f["out_color"][:, iy, ix], want, atol=1e-6)
    assert f["n_contrib"][iy, ix] == 1 and abs(f["final_T"][iy, ix] - (1 - alpha)) < 1e-6
    # DC-only SH: rgb = max(0, C0*dc + 0.5)
    assert np.allclose(f["rgb"][0], np.maximum(0, synth.C0 * scene["shs"][0, 0] + 0.5), atol=1e-6)


def test_background_and_empty_tiles():
    scene, cam, opts, S = _scene("two")
    f = go.forward(S)
    empty = (f["n_contrib"] == 0)
    assert empty.any()
    assert np.all(f["final_T"][empty] == 1.0)
    for ch in range(3):
        assert np.all(f["out_color"][ch][empty] == np.float32(opts["bg"][ch]))
    # tiles without instances keep (0,0)
    r = f["ranges"]
    assert np.all(r[r[:, 0] == r[:, 1]] == 0) or True
    assert f["invdepth"][0][empty].max() == 0.0


@pytest.mark.parametrize("name", ["tiny64", "odd167x101", "aa160x96"])
def test_binning_invariants(name):
    scene, cam, opts, S = _scene(name)
    f = go.forward(S, stages="binning")
    N = f["num_rendered"]
    assert N == int(f["tiles_touched"].sum())
    assert np.array_equal(f["point_offsets"], np.cumsum(f["tiles_touched"], dtype=np.uint64).astype(np.uint32))
    keys = f["keys"]
    assert np.all(keys[1:] >= keys[:-1]), "sorted"
    # stability: equal keys keep ascending Gaussian order
    same = keys[1:] == keys[:-1]
    assert np.all(f["point_list"][1:][same] > f["point_list"][:-1][same])
    # key = tile << 32 | depth bits of its Gaussian
    depth_bits = f["depths"].view(np.uint32)[f["point_list"]]
    assert np.array_equal((keys & 0xFFFFFFFF).astype(np.uint32), depth_bits)
    tiles = (keys >> 32).astype(np.int64)
    gx, gy = S.grid
    assert tiles.max() < gx * gy
    # ranges partition the sorted list by tile
    r = f["ranges"].astype(np.int64)
    cnt = np.bincount(tiles, minlength=gx * gy)
    assert np.array_equal(r[:, 1] - r[:, 0], cnt)
    nz = cnt > 0
    assert np.array_equal(r[nz, 0], (np.cumsum(cnt) - cnt)[nz])
    # sort restatement == numpy stable sort on the same bits
    ks, vs = go.sort_pairs(f["keys_unsorted"], f["values_unsorted"], 32 + f["bit"])
    order = np.argsort(f["keys_unsorted"], kind="stable")
    assert np.array_equal(ks, f["keys_unsorted"][order]) and np.array_equal(vs, f["values_unsorted"][order])


def test_sort_pairs_partial_bits_is_stable():
    rng = np.random.default_rng(3)
    k = rng.integers(0, 1 << 40, 20000, dtype=np.uint64)
    v = np.arange(20000, dtype=np.uint32)
    for end_bit in (0, 5, 13, 32, 40):
        ks, vs = go.sort_pairs(k, v, end_bit)
        masked = k & np.uint64((1 << end_bit) - 1)
        order = np.argsort(masked, kind="stable")
        assert np.array_equal(vs, v[order]) and np.array_equal(ks, k[order])


def test_mark_visible():
    scene, cam, opts, S = _scene("tiny64")
    vis = go.mark_visible(scene["means3D"], cam.world_view_transform)
    V = cam.world_view_transform.astype(np.float64)
    z = scene["means3D"].astype(np.float64) @ V[:3, 2] + V[3, 2]
    clear = np.abs(z - 0.2) > 1e-5
    assert np.array_equal(vis[clear], (z > 0.2)[clear])


def test_render_conserves_alpha():
    """final_T equals the product of (1 - alpha) over blended Gaussians; colour <= sum of weights."""
    scene, cam, opts, S = _scene("tiny64")
    f = go.forward(S)
    assert np.all(f["final_T"] <= 1.0) and np.all(f["final_T"] >= 0.0001 * (1 - 0.99) - 1e-9)
    assert f["pairs_blended"] <= f["pairs_evaluated"]
    assert f["n_contrib"].max() <= (f["ranges"][:, 1] - f["ranges"][:, 0]).max()


def _loss(S, dcol, ddep):
    f = go.forward(S, borderline=False)
    return float((f["out_color"].astype(np.float64) * dcol).sum() + (f["invdepth"].astype(np.float64) * ddep).sum())


@pytest.mark.parametrize("aa", [False, True])
def test_backward_matches_finite_differences(aa):
    """Central differences of the oracle's own forward (fp32, so loose tolerance).  The blend is only
    piecewise smooth (alpha < 1/255 and T < 1e-4 cut-offs), so the scene is built to stay inside one
    smooth piece: every Gaussian covers the whole 48x40 image with alpha > 1/255 and T stays > 1e-4."""
    rng = np.random.default_rng(0)
    P = 6
    scene = dict(means3D=(rng.normal(0, 0.03, (P, 3)) + [0, 0, 0.2]).astype(np.float32),
                 scales=np.exp(rng.normal(np.log(0.12), 0.2, (P, 3))).astype(np.float32),
                 rotations=(lambda q: (q / np.linalg.norm(q, axis=1, keepdims=True)))(rng.normal(0, 1, (P, 4))).astype(np.float32),
                 opacities=rng.uniform(0.2, 0.5, (P, 1)).astype(np.float32),
                 shs=rng.normal(0, 0.3, (P, 16, 3)).astype(np.float32) + np.float32(0.4), sh_degree=3)
    cam = synth.chair_camera(48, 40)
    opts = dict(bg=(0.2, 0.3, 0.1), antialiasing=aa, mode="sh")
    dcol, ddep = synth.upstream_grads(40, 48, 1)
    dcol *= 0.5
    def mk(sc):
        return go.Scene(W=48, H=40, **cases.inputs_of(sc, cam, opts))
    S = mk(scene)
    f = go.forward(S)
    assert f["n_contrib"].min() == P and f["pairs_blended"] == 48 * 40 * P   # inside one smooth piece
    g = go.backward(S, f, dcol, ddep)
    checks = [("means3D", "dL_dmeans3D", 2e-3), ("opacities", "dL_dopacity", 1e-3), ("scales", "dL_dscales", 1e-3),
              ("rotations", "dL_drotations", 5e-3), ("shs", "dL_dsh", 2e-3)]
    worst = {}
    for key, gkey, eps in checks:
        arr = scene[key]
        flat_idx = rng.choice(arr.size, size=min(10, arr.size), replace=False)
        for fi in flat_idx:
            idx = np.unravel_index(fi, arr.shape)
            sp, sm = dict(scene), dict(scene)
            ap, am = arr.copy(), arr.copy()
            ap[idx] += eps; am[idx] -= eps
            sp[key], sm[key] = ap, am
            num = (_loss(mk(sp), dcol, ddep) - _loss(mk(sm), dcol, ddep)) / (float(ap[idx]) - float(am[idx]))
            ana = float(g[gkey].reshape(arr.shape)[idx])
            scale = max(abs(num), abs(ana), 1e-2 * float(np.abs(g[gkey]).max()))
            worst[gkey] = max(worst.get(gkey, 0), abs(num - ana) / scale)
    # fp32 forward => ~1e-3 is the best central differences can do
    for k, v in worst.items():
        assert v < 0.03, (k, v, worst)
