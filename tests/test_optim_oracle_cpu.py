"""CPU: the optimizer oracle (oracle/optim_oracle.py) against the reference's own optimizer -- torch.optim.Adam
configured as /root/reference/scene/gaussian_model.py:178-201 does -- and the densification statistics against
the torch expressions of /root/reference/train.py:166-167, scene/gaussian_model.py:471-473."""
import numpy as np
import torch

from oracle import optim_oracle as oo

# the reference's learning rates (arguments/__init__.py OptimizationParams): position 1.6e-4 * extent, feature 2.5e-3
# (f_rest: /20), opacity 0.025, scaling 0.005, rotation 0.001
GROUPS = [("xyz", 3, 1.6e-4 * 5.0), ("f_dc", 3, 2.5e-3), ("f_rest", 45, 2.5e-3 / 20.0), ("opacity", 1, 0.025),
          ("scaling", 3, 0.005), ("rotation", 4, 0.001)]


def test_adam_matches_torch_optim_adam_over_steps():
    rng = np.random.default_rng(0)
    P = 257
    params = {n: rng.standard_normal((P, c)).astype(np.float32) for n, c, _ in GROUPS}
    tp = {n: torch.nn.Parameter(torch.from_numpy(params[n].copy())) for n, _, _ in GROUPS}
    opt = torch.optim.Adam([{"params": [tp[n]], "lr": lr, "name": n} for n, _, lr in GROUPS], lr=0.0, eps=1e-15)
    m = {n: np.zeros_like(params[n]) for n, _, _ in GROUPS}
    v = {n: np.zeros_like(params[n]) for n, _, _ in GROUPS}
    for step in range(1, 8):
        grads = {n: (rng.standard_normal((P, c)) * 10.0 ** rng.integers(-6, 1)).astype(np.float32) for n, c, _ in GROUPS}
        if step == 3:
            grads["xyz"][::2] = 0.0          # rows the view did not see still move by momentum (dense Adam)
        for n, _, lr in GROUPS:
            tp[n].grad = torch.from_numpy(grads[n].copy())
            oo.adam_step(params[n], grads[n], m[n], v[n], step, lr)
        opt.step()
        for n, _, lr in GROUPS:
            # tolerances: fp32 rounding of each formula, relative to the scale of its inputs (m and p can cancel)
            gmax = np.abs(grads[n]).max()
            want = tp[n].detach().numpy()
            st = opt.state[tp[n]]
            assert np.abs(m[n] - st["exp_avg"].numpy()).max() <= 1e-6 * max(gmax, np.abs(m[n]).max()), (n, step)
            assert np.abs(v[n] - st["exp_avg_sq"].numpy()).max() <= 1e-6 * max(gmax ** 2, np.abs(v[n]).max()), (n, step)
            # one step moves a parameter by at most ~lr: 1e-4 of that, plus fp32 rounding of the parameter itself
            assert np.abs(params[n] - want).max() <= 1e-4 * lr + 2e-7 * np.abs(want).max(), (n, step)


def test_adam_per_column_lr_equals_two_groups():
    # the SH row [P,48] = f_dc (3, feature_lr) + f_rest (45, feature_lr/20) in one tensor
    rng = np.random.default_rng(1)
    P = 64
    sh = rng.standard_normal((P, 48)).astype(np.float32)
    g = rng.standard_normal((P, 48)).astype(np.float32)
    a, b = sh[:, :3].copy(), sh[:, 3:].copy()
    lr = np.concatenate([np.full(3, 2.5e-3), np.full(45, 2.5e-3 / 20)])
    m, v = np.zeros_like(sh), np.zeros_like(sh)
    oo.adam_step(sh, g, m, v, 1, lr)
    oo.adam_step(a, g[:, :3].copy(), np.zeros_like(a), np.zeros_like(a), 1, 2.5e-3)
    oo.adam_step(b, g[:, 3:].copy(), np.zeros_like(b), np.zeros_like(b), 1, 2.5e-3 / 20)
    assert np.array_equal(sh[:, :3], a) and np.array_equal(sh[:, 3:], b)


def test_adam_visible_mask_leaves_other_rows_untouched():
    rng = np.random.default_rng(2)
    p = rng.standard_normal((50, 4)).astype(np.float32)
    g = rng.standard_normal((50, 4)).astype(np.float32)
    m, v = rng.random((50, 4)).astype(np.float32), rng.random((50, 4)).astype(np.float32)
    vis = rng.random(50) < 0.4
    p0, m0, v0 = p.copy(), m.copy(), v.copy()
    pd, md, vd = p.copy(), m.copy(), v.copy()
    oo.adam_step(p, g, m, v, 5, 0.01, visible=vis)
    oo.adam_step(pd, g, md, vd, 5, 0.01)
    assert np.array_equal(p[~vis], p0[~vis]) and np.array_equal(m[~vis], m0[~vis]) and np.array_equal(v[~vis], v0[~vis])
    assert np.array_equal(p[vis], pd[vis]) and np.array_equal(m[vis], md[vis])


def test_densification_stats_match_reference_expressions():
    rng = np.random.default_rng(3)
    P = 1000
    g2 = rng.standard_normal((P, 3)).astype(np.float32)
    radii = rng.integers(-1, 40, P).astype(np.int32)
    radii[rng.random(P) < 0.5] = 0
    accum = rng.random((P, 1)).astype(np.float32)
    denom = rng.integers(0, 5, (P, 1)).astype(np.float32)
    maxr = rng.integers(0, 30, P).astype(np.float32)
    t_acc, t_den, t_max = torch.from_numpy(accum.copy()), torch.from_numpy(denom.copy()), torch.from_numpy(maxr.copy())
    tg, tr = torch.from_numpy(g2), torch.from_numpy(radii)
    vf = tr > 0
    t_max[vf] = torch.max(t_max[vf], tr[vf])                                       # train.py:166
    t_acc[vf] += torch.norm(tg[vf, :2], dim=-1, keepdim=True)                      # gaussian_model.py:472
    t_den[vf] += 1                                                                 # gaussian_model.py:473
    oo.densification_stats(g2, radii, accum, denom, maxr)
    assert np.allclose(accum, t_acc.numpy(), rtol=1e-6, atol=0)
    assert np.array_equal(denom, t_den.numpy()) and np.array_equal(maxr, t_max.numpy())


def test_sparse_gaussian_adam_has_no_cpu_path():
    import pytest
    import torch
    from gsr_b200.optim import SparseGaussianAdam
    p = torch.nn.Parameter(torch.zeros(4, 3))
    opt = SparseGaussianAdam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    assert isinstance(opt, torch.optim.Adam)
    p.grad = torch.ones_like(p)
    with pytest.raises(RuntimeError, match="no CPU path"):
        opt.step(torch.ones(4, dtype=torch.bool), 4)
