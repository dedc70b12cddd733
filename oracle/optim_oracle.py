"""CPU oracle of the optimizer step and the densification statistics (numpy, fp32).

TEST INFRASTRUCTURE ONLY: imported by tests/ and bench.py's cpu_baseline leg; the product never imports it.

Restates
  * torch.optim.Adam as the reference configures it (/root/reference/scene/gaussian_model.py:178-201:
    Adam(l, lr=0.0, eps=1e-15), six groups with their own lr, default betas (0.9, 0.999), weight_decay 0, amsgrad
    off) and steps it (/root/reference/train.py:176-186).  torch is a third-party dependency of the reference
    (environment.yml pins pytorch 1.12.1; this image has torch 2.11).  The algorithm follows torch/optim/adam.py
    `_single_tensor_adam`:  step += 1;  exp_avg.lerp_(grad, 1-beta1);  exp_avg_sq.mul_(beta2).addcmul_(grad, grad,
    value=1-beta2);  bias corrections in Python floats;  denom = exp_avg_sq.sqrt() / sqrt(bias_correction2) + eps;
    param.addcdiv_(exp_avg, denom, value=-lr/bias_correction1).
    Pinned by tests/test_optim_oracle_cpu.py against torch.optim.Adam itself (the dependency is importable here).
  * the densification statistics of /root/reference/train.py:166-167 and scene/gaussian_model.py:471-473.
"""
import math

import numpy as np

F = np.float32


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-15, visible=None):
    """One Adam step on fp32 arrays [rows, cols] (in place).  lr: scalar or per-column array.  visible: bool[rows]
    or None; rows that are not visible keep parameters and moments (the sparse variant of train.py:179-181)."""
    p = param.reshape(param.shape[0], -1)
    g = grad.reshape(p.shape)
    m = exp_avg.reshape(p.shape)
    v = exp_avg_sq.reshape(p.shape)
    rows = slice(None) if visible is None else np.asarray(visible, bool)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = (np.broadcast_to(np.asarray(lr, np.float64), (p.shape[1],)) / bc1).astype(F)
    bc2_sqrt = F(math.sqrt(bc2))
    gg, mm, vv, pp = g[rows], m[rows], v[rows], p[rows]
    mm = (mm + F(1.0 - beta1) * (gg - mm)).astype(F)
    vv = (vv * F(beta2) + (F(1.0 - beta2) * gg) * gg).astype(F)
    denom = (np.sqrt(vv) / bc2_sqrt + F(eps)).astype(F)
    pp = (pp - step_size[None, :] * (mm / denom)).astype(F)
    m[rows], v[rows], p[rows] = mm, vv, pp


def densification_stats(grad_means2D, radii, xyz_gradient_accum, denom, max_radii2D):
    """train.py:166-167 + gaussian_model.py:471-473 (in place)."""
    vis = radii > 0
    g = grad_means2D[vis, :2].astype(F)
    xyz_gradient_accum.reshape(-1)[vis] += np.sqrt(g[:, 0] * g[:, 0] + g[:, 1] * g[:, 1]).astype(F)
    denom.reshape(-1)[vis] += F(1)
    max_radii2D[vis] = np.maximum(max_radii2D[vis], radii[vis].astype(F))
