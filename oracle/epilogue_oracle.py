"""CPU oracle of the render epilogue (TEST INFRASTRUCTURE ONLY: tests/, smoke(), bench cpu legs).

Restates /root/reference/gaussian_renderer/__init__.py:112-119 in numpy:
    y[c] = sum_k x[k] * E[k][c] + E[c][3]   when an exposure [3,4] is given        (:113-115)
    y = clamp(y, 0, 1)                                                               (:119)
and the gradients torch autograd derives for those lines (clamp passes the gradient where 0 <= y <= 1).
Pinned against the reference's own torch expression + autograd in tests/test_epilogue_cpu.py."""
import numpy as np


def forward(image, exposure=None):
    x = np.asarray(image, np.float32)
    if exposure is None:
        return np.clip(x, 0.0, 1.0), x
    E = np.asarray(exposure, np.float32)
    y = np.empty_like(x)
    for c in range(3):
        y[c] = (x[0] * E[0, c] + x[1] * E[1, c] + x[2] * E[2, c]) + E[c, 3]
    return np.clip(y, 0.0, 1.0), y


def backward(image, exposure, dL_dout):
    """-> (dL_dimage [3,H,W] fp32, dL_dexposure [3,4] fp64-accumulated then fp32, or None)."""
    x = np.asarray(image, np.float32)
    _, y = forward(x, exposure)
    g = np.where((y >= 0.0) & (y <= 1.0), np.asarray(dL_dout, np.float32), np.float32(0.0)).astype(np.float32)
    if exposure is None:
        return g, None
    E = np.asarray(exposure, np.float32)
    dx = np.empty_like(x)
    for k in range(3):
        dx[k] = g[0] * E[k, 0] + g[1] * E[k, 1] + g[2] * E[k, 2]
    dE = np.zeros((3, 4), np.float64)
    for k in range(3):
        for c in range(3):
            dE[k, c] = np.sum(x[k].astype(np.float64) * g[c])
    for c in range(3):
        dE[c, 3] = np.sum(g[c].astype(np.float64))
    return dx, dE.astype(np.float32)
