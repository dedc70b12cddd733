"""Seeded inputs of the loss-stage tests (shared by the CPU and GPU tests and the golden generator)."""
import numpy as np

# name -> (B, CH, H, W, seed, kind)
CASES = {
    "rand_1x3x64x64": (1, 3, 64, 64, 11, "rand"),
    "rand_2x5x37x75": (2, 5, 37, 75, 12, "rand"),       # ragged: neither side a multiple of the 32-pixel tile
    "tiny_1x1x7x5": (1, 1, 7, 5, 13, "rand"),           # smaller than the 11-tap window
    "smooth_1x3x96x160": (1, 3, 96, 160, 14, "smooth"),  # image-like: render vs. slightly different target
    "edge_1x3x33x33": (1, 3, 33, 33, 15, "rand"),        # one pixel past a tile boundary
}
GOLDEN = ["rand_1x3x64x64", "rand_2x5x37x75", "tiny_1x1x7x5", "smooth_1x3x96x160"]


def make(name):
    B, CH, H, W, seed, kind = CASES[name]
    rng = np.random.default_rng(seed)
    if kind == "rand":
        a = rng.random((B, CH, H, W), dtype=np.float32)
        b = rng.random((B, CH, H, W), dtype=np.float32)
    else:
        yy, xx = np.meshgrid(np.linspace(0, 1, H, dtype=np.float32), np.linspace(0, 1, W, dtype=np.float32), indexing="ij")
        base = np.stack([0.5 + 0.4 * np.sin(6.0 * xx + c) * np.cos(5.0 * yy - c) for c in range(CH)]).astype(np.float32)
        a = np.clip(base[None] + 0.05 * rng.standard_normal((B, CH, H, W)).astype(np.float32), 0, 1)
        b = np.clip(base[None] + 0.02 * rng.standard_normal((B, CH, H, W)).astype(np.float32), 0, 1)
    dl = rng.standard_normal((B, CH, H, W)).astype(np.float32)
    return np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32), dl
