"""The hand-written onesweep radix sort alone (gsr_sort_pairs) against torch's stable sort."""
import numpy as np
import pytest

import cases  # noqa: F401

pytestmark = pytest.mark.gpu


def _check(n, end_bit, seed, key_bits=64, runs=False):
    import torch
    from gsr_b200 import capi
    g = torch.Generator(device="cuda").manual_seed(seed)
    if n == 0:
        keys = torch.empty(0, dtype=torch.int64, device="cuda")
    elif runs:   # the rasterizer's shape: runs of equal depth bits, few distinct tiles
        depth = torch.randint(0, 1 << 31, (n // 6 + 1,), generator=g, device="cuda").repeat_interleave(6)[:n]
        tile = torch.randint(0, 8160, (n,), generator=g, device="cuda")
        keys = (tile << 32) | depth
    else:
        keys = torch.randint(0, 1 << 62, (n,), generator=g, device="cuda")
        if key_bits < 62:
            keys = keys & ((1 << key_bits) - 1)
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    masked = keys & ((1 << end_bit) - 1) if end_bit < 63 else keys
    order = torch.sort(masked, stable=True).indices
    want_k, want_v = keys[order], vals[order]
    ko, vo = capi.sort_pairs(keys.clone(), vals.clone(), end_bit)
    torch.cuda.synchronize()
    assert torch.equal(vo, want_v), (n, end_bit)
    assert torch.equal(ko, want_k), (n, end_bit)


@pytest.mark.parametrize("n", [0, 1, 2, 31, 32, 33, 255, 4095, 4096, 4097, 8191, 100_000, 1_000_003])
def test_sort_sizes(cuda, n):
    _check(n, 45, seed=n)


@pytest.mark.parametrize("end_bit", [1, 7, 8, 9, 16, 33, 44, 45, 47, 62])
def test_sort_bit_ranges(cuda, end_bit):
    _check(50_000, end_bit, seed=end_bit)


def test_sort_duplicates_are_stable(cuda):
    _check(300_000, 45, seed=1, key_bits=6)      # 64 distinct keys: long equal runs
    _check(300_000, 45, seed=2, runs=True)


def test_sort_full_size(cuda):
    _check(20_000_000, 45, seed=7, runs=True)    # BASELINE config-2 scale
