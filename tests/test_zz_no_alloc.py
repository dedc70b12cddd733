"""gsr_sort_pairs_static -- the 64-bit sort for callers that cannot allocate device memory.

Some users of libgsr_b200.so are given their input and output buffers by a host application and have
no device allocator of their own (a plugin under an engine that owns the heap, a frame loop that bans
allocator calls once it runs, an FFI binding without torch).  For them the sort's histogram / ticket /
look-back status block is an array built into the library rather than a `temp` argument.  This file
checks, through the C ABI, that the variant returns exactly what gsr_sort_pairs returns (both are
integer sorts, so the comparison is bit-for-bit), on two calls in a row -- the second call finds the
first call's leftovers in the shared array and must still be right -- and that sizes beyond the
built-in maximum are refused with a message instead of overrunning the array.

The file name sorts last on purpose: pytest -x stops at the first failure, and a problem here should
not hide the results of the parity tests.  It passes alone in a fresh process (the first call then
does the per-device address look-up) and after the other GPU tests.
"""
import pytest

import cases  # noqa: F401  (puts the package on sys.path, like the other GPU tests)

pytestmark = pytest.mark.gpu


def _keys(n, seed, runs):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    if runs:   # the rasterizer's key shape: tile id in the high word, runs of equal depth bits below
        depth = torch.randint(0, 1 << 31, (n // 6 + 1,), generator=g, device="cuda").repeat_interleave(6)[:n]
        tile = torch.randint(0, 8160, (n,), generator=g, device="cuda")
        return (tile << 32) | depth
    return torch.randint(0, 1 << 62, (n,), generator=g, device="cuda")


def test_static_scratch_sort_matches_gsr_sort_pairs_twice_in_a_row(cuda):
    import torch
    from gsr_b200 import capi
    n_max = int(capi.lib().gsr_sort_static_max_pairs())
    assert n_max == 1 << 20
    # two back-to-back calls of different size, key shape and pass count (45 bits = 6 passes, result copied
    # back from the ping-pong input; 33 bits = 5 passes, result already in the output), then the maximum size
    for n, end_bit, seed, runs in [(n_max - 7, 45, 11, True), (300_001, 33, 12, False), (n_max, 62, 13, False), (1, 45, 14, False)]:
        keys = _keys(n, seed, runs)
        vals = torch.arange(n, dtype=torch.int32, device="cuda")
        want_k, want_v = capi.sort_pairs(keys.clone(), vals.clone(), end_bit)          # existing entry point
        kin, vin = keys.clone(), vals.clone()
        ko, vo = torch.empty_like(keys), torch.empty_like(vals)                        # caller-owned outputs
        before = torch.cuda.memory_allocated()
        capi.sort_pairs_static(kin, vin, ko, vo, end_bit)
        assert torch.cuda.memory_allocated() == before                                 # nothing drawn from torch
        torch.cuda.synchronize()
        assert torch.equal(ko, want_k), (n, end_bit)
        assert torch.equal(vo, want_v), (n, end_bit)
        # and against an independent stable sort, so that a shared bug cannot hide
        masked = keys & ((1 << end_bit) - 1)
        assert torch.equal(vo.long(), torch.sort(masked, stable=True).indices)


def test_static_scratch_sort_refuses_sizes_beyond_the_built_in_array(cuda):
    import torch
    from gsr_b200 import capi
    n = int(capi.lib().gsr_sort_static_max_pairs()) + 1
    keys = torch.zeros(n, dtype=torch.int64, device="cuda")
    vals = torch.zeros(n, dtype=torch.int32, device="cuda")
    ko, vo = torch.full_like(keys, -1), torch.full_like(vals, -1)
    with pytest.raises(RuntimeError, match="GSR_SORT_STATIC_MAX_PAIRS"):
        capi.sort_pairs_static(keys, vals, ko, vo, 45)
    torch.cuda.synchronize()
    assert bool((ko == -1).all()) and bool((vo == -1).all())                           # nothing was launched
    capi.sort_pairs_static(keys[:0], vals[:0], ko[:0], vo[:0], 45)                     # n = 0 is a no-op
