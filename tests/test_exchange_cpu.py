"""Host logic of the fused exchange step (include/gsr_dist.h, gsr_b200/exchange.py) without a GPU: shard table, segment
table, argument validation, and -- world size 2 over gloo -- that "reduce-scatter -> Adam on the owner's shard ->
all-gather" over the kernel's own shard / segment tables equals "all-reduce -> Adam on every replica" (the semantics of
train.py:163-186 the kernel must keep)."""
import ctypes
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cases  # noqa: F401
from gsr_b200 import exchange
from oracle import optim_oracle


def test_shards_tile_the_buffer():
    for n in (0, 4, 64, 4096, 59 * 1000 + 4, 106_200_000):
        n -= n % 4
        for world in (1, 2, 3, 4, 8, 16):
            edges = [exchange.shard(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            for (b0, e0), (b1, e1) in zip(edges, edges[1:]):
                assert e0 == b1
            assert all((b % 128 == 0 or b == n) and (e % 128 == 0 or e == n) and b <= e for b, e in edges)
            sizes = [e - b for b, e in edges]
            assert max(sizes) <= (n // 4 + world - 1) // world * 4 + 128       # ceil(n / world), rounded up to 512 bytes


def test_layout_matches_flat_gaussians():
    from gsr_b200.dist import FlatGaussians, SEGMENTS
    P = 1001
    offsets, n_param, n_grad = exchange.layout(P)
    f = FlatGaussians(P, "cpu")
    assert {k: offsets[k] for k, _ in SEGMENTS} == f.offsets and n_param == f.bucket.numel()
    assert n_grad > n_param and all(offsets[k] % 64 == 0 for k in offsets)
    assert offsets["stat_radii"] + P <= n_grad


def test_argument_validation_without_gpu():
    L = exchange._lib()
    assert L.gsr_exchange_signal_words() >= 148 * 16
    a = exchange.ExchangeArgs()
    a.rank, a.world, a.n = 2, 2, 64
    assert L.gsr_exchange_step(ctypes.byref(a), None) != 0 and b"rank" in L.gsr_last_error()
    a.rank, a.world, a.n = 0, 1, 66
    assert L.gsr_exchange_step(ctypes.byref(a), None) != 0 and b"multiple of 4" in L.gsr_last_error()
    a.n = 64
    segs = (exchange.ExchangeSegment * 1)(exchange.ExchangeSegment(0, 64, exchange.ADAM, 3, 3, 0.1, 0.1))
    a.segments, a.n_segments = segs, 1
    assert L.gsr_exchange_step(ctypes.byref(a), None) != 0 and b"peer pointer" in L.gsr_last_error()


def _worker(rank, world, port, P, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        offsets, n_param, n_grad = exchange.layout(P)
        rng = np.random.default_rng(100)
        params = rng.standard_normal(n_param).astype(np.float32)          # identical replicas
        g_rank = np.random.default_rng(200 + rank).standard_normal(n_param).astype(np.float32)
        # (a) what every replica would do after an all-reduce: dense Adam over the whole bucket
        g_sum = torch.from_numpy(g_rank.copy())
        dist.all_reduce(g_sum)
        g_sum = g_sum.numpy()
        lr = 0.01
        p_a, m_a, v_a = params.copy(), np.zeros(n_param, np.float32), np.zeros(n_param, np.float32)
        optim_oracle.adam_step(p_a, g_sum, m_a, v_a, 1, lr)
        # (b) the fused step's data flow on the kernel's own shard table
        b, e = exchange.shard(n_grad, rank, world)
        b, e = min(b, n_param), min(e, n_param)
        shards = [exchange.shard(n_grad, r, world) for r in range(world)]
        mine = torch.zeros(e - b)
        for r, (rb, re_) in enumerate(shards):                             # reduce-scatter
            rb, re_ = min(rb, n_param), min(re_, n_param)
            piece = torch.from_numpy(g_rank[rb:re_].copy())
            dist.reduce(piece, dst=r)
            if r == rank:
                mine = piece
        p_b = params.copy()
        ps, ms, vs = p_b[b:e].copy(), np.zeros(e - b, np.float32), np.zeros(e - b, np.float32)
        optim_oracle.adam_step(ps, mine.numpy(), ms, vs, 1, lr)            # Adam on the owner's shard only
        for r, (rb, re_) in enumerate(shards):                             # all-gather of the updated parameters
            rb, re_ = min(rb, n_param), min(re_, n_param)
            piece = torch.from_numpy(ps.copy()) if r == rank else torch.zeros(re_ - rb)
            dist.broadcast(piece, src=r)
            p_b[rb:re_] = piece.numpy()
        out[rank] = bool(np.array_equal(p_a, p_b)) and bool(np.array_equal(m_a[b:e], ms))
    finally:
        dist.destroy_process_group()


def test_sharded_adam_equals_replicated_adam_gloo_world2():
    world, port = 2, 29641
    with mp.Manager() as man:
        out = man.dict()
        mp.spawn(_worker, args=(world, port, 777, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}
