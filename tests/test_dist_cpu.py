"""View-parallel plumbing on CPU: world_size-2 gloo.  N ranks x V views must reproduce what one
process accumulating the same N*V views sequentially gets (gradient bucket SUM, densification stats
SUM / MAX with per-view norms taken before the reduction)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cases  # noqa: F401
from gsr_b200.dist import DensificationStats, FlatGaussians, FLOATS_PER_GAUSSIAN, SEGMENTS

P = 257   # deliberately not a multiple of the segment padding


def _fake_view(view, P):
    """Deterministic stand-in for one view's rasterizer backward outputs."""
    g = torch.Generator().manual_seed(1000 + view)
    grads = dict(dL_dmeans3D=torch.randn(P, 3, generator=g), dL_dsh=torch.randn(P, 16, 3, generator=g),
                 dL_dopacity=torch.randn(P, 1, generator=g), dL_dscales=torch.randn(P, 3, generator=g),
                 dL_drotations=torch.randn(P, 4, generator=g))
    radii = torch.randint(0, 40, (P,), generator=g, dtype=torch.int32)
    radii[torch.rand(P, generator=g) < 0.4] = 0
    g2d = torch.randn(P, 3, generator=g)
    return grads, radii, g2d


def _worker(rank, world, port, views_per_rank, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    G = FlatGaussians(P, torch.device("cpu"))
    S = DensificationStats(P, torch.device("cpu"))
    for step in range(2):
        G.bucket.zero_()
        for v in range(views_per_rank):
            grads, radii, g2d = _fake_view(step * 100 + rank * views_per_rank + v, P)
            for k, view in G.grad_views().items():
                view += grads[k]
            S.add_view(g2d, radii)
        G.allreduce_from({k: v for k, v in G.grad_views().items()})
        S.allreduce()
        if step == 0:
            first = G.bucket.clone()
    torch.save(dict(bucket=G.bucket, first=first, accum=S.xyz_gradient_accum, denom=S.denom, radii=S.max_radii2D),
               os.path.join(out_dir, "rank%d.pt" % rank))
    dist.destroy_process_group()


def test_two_rank_allreduce_equals_sequential(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world, vpr = 2, 2
    mp.spawn(_worker, args=(world, port, vpr, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = (torch.load(os.path.join(tmp_path, "rank%d.pt" % r)) for r in range(world))
    for k in r0:
        assert torch.equal(r0[k], r1[k]), "ranks disagree on " + k
    # sequential single-process reference over the same views
    G = FlatGaussians(P, torch.device("cpu"))
    S = DensificationStats(P, torch.device("cpu"))
    for step in range(2):
        G.bucket.zero_()
        for v in range(world * vpr):
            grads, radii, g2d = _fake_view(step * 100 + v, P)
            for k, view in G.grad_views().items():
                view += grads[k]
            S.add_view(g2d, radii)
        S.allreduce()
    assert torch.allclose(r0["bucket"], G.bucket, rtol=1e-6, atol=1e-6)
    assert torch.allclose(r0["accum"], S.xyz_gradient_accum, rtol=1e-6, atol=1e-6)
    assert torch.equal(r0["denom"], S.denom) and torch.equal(r0["radii"], S.max_radii2D)


def test_flat_layout():
    G = FlatGaussians(P, torch.device("cpu"))
    assert FLOATS_PER_GAUSSIAN == 59
    seen = torch.zeros_like(G.bucket)
    for name, n in SEGMENTS:
        v = G._view(G.bucket, name)
        assert v.numel() == P * n and v.is_contiguous() and v.data_ptr() % 16 == 0
        v += 1
    assert int(G.bucket.sum()) == 59 * P
    sc = dict(means3D=np.random.rand(P, 3).astype(np.float32), shs=np.random.rand(P, 16, 3).astype(np.float32),
              opacities=np.random.rand(P, 1).astype(np.float32), scales=np.random.rand(P, 3).astype(np.float32),
              rotations=np.random.rand(P, 4).astype(np.float32))
    G2 = FlatGaussians.from_numpy(sc, torch.device("cpu"))
    assert np.array_equal(G2.shs.numpy(), sc["shs"]) and np.array_equal(G2.rotations.numpy(), sc["rotations"])
    leaves = G2.as_leaf_params()
    assert all(p.requires_grad and p.is_leaf for p in leaves.values())


def test_balance_views_equalises_rank_sums():
    from gsr_b200.dist import balance_views
    rng = np.random.default_rng(3)
    costs = rng.uniform(5e6, 9e6, 64)
    naive = [list(range(r * 8, r * 8 + 8)) for r in range(8)]
    bal = balance_views(costs, 8, 8)
    assert sorted(i for v in bal for i in v) == list(range(64)) and all(len(v) == 8 for v in bal)
    spread = lambda a: max(costs[v].sum() for v in a) / costs.sum() * 8
    assert spread(bal) < 1.01 < spread(naive)
    assert balance_views(costs, 8, 8) == bal                      # deterministic
    assert balance_views([3, 1, 2], 2, 2) == [[0], [1, 2]]
