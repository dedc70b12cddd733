"""The fused exchange kernel (include/gsr_dist.h, csrc/exchange.cu) on the GPU.

World size 1 degenerates to the plain Adam step: bit-identical to gsr_adam_step.  With >= 2 visible GPUs the test spawns
one process per GPU (NCCL only for the rendezvous of torch's symmetric memory) and checks the N-rank step -- peer
loads/stores and, where the box offers it, the multicast (multimem) path -- against all-reduce + dense Adam, and the
SUM / MAX statistic segments against torch.distributed all-reduces."""
import os

import numpy as np
import pytest
import torch

import cases  # noqa: F401
from gsr_b200 import capi, exchange, synth
from gsr_b200.dist import FlatGaussians
from gsr_b200.optim import GaussianAdam

pytestmark = pytest.mark.gpu


def _random_grads(flat, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    # the padding between the segments stays zero (no kernel ever writes a gradient there): a zero gradient with zero
    # moments leaves a parameter unchanged, so whole buffers can be compared
    flat.bucket.zero_()
    for v in flat.grad_views().values():
        v.copy_(torch.randn(v.shape, device="cuda", generator=g))
        v[::3] = 0      # a share of exactly-zero rows (Gaussians no view touched), as in training


def test_world1_equals_adam_step(cuda):
    P = 20_001
    scene = synth.trained_like_scene(P, 5, s0=0.03)
    ref = FlatGaussians.from_numpy(scene, cuda)
    sym = exchange.SymmetricGaussians.from_numpy(scene, cuda)
    assert torch.equal(ref.params, sym.params)
    opt_ref, opt = GaussianAdam(ref), exchange.FusedExchange(sym)
    for step in range(3):
        _random_grads(ref, 10 + step)
        sym.bucket.copy_(ref.bucket)
        opt_ref.step()
        opt.step()
        torch.cuda.synchronize()
        assert torch.equal(ref.params, sym.params), "step %d" % step
    b, e = opt.shard
    assert (b, e) == (0, sym.n_grad)
    assert torch.equal(opt_ref.exp_avg, opt.exp_avg[:ref.params.numel()])
    assert torch.equal(opt_ref.exp_avg_sq, opt.exp_avg_sq[:ref.params.numel()])


def _worker(rank, world, port, use_mc, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        P = 50_003
        scene = synth.trained_like_scene(P, 5, s0=0.03)
        ref = FlatGaussians.from_numpy(scene, dev)
        sym = exchange.SymmetricGaussians.from_numpy(scene, dev)
        opt_ref = GaussianAdam(ref)
        opt = exchange.FusedExchange(sym, use_multicast=use_mc)
        if use_mc and not opt.use_multicast:
            out[rank] = "no-multicast"
            return
        ok = True
        for step in range(3):
            g = torch.Generator(device=dev).manual_seed(100 * step + rank)
            ref.bucket.zero_()
            for v in ref.grad_views().values():
                v.copy_(torch.randn(v.shape, device=dev, generator=g))
                v[::5] = 0
            sym.bucket.copy_(ref.bucket)
            acc, den, rad = sym.stat_views()
            acc.copy_(torch.rand(P, 1, device=dev, generator=g)); den.copy_((torch.rand(P, 1, device=dev, generator=g) > 0.5).float())
            rad.copy_(torch.randint(0, 200, (P,), device=dev, generator=g).float())
            want = [acc.clone(), den.clone(), rad.clone()]
            dist.all_reduce(want[0]); dist.all_reduce(want[1]); dist.all_reduce(want[2], op=dist.ReduceOp.MAX)
            # reference: gradients summed over ranks, then dense Adam on every replica.  The kernel's peer path adds in
            # rank order, so the reference does too (bit-exact for every world size); the multicast path's order is
            # the switch's: compared with a tolerance, and an element whose sum nearly cancels may even take the
            # opposite Adam step (sign of a ~1e-8 gradient), which is counted, not forbidden.
            parts = [torch.empty_like(ref.bucket) for _ in range(world)]
            dist.all_gather(parts, ref.bucket)
            total = parts[0].clone()
            for w_ in range(1, world):
                total += parts[w_]
            ref.bucket.copy_(total)
            opt_ref.step()
            opt.step(stats=True)
            torch.cuda.synchronize(dev)
            dist.barrier()
            if not opt.use_multicast:
                same = torch.equal(ref.params, sym.params)
            else:
                d = (ref.params - sym.params).abs()
                off = d > 1e-5 * (1.0 + ref.params.abs())
                same = bool(off.float().mean() < 1e-3) and bool(d.max() <= 2 * 0.025 * (step + 1) + 1e-6)
            exact_sum = world == 2
            acc_ok = torch.equal(acc, want[0]) if exact_sum else bool(((acc - want[0]).abs() <= 1e-5 * want[0].abs() + 1e-6).all())
            ok = ok and same and acc_ok and torch.equal(den, want[1]) and torch.equal(rad, want[2])
        # plain all-reduce mode
        g = torch.Generator(device=dev).manual_seed(999 + rank)
        sym.bucket.copy_(torch.randn(sym.bucket.numel(), device=dev, generator=g))
        want = sym.bucket.clone()
        dist.all_reduce(want)
        before = sym.params.clone()
        opt.allreduce()
        torch.cuda.synchronize(dev)
        dist.barrier()
        close = torch.equal(want, sym.bucket) if world == 2 else bool(((want - sym.bucket).abs() <= 1e-5 * want.abs() + 1e-6).all())
        ok = ok and close and torch.equal(before, sym.params)
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("use_mc", [False, True], ids=["peer", "multicast"])
def test_multi_gpu_step_equals_allreduce_plus_adam(cuda, use_mc):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs on the box (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    with mp.Manager() as man:
        out = man.dict()
        mp.spawn(_worker, args=(world, 29733 + int(use_mc), use_mc, out), nprocs=world, join=True)
        res = dict(out)
    if use_mc and all(v == "no-multicast" for v in res.values()):
        pytest.skip("symmetric memory on this box has no multicast mapping")
    assert res == {r: True for r in range(world)}, res
