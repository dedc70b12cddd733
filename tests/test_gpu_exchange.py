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


def _densify_inputs(P, dev, seed):
    """A densification event: about a fifth of the Gaussians pruned, some hundred new ones appended."""
    g = torch.Generator(device=dev).manual_seed(seed)
    keep = torch.rand(P, device=dev, generator=g) > 0.2
    n_new = 777
    shapes = dict(xyz=(n_new, 3), shs=(n_new, 16, 3), opacity=(n_new, 1), scales=(n_new, 3), rotations=(n_new, 4))
    new_rows = {k: torch.randn(v, device=dev, generator=g) * 0.1 for k, v in shapes.items()}
    return keep, new_rows


def _dense_remap(ref, opt_ref, keep, new_rows):
    """What the reference does to a dense optimizer at a densification (scene/gaussian_model.py:316-331 keeps the masked
    rows of the parameters and both moments, :344-363 appends the new rows with zero moments), on the flat layout."""
    P2 = int(keep.sum().item()) + new_rows["xyz"].shape[0]
    ref2 = FlatGaussians(P2, ref.device)
    opt2 = GaussianAdam(ref2, lrs=opt_ref.lrs, betas=opt_ref.betas, eps=opt_ref.eps)
    opt2.step_count = opt_ref.step_count
    for name in new_rows:
        ref2._view(ref2.params, name).copy_(torch.cat([ref._view(ref.params, name)[keep], new_rows[name]]))
        for src, dst in ((opt_ref.exp_avg, opt2.exp_avg), (opt_ref.exp_avg_sq, opt2.exp_avg_sq)):
            kept = ref._view(src, name)[keep]
            ref2._view(dst, name).copy_(torch.cat([kept, torch.zeros_like(new_rows[name])]))
    return ref2, opt2


def test_world1_remap_across_a_densification(cuda):
    """FusedExchange.remap: after pruning and appending rows the fused step keeps tracking the dense Adam whose state the
    reference's _prune_optimizer / cat_tensors_to_optimizer edited -- parameters and both moments, bit for bit."""
    P = 20_001
    scene = synth.trained_like_scene(P, 5, s0=0.03)
    ref = FlatGaussians.from_numpy(scene, cuda)
    sym = exchange.SymmetricGaussians.from_numpy(scene, cuda)
    opt_ref, opt = GaussianAdam(ref), exchange.FusedExchange(sym)
    opt.set_lr("xyz", 3e-4); opt_ref.set_lr("xyz", 3e-4)
    for step in range(2):
        _random_grads(ref, 20 + step)
        sym.bucket.copy_(ref.bucket)
        opt_ref.step(); opt.step()
    keep, new_rows = _densify_inputs(P, cuda, 3)
    ref, opt_ref = _dense_remap(ref, opt_ref, keep, new_rows)
    opt = opt.remap(keep, new_rows)
    sym = opt.flat
    assert sym.P == ref.P == int(keep.sum()) + 777 and opt.step_count == 2 and opt.lrs["xyz"] == 3e-4
    assert torch.equal(ref.params, sym.params)
    assert float(sym.grad_all.abs().max()) == 0.0           # gradients and statistic segments start over
    for step in range(2):
        _random_grads(ref, 30 + step)
        sym.bucket.copy_(ref.bucket)
        opt_ref.step(); opt.step()
        torch.cuda.synchronize()
        assert torch.equal(ref.params, sym.params), "step %d after the remap" % step
    n = ref.params.numel()
    assert torch.equal(opt_ref.exp_avg, opt.exp_avg[:n]) and torch.equal(opt_ref.exp_avg_sq, opt.exp_avg_sq[:n])
    # prune only / append nothing, and the identity remap
    opt2 = opt.remap(None, None)
    assert torch.equal(opt2.flat.params, sym.params) and torch.equal(opt2.exp_avg, opt.exp_avg)


def _reference_sum(ref, sym, opt, world, dev):
    """Leaves the rank-summed gradients in ref.bucket, added in the order the fused step adds them, and every rank's
    own gradients in sym.bucket.  Peer path: rank order (the kernel's loop over peers).  Multicast path: the order is the
    switch's, so the sums are taken by the exchange's plain all-reduce mode on the same addresses (multimem.ld_reduce
    without the optimizer) and the rank-local gradients are put back afterwards; the all-reduce mode itself is checked
    against dist.all_reduce at the end of the worker."""
    import torch.distributed as dist
    if opt.use_multicast:
        opt.allreduce()
        torch.cuda.synchronize(dev)
        dist.barrier()
        switch_sum = sym.bucket.clone()
        sym.bucket.copy_(ref.bucket)
        ref.bucket.copy_(switch_sum)
        torch.cuda.synchronize(dev)
        dist.barrier()
        return
    parts = [torch.empty_like(ref.bucket) for _ in range(world)]
    dist.all_gather(parts, ref.bucket)
    total = parts[0].clone()
    for w_ in range(1, world):
        total += parts[w_]
    ref.bucket.copy_(total)


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
            # reference: gradients summed over ranks, then dense Adam on every replica -- with the sums taken in the
            # kernel's own order (_reference_sum), so parameters must come out bit for bit on both data paths
            _reference_sum(ref, sym, opt, world, dev)
            opt_ref.step()
            opt.step(stats=True)
            torch.cuda.synchronize(dev)
            dist.barrier()
            same = torch.equal(ref.params, sym.params)
            exact_sum = world == 2
            acc_ok = torch.equal(acc, want[0]) if exact_sum else bool(((acc - want[0]).abs() <= 1e-5 * want[0].abs() + 1e-6).all())
            ok = ok and same and acc_ok and torch.equal(den, want[1]) and torch.equal(rad, want[2])
        # a densification event in the middle of training (FusedExchange.remap), then two more steps: the sharded moments
        # must land on the right rows of the new layout on every rank.  Exact on the peer path.
        keep, new_rows = _densify_inputs(P, dev, 4)          # same seed on every rank: the replicas stay identical
        ref, opt_ref = _dense_remap(ref, opt_ref, keep, new_rows)
        opt = opt.remap(keep, new_rows)
        sym = opt.flat
        ok = ok and torch.equal(ref.params, sym.params)
        for step in range(3, 5):
            g = torch.Generator(device=dev).manual_seed(100 * step + rank)
            ref.bucket.zero_()
            for v in ref.grad_views().values():
                v.copy_(torch.randn(v.shape, device=dev, generator=g))
                v[::5] = 0
            sym.bucket.copy_(ref.bucket)
            _reference_sum(ref, sym, opt, world, dev)
            opt_ref.step()
            opt.step()
            torch.cuda.synchronize(dev)
            dist.barrier()
            ok = ok and torch.equal(ref.params, sym.params)
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
        n = ref.params.numel()                              # the owner's moment shard against the dense moments
        b, e = opt.shard
        e = min(e, n)
        if e > b:
            ok = ok and torch.equal(opt_ref.exp_avg[b:e], opt.exp_avg[:e - b]) and torch.equal(opt_ref.exp_avg_sq[b:e], opt.exp_avg_sq[:e - b])
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
