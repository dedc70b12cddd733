"""Device time of the gradient exchange alone on N GPUs (torchrun): this library's fused kernel (in-place all-reduce mode and
reduce-scatter -> sharded Adam -> all-gather mode, peer and multicast data paths) against ncclAllReduce.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/exchange_probe.py [P]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist
from gsr_b200 import exchange

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", device_id=dev)
P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
G = exchange.SymmetricGaussians(P, dev)
G.params.normal_()


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


res = {"world": world, "P": P, "bucket_bytes": int(G.bucket.numel() * 4)}
for mc in (False, True):
    fx = exchange.FusedExchange(G, lrs={k: 0.0 for k in ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")}, use_multicast=mc)
    if mc and not fx.use_multicast:
        res["multicast"] = "unavailable"
        continue
    tag = "multicast" if mc else "peer"
    G.grad_all.zero_()
    res[tag + "_allreduce_ms"] = timed(fx.allreduce)
    res[tag + "_step_ms"] = timed(lambda: fx.step(stats=True))
G.bucket.zero_()
res["nccl_allreduce_ms"] = timed(lambda: dist.all_reduce(G.bucket))
res["nccl_reduce_scatter_plus_all_gather_ms"] = None
b, e = exchange.shard(G.n_param, rank, world)
if (e - b) * world == G.n_param:
    shard_buf = torch.empty(e - b, device=dev)
    def rs_ag():
        dist.reduce_scatter_tensor(shard_buf, G.bucket)
        dist.all_gather_into_tensor(G.params, shard_buf)
    res["nccl_reduce_scatter_plus_all_gather_ms"] = timed(rs_ag)
if rank == 0:
    for k in list(res):
        if k.endswith("_ms") and res[k]:
            res[k.replace("_ms", "_GBps_algo")] = res["bucket_bytes"] / res[k] / 1e6
    print(json.dumps(res))
dist.destroy_process_group()
