"""View-parallel multi-GPU training plumbing; one process per GPU.

The reference is single-GPU (utils/general_utils.py:133 pins cuda:0).  Training partitions over
camera views: Gaussians are replicated on every GPU, each rank renders its own views through the
unchanged rasterizer boundary, and the parameter gradients -- 59 fp32 per Gaussian
(xyz 3, SH 48 = f_dc 3 + f_rest 45, opacity 1, scale 3, rotation 4; scene/gaussian_model.py:178-201)
-- are summed with ONE NCCL all-reduce over a flat bucket before the optimizer step.
Densification statistics are per-view quantities reduced separately (train.py:163-174,
scene/gaussian_model.py:471-473): grad-norm accumulators and visibility counts with SUM,
max_radii2D with MAX.

The rasterizer's backward writes its gradient rows straight into views of the bucket (no copy).
"""
import numpy as np
import torch
import torch.distributed as dist

SEGMENTS = (("xyz", 3), ("shs", 48), ("opacity", 1), ("scales", 3), ("rotations", 4))
FLOATS_PER_GAUSSIAN = sum(n for _, n in SEGMENTS)   # 59


def _pad(n, a=64):
    return (n + a - 1) // a * a


class FlatGaussians:
    """Replicated Gaussian parameters + one flat, segment-aligned gradient bucket."""

    def __init__(self, P, device):
        self.P, self.device = P, device
        self.offsets, off = {}, 0
        for name, n in SEGMENTS:
            self.offsets[name] = off
            off += _pad(P * n)
        self.bucket = torch.zeros(off, dtype=torch.float32, device=device)
        self.params = torch.zeros(off, dtype=torch.float32, device=device)

    def _view(self, flat, name):
        n = dict(SEGMENTS)[name]
        v = flat[self.offsets[name]: self.offsets[name] + self.P * n]
        return v.view(self.P, 16, 3) if name == "shs" else v.view(self.P, n)

    @classmethod
    def from_numpy(cls, scene, device):
        P = scene["means3D"].shape[0]
        self = cls(P, device)
        src = dict(xyz=scene["means3D"], shs=scene["shs"], opacity=scene["opacities"], scales=scene["scales"],
                   rotations=scene["rotations"])
        for name, _ in SEGMENTS:
            self._view(self.params, name).copy_(torch.from_numpy(np.ascontiguousarray(src[name])).reshape(
                self._view(self.params, name).shape))
        return self

    # parameter views (rasterizer inputs)
    xyz = property(lambda s: s._view(s.params, "xyz"))
    shs = property(lambda s: s._view(s.params, "shs"))
    opacity = property(lambda s: s._view(s.params, "opacity"))
    scales = property(lambda s: s._view(s.params, "scales"))
    rotations = property(lambda s: s._view(s.params, "rotations"))

    def grad_views(self):
        """Views of the bucket keyed like capi.backward's outputs."""
        return dict(dL_dmeans3D=self._view(self.bucket, "xyz"), dL_dsh=self._view(self.bucket, "shs"),
                    dL_dopacity=self._view(self.bucket, "opacity"), dL_dscales=self._view(self.bucket, "scales"),
                    dL_drotations=self._view(self.bucket, "rotations"))

    def allreduce_from(self, grads):
        """Sum this step's gradients over ranks.  `grads` are capi.backward outputs; rows that were
        not written into the bucket directly are copied in first."""
        for k, v in self.grad_views().items():
            if grads[k].data_ptr() != v.data_ptr():
                v.copy_(grads[k].view(v.shape))
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.bucket, op=dist.ReduceOp.SUM)
        return self.bucket

    def as_leaf_params(self):
        """Leaf tensors (requires_grad) aliasing the parameter storage, for the autograd API path."""
        return {name: self._view(self.params, name).detach().requires_grad_(True) for name, _ in SEGMENTS}

    def allreduce_param_grads(self, params):
        if not (dist.is_initialized() and dist.get_world_size() > 1):
            return
        for name, _ in SEGMENTS:
            g = params[name].grad
            if g is not None:
                self._view(self.bucket, name).copy_(g.view(self._view(self.bucket, name).shape))
        dist.all_reduce(self.bucket, op=dist.ReduceOp.SUM)
        for name, _ in SEGMENTS:
            if params[name].grad is not None:
                params[name].grad.copy_(self._view(self.bucket, name).view(params[name].grad.shape))


def balance_views(costs, world, per_rank=None):
    """Deals the views of a step to `world` ranks so that the ranks' summed costs are as equal as a greedy pass gets them:
    views in decreasing cost order, each to the rank with the smallest sum that still has room (LPT with equal counts).
    `costs[i]` is any predictor of view i's time -- the instance count N of the last time that camera was rendered works
    (the step ends with the slowest rank, SURVEY.md section 8(e): "assign views to GPUs by predicted N").  Deterministic:
    every rank computes the same assignment from the same costs.  Returns a list of `world` lists of view indices."""
    n = len(costs)
    per_rank = per_rank if per_rank is not None else (n + world - 1) // world
    assert per_rank * world >= n
    order = sorted(range(n), key=lambda i: (-float(costs[i]), i))
    sums, out = [0.0] * world, [[] for _ in range(world)]
    for i in order:
        r = min((r for r in range(world) if len(out[r]) < per_rank), key=lambda r: (sums[r], r))
        out[r].append(i)
        sums[r] += float(costs[i])
    return [sorted(v) for v in out]


class ViewStreams:
    """Renders the views of one step (forward + backward, gradients accumulated in one bucket) on several CUDA streams.

    The views of a step are independent until their gradients meet in the bucket, so view v runs on stream v % S: the
    short latency-bound stages of one view (scans, the sorts of ~1 M keys, the mid-forward count readback the host waits
    for) execute under the blend kernels of its neighbour.  One host thread drives all streams; ordering is by events:
      * side streams start after everything already queued on the caller's stream (previous step's all-reduce / Adam);
      * view 0's backward OVERWRITES the bucket (no separate zero fill), so no other backward may start before it ends;
        later backwards only add (RED), in any order;
      * the caller's stream continues after the last kernel of every side stream.
    """

    def __init__(self, device, n_streams):
        self.side = [torch.cuda.Stream(device=device) for _ in range(max(1, n_streams) - 1)]

    def run(self, n_views, forward, backward):
        """forward(v) -> state and backward(v, state, accumulate) -> grads are called under the view's stream."""
        main = torch.cuda.current_stream()
        streams = [main] + self.side
        if self.side:
            start = torch.cuda.Event()
            start.record(main)
            for s in self.side:
                s.wait_event(start)
        first_done, out = None, None
        for v in range(n_views):
            s = streams[v % len(streams)]
            with torch.cuda.stream(s):
                st = forward(v)
                if v > 0 and s is not main and first_done is not None:
                    s.wait_event(first_done)
                out = (st, backward(v, st, v > 0))
                if v == 0 and self.side:
                    first_done = torch.cuda.Event()
                    first_done.record(s)
        for s in self.side:
            done = torch.cuda.Event()
            done.record(s)
            main.wait_event(done)
        return out


class DensificationStats:
    """Densification statistics (scene/gaussian_model.py:471-473, train.py:166) for view-parallel
    steps.  Each rank accumulates this step's views into step-local deltas; `allreduce()` reduces the
    deltas over ranks (SUM for the grad-norm accumulator and the visibility count, MAX for the radii)
    and folds them into the replicated running state, so N ranks x V views give exactly what one GPU
    rendering the same N*V views sequentially would have accumulated.  Norms are taken per view,
    BEFORE any reduction (summing 2-D gradients first would change densification)."""

    def __init__(self, P, device):
        z = lambda *s: torch.zeros(*s, device=device)
        self.xyz_gradient_accum, self.denom, self.max_radii2D = z(P, 1), z(P, 1), z(P)
        self._d_accum, self._d_denom, self._d_radii = z(P, 1), z(P, 1), z(P)

    def add_view(self, grad_means2D, radii):
        if grad_means2D.is_cuda:
            # one fused launch (csrc/adam.cu: densification_stats_kernel) instead of three indexed updates
            from . import capi
            capi.densification_stats(grad_means2D.contiguous(), radii.contiguous(), self._d_accum, self._d_denom, self._d_radii)
            return
        vis = radii > 0      # host tensors (the gloo CPU tests of the reduction logic)
        self._d_accum[vis] += torch.norm(grad_means2D[vis, :2], dim=-1, keepdim=True)
        self._d_denom[vis] += 1
        self._d_radii[vis] = torch.max(self._d_radii[vis], radii[vis].float())

    def allreduce(self):
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self._d_accum, op=dist.ReduceOp.SUM)
            dist.all_reduce(self._d_denom, op=dist.ReduceOp.SUM)
            dist.all_reduce(self._d_radii, op=dist.ReduceOp.MAX)
        self.xyz_gradient_accum += self._d_accum
        self.denom += self._d_denom
        self.max_radii2D = torch.max(self.max_radii2D, self._d_radii)
        self._d_accum.zero_(); self._d_denom.zero_(); self._d_radii.zero_()
