"""The optimizer the reference attaches to the Gaussians, on the flat bucket of gsr_b200.dist.FlatGaussians.

Mirrors GaussianModel.training_setup (/root/reference/scene/gaussian_model.py:178-201): Adam(lr=0, eps=1e-15) with the
groups xyz / f_dc / f_rest / opacity / scaling / rotation and their learning rates
(/root/reference/arguments/__init__.py OptimizationParams), stepped as /root/reference/train.py:176-186 does.
One kernel launch updates all 59 floats of every Gaussian (csrc/adam.cu); `step(visible=...)` is the sparse form
train.py:179-181 asks for (`optimizer.step(visible, N)`), `zero_grad=True` clears the consumed gradients in the same
pass (optimizer.zero_grad()).  No CPU fallback.
"""
import torch

from . import capi
from .dist import SEGMENTS

DEFAULT_LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20.0, opacity=0.025, scaling=0.005, rotation=0.001)


class GaussianAdam:
    def __init__(self, flat, lrs=None, spatial_lr_scale=1.0, betas=(0.9, 0.999), eps=1e-15):
        self.flat, self.betas, self.eps, self.step_count = flat, betas, eps, 0
        self.lrs = dict(DEFAULT_LRS)
        self.lrs.update(lrs or {})
        self.lrs["xyz"] = self.lrs["xyz"] * spatial_lr_scale
        self.exp_avg = torch.zeros_like(flat.params)
        self.exp_avg_sq = torch.zeros_like(flat.params)

    def set_lr(self, name, lr):
        """update_learning_rate (gaussian_model.py:203-209) sets the xyz group's rate every iteration."""
        self.lrs[name] = float(lr)

    def _segments(self):
        f = self.flat
        v = lambda buf, name: f._view(buf, name).reshape(f.P, -1)
        lr_of = dict(xyz=("xyz", "xyz", 3), shs=("f_dc", "f_rest", 3), opacity=("opacity", "opacity", 1),
                     scales=("scaling", "scaling", 3), rotations=("rotation", "rotation", 4))
        segs = []
        for name, _ in SEGMENTS:
            a, b, split = lr_of[name]
            segs.append(dict(param=v(f.params, name), grad=v(f.bucket, name), exp_avg=v(self.exp_avg, name),
                             exp_avg_sq=v(self.exp_avg_sq, name), lr_a=self.lrs[a], lr_b=self.lrs[b], split_col=split))
        return segs

    def step(self, visible=None, zero_grad=False):
        self.step_count += 1
        capi.adam_step(self._segments(), self.step_count, self.betas[0], self.betas[1], self.eps, visible=visible,
                       zero_grad=zero_grad)

    def zero_grad(self):
        self.flat.bucket.zero_()


class SparseGaussianAdam(torch.optim.Adam):
    """The optimizer interface the reference's training loop already knows:
        self.optimizer = SparseGaussianAdam(l, lr=0.0, eps=1e-15)        /root/reference/scene/gaussian_model.py:194-198
        gaussians.optimizer.step(visible, radii.shape[0])                /root/reference/train.py:179-181
    A torch.optim.Adam subclass (so the densification code that edits `optimizer.state[param]["exp_avg"/"exp_avg_sq"]`
    and `param_groups`, gaussian_model.py:316-392, works unchanged) whose step updates ONLY the rows of visible
    Gaussians, all parameter groups in one launch (gsr_adam_step).  The update is torch.optim.Adam's (bias-corrected,
    the reference's default optimizer) restricted to those rows; rows outside the mask keep parameter and moments bit
    for bit.  Not exported from diff_gaussian_rasterization by default, exactly like the reference's own rasterizer
    package (exporting it switches train.py to its separate_sh path); GSR_EXPORT_SPARSE_ADAM=1 opts in.  No CPU path."""

    def __init__(self, params, lr, eps):
        super().__init__(params=params, lr=lr, eps=eps)

    @torch.no_grad()
    def step(self, visibility, N):
        by_step, writeback = {}, []
        for group in self.param_groups:
            assert len(group["params"]) == 1, "more than one tensor in group"
            param = group["params"][0]
            if param.grad is None:
                continue
            if not param.is_cuda:
                raise RuntimeError("SparseGaussianAdam needs CUDA parameters: there is no CPU path")
            state = self.state[param]
            if "exp_avg" not in state:                       # lazy init, as torch.optim.Adam does
                state["exp_avg"] = torch.zeros_like(param, memory_format=torch.preserve_format)
                state["exp_avg_sq"] = torch.zeros_like(param, memory_format=torch.preserve_format)
            step = int(state["step"]) + 1 if "step" in state else 1
            state["step"] = torch.tensor(float(step))
            if param.shape[0] != N:
                raise RuntimeError("SparseGaussianAdam.step: parameter group '%s' has %d rows, N = %d"
                                   % (group.get("name", "?"), param.shape[0], N))
            # on the separate_sh path the gradients of _features_dc / _features_rest are slices of the [P,16,3]
            # gradient the rasterizer returned for torch.cat([dc, shs]) (not contiguous): pack them once
            grad = param.grad if param.grad.is_contiguous() else param.grad.contiguous()
            v2 = lambda t: t.view(N, -1)
            tensors = (param.data, state["exp_avg"], state["exp_avg_sq"])
            if all(x.is_contiguous() for x in tensors):
                work = tensors
            else:
                # Until the first densification `_xyz` is column-major: create_from_pcd wraps
                # torch.tensor(np.vstack([x, y, z]).T), which keeps numpy's strides (dataset_readers.py:120-127,
                # gaussian_model.py:150), and zeros_like(preserve_format) gives the moments the same layout.  The
                # kernel walks row-major rows, so such a group steps on packed copies that are written back.
                work = tuple(x.contiguous() for x in tensors)
                writeback.append((tensors, work))
            seg = dict(param=v2(work[0]), grad=v2(grad), exp_avg=v2(work[1]), exp_avg_sq=v2(work[2]), lr=group["lr"])
            by_step.setdefault((step, tuple(group["betas"]), group["eps"]), []).append(seg)
        vis = visibility if visibility is None or visibility.dtype in (torch.bool, torch.uint8) else visibility != 0
        for (step, betas, eps), segs in by_step.items():     # one launch unless groups have diverged step counts
            capi.adam_step(segs, step, betas[0], betas[1], eps, visible=None if vis is None else vis.contiguous())
        for tensors, work in writeback:
            for dst, src in zip(tensors, work):
                dst.copy_(src)
