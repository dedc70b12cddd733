"""The exchange step of view-parallel training as one fused kernel per rank (include/gsr_dist.h, csrc/exchange.cu):
reduce-scatter of the gradient bucket -> Adam on the owner's shard -> all-gather of the updated parameters, plus the
SUM / MAX all-reduce of the densification statistics, over NVLink / NVSwitch peer memory.

Replaces `dist.all_reduce(bucket)` + `GaussianAdam.step()` on every rank + `DensificationStats.allreduce()` (three more
collectives) of gsr_b200/dist.py -- the semantics of /root/reference/train.py:163-186 applied to the sum of the step's
views.  torch.distributed is plumbing only: its symmetric-memory allocator maps every rank's buffers into every
process (and, on NVSwitch systems, gives the multicast mapping multimem.ld_reduce / multimem.st need); the kernel and
its barriers are this library's.  No fallback: without the library or without peer access this raises.
"""
import ctypes
from ctypes import POINTER, Structure, c_double, c_float, c_int, c_size_t, c_void_p

import torch
import torch.distributed as dist

from . import capi
from .dist import SEGMENTS, FlatGaussians, _pad
from .optim import DEFAULT_LRS

MAX_RANKS = 16
ADAM, SUM, MAX = 0, 1, 2
STAT_SEGMENTS = (("stat_accum", SUM), ("stat_denom", SUM), ("stat_radii", MAX))


class ExchangeSegment(Structure):
    _fields_ = [("begin", c_size_t), ("end", c_size_t), ("kind", c_int), ("cols", c_int), ("split_col", c_int),
                ("lr_a", c_float), ("lr_b", c_float)]


class ExchangeArgs(Structure):
    _fields_ = [("rank", c_int), ("world", c_int), ("n", c_size_t), ("grad_peer", c_void_p * MAX_RANKS),
                ("param_peer", c_void_p * MAX_RANKS), ("grad_mc", c_void_p), ("param_mc", c_void_p),
                ("signal_peer", c_void_p * MAX_RANKS), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("segments", POINTER(ExchangeSegment)), ("n_segments", c_int), ("step", c_int),
                ("beta1", c_double), ("beta2", c_double), ("eps", c_double)]


_bound = False


def _lib():
    global _bound
    L = capi.lib()
    if not _bound:
        L.gsr_exchange_signal_words.restype = c_size_t
        L.gsr_exchange_shard.argtypes = [c_size_t, c_int, c_int, POINTER(c_size_t), POINTER(c_size_t)]
        L.gsr_exchange_step.restype = c_int
        L.gsr_exchange_step.argtypes = [POINTER(ExchangeArgs), c_void_p]
        _bound = True
    return L


def shard(n, rank, world):
    """Floats [begin, end) of the flat buffers that `rank` owns (gsr_exchange_shard)."""
    b, e = c_size_t(0), c_size_t(0)
    _lib().gsr_exchange_shard(n, rank, world, ctypes.byref(b), ctypes.byref(e))
    return b.value, e.value


def layout(P):
    """Offsets (floats) of the parameter segments followed by the three statistic segments; total length."""
    off, offsets = 0, {}
    for name, n in SEGMENTS:
        offsets[name] = off
        off += _pad(P * n)
    n_param = off
    for name, _ in STAT_SEGMENTS:
        offsets[name] = off
        off += _pad(P)
    return offsets, n_param, off


class SymmetricGaussians(FlatGaussians):
    """FlatGaussians whose gradient bucket (+ the densification-statistic deltas behind it) and parameter buffer live in
    symmetric memory: every rank's copy is mapped into every process of `group`.  With world size 1 (or no process
    group) the buffers are ordinary device tensors."""

    def __init__(self, P, device, group=None):
        self.P, self.device = P, device
        self.group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world = dist.get_world_size(self.group) if self.group is not None else 1
        self.rank = dist.get_rank(self.group) if self.group is not None else 0
        self.offsets, self.n_param, self.n_grad = layout(P)
        n_sig = int(_lib().gsr_exchange_signal_words())
        self._n_total = self.n_grad + self.n_param + n_sig         # [gradients + statistics | parameters | signal words]
        if self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            try:
                symm_mem.enable_symm_mem_for_group(self.group.group_name)
            except Exception as e:                                   # noqa: BLE001 -- re-raised unless it is the benign case
                # a second SymmetricGaussians on the same group (FusedExchange.remap after a densification) finds the
                # group enabled already; anything else is a real set-up failure and must not be swallowed
                if "already" not in str(e).lower():
                    raise RuntimeError("symmetric memory cannot be enabled for process group %r: %s"
                                       % (self.group.group_name, e)) from e
            self._storage = symm_mem.empty(self._n_total, dtype=torch.float32, device=device)
            self._storage.zero_()
            self._hdl = symm_mem.rendezvous(self._storage, self.group.group_name)
            peers = [int(p) for p in self._hdl.buffer_ptrs]
            mc = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
        else:
            self._storage = torch.zeros(self._n_total, dtype=torch.float32, device=device)
            self._hdl, peers, mc = None, [self._storage.data_ptr()], 0
        self.grad_all = self._storage[:self.n_grad]
        self.bucket = self._storage[:self.n_param]
        self.params = self._storage[self.n_grad:self.n_grad + self.n_param]
        self.peer_grad = peers
        self.peer_param = [p + 4 * self.n_grad for p in peers]
        self.peer_signal = [p + 4 * (self.n_grad + self.n_param) for p in peers]
        self.mc_grad = mc
        self.mc_param = mc + 4 * self.n_grad if mc else 0
        torch.cuda.synchronize(device)
        if self.world > 1:
            dist.barrier(self.group)            # every rank's buffers are zeroed before anybody signals into them

    def stat_views(self):
        """Step-local densification-statistic deltas (rank-local sums over this rank's views): [P,1], [P,1], [P]."""
        v = lambda name: self._storage[self.offsets[name]: self.offsets[name] + self.P]
        return v("stat_accum").view(self.P, 1), v("stat_denom").view(self.P, 1), v("stat_radii")


class FusedExchange:
    """Per-step driver of gsr_exchange_step on a SymmetricGaussians.

    step()          reduce-scatter -> sharded Adam -> all-gather (+ statistics): after the kernel every rank holds the
                    updated parameters; the gradient bucket is consumed (the next step's first backward overwrites it).
    allreduce()     in-place SUM all-reduce of the bucket only (no optimizer), the exchange `dist.all_reduce` did.
    Adam hyper-parameters and learning rates are GaussianAdam's (gaussian_model.py:178-201).  Moments live on the shard's
    owner only: 2 x 59 P / W floats per GPU."""

    def __init__(self, flat, lrs=None, spatial_lr_scale=1.0, betas=(0.9, 0.999), eps=1e-15, use_multicast=None):
        """use_multicast: True / False force the data path (multimem through the switch / peer loads and stores);
        None (default) measures both once on the zeroed gradient bucket and keeps the faster one -- which one wins
        depends on the world size (2 GPUs: peer, 1.12 vs 1.85 ms for 708 MB; profiles/r02_summary.md)."""
        assert isinstance(flat, SymmetricGaussians)
        self.flat, self.betas, self.eps, self.step_count = flat, betas, eps, 0
        self.lrs = dict(DEFAULT_LRS)
        self.lrs.update(lrs or {})
        self.lrs["xyz"] = self.lrs["xyz"] * spatial_lr_scale
        b, e = shard(flat.n_grad, flat.rank, flat.world)
        self.shard = (b, e)
        self.exp_avg = torch.zeros(max(e - b, 4), dtype=torch.float32, device=flat.device)
        self.exp_avg_sq = torch.zeros_like(self.exp_avg)
        self.calibration_ms = None
        if use_multicast is None and flat.world > 1 and flat.mc_grad:
            self.use_multicast = self._calibrate()
        else:
            self.use_multicast = bool(use_multicast and flat.mc_grad)

    def _calibrate(self):
        """Times the in-place all-reduce of the (zeroed) bucket on both data paths; every rank takes the same decision
        (MAX over ranks of the device times)."""
        f = self.flat
        f.grad_all.zero_()
        times = []
        for mc in (False, True):
            self.use_multicast = mc
            for _ in range(2):
                self._launch(False, False)
            torch.cuda.synchronize(f.device)
            dist.barrier(f.group)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(4):
                self._launch(False, False)
            b.record()
            torch.cuda.synchronize(f.device)
            t = torch.tensor([a.elapsed_time(b) / 4], device=f.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=f.group)
            times.append(float(t.item()))
        self.calibration_ms = {"peer": times[0], "multicast": times[1]}
        return times[1] < times[0]

    def set_lr(self, name, lr):
        self.lrs[name] = float(lr)

    def _segments(self, adam, stats):
        f = self.flat
        lr_of = dict(xyz=("xyz", "xyz", 3), shs=("f_dc", "f_rest", 3), opacity=("opacity", "opacity", 1),
                     scales=("scaling", "scaling", 3), rotations=("rotation", "rotation", 4))
        cols = dict(SEGMENTS)
        segs = []
        for name, _ in SEGMENTS:
            a, b, split = lr_of[name]
            begin = f.offsets[name]
            segs.append(ExchangeSegment(begin, begin + _pad(f.P * cols[name]), ADAM if adam else SUM, cols[name], split,
                                        self.lrs[a], self.lrs[b]))
        if stats:
            for name, kind in STAT_SEGMENTS:
                begin = f.offsets[name]
                segs.append(ExchangeSegment(begin, begin + _pad(f.P), kind, 1, 1, 0.0, 0.0))
        return (ExchangeSegment * len(segs))(*segs)

    def _launch(self, adam, stats):
        f = self.flat
        segs = self._segments(adam, stats)
        a = ExchangeArgs()
        a.rank, a.world, a.n = f.rank, f.world, f.n_grad
        for w in range(f.world):
            a.grad_peer[w], a.param_peer[w], a.signal_peer[w] = f.peer_grad[w], f.peer_param[w], f.peer_signal[w]
        a.grad_mc = f.mc_grad if self.use_multicast else None
        a.param_mc = f.mc_param if self.use_multicast else None
        a.exp_avg, a.exp_avg_sq = self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr()
        a.segments, a.n_segments = segs, len(segs)
        a.step, a.beta1, a.beta2, a.eps = max(self.step_count, 1), self.betas[0], self.betas[1], self.eps
        with torch.cuda.device(f.device):
            capi._check(_lib().gsr_exchange_step(ctypes.byref(a), capi._stream()))

    def step(self, stats=False):
        self.step_count += 1
        self._launch(True, stats)

    def allreduce(self, stats=False):
        self._launch(False, stats)

    def remap(self, keep_mask=None, new_rows=None):
        """Carries the sharded optimizer across a densification event; returns a NEW FusedExchange on a NEW
        SymmetricGaussians of P' = keep_mask.sum() + n_new rows (this object and its `flat` are dead afterwards).

        The reference edits the optimizer state row by row when the set of Gaussians changes
        (/root/reference/train.py:163-174): `_prune_optimizer` keeps the masked rows of every parameter and of its
        exp_avg / exp_avg_sq (/root/reference/scene/gaussian_model.py:316-331), `cat_tensors_to_optimizer` appends the new
        rows with ZERO moments (:344-363), and `densification_postfix` / `prune_points` start the three statistic
        accumulators over at the new size (:381-392).  Here the moments are sharded by flat-buffer offset and the offsets
        depend on P, so the shards are first brought back to row space (every rank contributes its shard to a zeroed
        full-length buffer, summed over the group -- x + 0 is exact), pruned / extended exactly as above, laid out for P',
        and re-sharded.  The step count (Adam's bias correction), learning rates and the chosen data path carry over.

        keep_mask  bool[P] (True = the Gaussian stays), None = keep all.
        new_rows   {"xyz": [n,3], "shs": [n,16,3], "opacity": [n,1], "scales": [n,3], "rotations": [n,4]} appended after
                   the kept rows, None = nothing appended.
        COLLECTIVE: every rank of the group calls it with the same mask and rows (the replicas are identical)."""
        f = self.flat
        dev = f.device
        keep = torch.ones(f.P, dtype=torch.bool, device=dev) if keep_mask is None else keep_mask.to(device=dev, dtype=torch.bool)
        assert keep.numel() == f.P
        names = [name for name, _ in SEGMENTS]
        n_new = 0
        if new_rows:
            assert sorted(new_rows) == sorted(names), "new_rows needs exactly the segments %s" % (names,)
            n_new = int(new_rows["xyz"].shape[0])
            assert all(int(new_rows[k].shape[0]) == n_new for k in names)
        kept = int(keep.sum().item())
        P2 = kept + n_new
        torch.cuda.synchronize(dev)
        if f.world > 1:
            dist.barrier(f.group)                          # nobody is still inside the previous exchange kernel
        b, e = self.shard
        full = torch.zeros(2, f.n_grad, dtype=torch.float32, device=dev)
        full[0, b:e] = self.exp_avg[:e - b]
        full[1, b:e] = self.exp_avg_sq[:e - b]
        if f.world > 1:
            dist.all_reduce(full, op=dist.ReduceOp.SUM, group=f.group)
        flat2 = SymmetricGaussians(P2, dev, f.group)       # zeroed: gradients, statistic segments, parameters
        full2 = torch.zeros(2, flat2.n_grad, dtype=torch.float32, device=dev)
        cols = dict(SEGMENTS)
        for name in names:
            old = f._view(f.params, name)[keep]
            if n_new:
                old = torch.cat([old, new_rows[name].to(device=dev, dtype=torch.float32).reshape((n_new,) + tuple(old.shape[1:]))])
            flat2._view(flat2.params, name).copy_(old)
            c, o, o2 = cols[name], f.offsets[name], flat2.offsets[name]
            for k in range(2):                             # new rows keep the zeros (cat_tensors_to_optimizer)
                full2[k, o2:o2 + kept * c].view(kept, c).copy_(full[k, o:o + f.P * c].view(f.P, c)[keep])
        opt = FusedExchange(flat2, lrs=self.lrs, spatial_lr_scale=1.0, betas=self.betas, eps=self.eps,
                            use_multicast=self.use_multicast)
        opt.calibration_ms = self.calibration_ms
        opt.step_count = self.step_count
        b2, e2 = opt.shard
        opt.exp_avg[:e2 - b2] = full2[0, b2:e2]
        opt.exp_avg_sq[:e2 - b2] = full2[1, b2:e2]
        torch.cuda.synchronize(dev)
        if f.world > 1:
            dist.barrier(f.group)
        return opt
