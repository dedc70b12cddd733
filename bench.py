#!/usr/bin/env python
"""bench.py -- fwd+bwd renders/s of the differentiable Gaussian-splatting rasterizer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c1|c3|c4]

Workload is BASELINE.json configs[1] ("c2"): 3M SH-3 Gaussians, single-view forward + backward at 1920x1080.
One "step" = the V = 8 views of this rank (distinct cameras), each view ALONE on one CUDA stream, one after another
(nothing of two views overlaps), gradients summed in one flat bucket; with N > 1 ranks the step ends with the gradient
exchange (SUM over ranks of the 59 fp32/Gaussian bucket).  Fixed work per GPU => weak scaling, `value` = views/s over all
ranks; at N = 1 that is 1 / (single-view forward+backward time).  Prints ONE JSON line (rank 0).  DESIGN.md section 6.

  value            device-timed (CUDA events around the K steps) through the C ABI, inputs resident in HBM
  e2e              the same step through the public Python API (diff_gaussian_rasterization autograd Function) with the
                   per-view inputs (camera + target image) copied from pinned host memory and the loss read back to the
                   host inside the timed region; `stock_autograd` = the same without the grad_sink extension (gradients
                   returned to autograd, what the reference's unchanged train.py does)
  with_optimizer   the step followed by the densification statistics and the Adam step (N > 1: the fused exchange kernel
                   reduce-scatter -> sharded Adam -> all-gather; N = 1: gsr_adam_step)
  overlapped_step  the same 8 views spread over 3 CUDA streams (gsr_b200.dist.ViewStreams)
  batched_step     the same 8 views through gsr_forward_batch / gsr_backward_batch (one per-Gaussian launch for all cameras)
  c3_strong        BASELINE.json configs[2] as written: 1.8M Gaussians, 979x546, 8 views per step IN TOTAL split over the
                   N ranks, gradient exchange every step (strong scaling), with and without the optimizer
  roofline         the dominant kernel (backward blend, FP32-pipe bound) + roofline_stages for every stage
  cpu_baseline     the CPU oracle (a port of the reference algorithm) on one view of the workload: all host threads, and
                   1 thread; plus the fork's own forward-only host renderer restated (gsiron_host, 1 thread, C1)
  --impl reference : the reference algorithm's CPU implementation (the oracle port), all host threads, one view per step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gaussian-splatting-npu_b200")
for _p in (ROOT, PKG, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from gsr_b200 import synth  # noqa: E402

WORKLOADS = {
    # name: (P, W, H, seed, s0, antialiasing, n_cams)
    "c1": dict(P=100_000, W=800, H=800, seed=1001, kind="random_init", aa=False,
               desc="nerf_synthetic lego camera 800x800, 100K random-init SH3 Gaussians"),
    "c2": dict(P=3_000_000, W=1920, H=1080, seed=2002, s0=0.004, kind="trained", aa=False,
               desc="Mip-NeRF360-shape: 3M SH3 Gaussians, single view 1920x1080"),
    "c3": dict(P=1_800_000, W=979, H=546, seed=3003, s0=0.006, kind="trained", aa=False, extent=1.5,
               desc="T&T-shape: 1.8M SH3 Gaussians, 979x546"),
    "c5": dict(P=10_000_000, W=1920, H=1080, seed=5009, s0=0.004, kind="trained", aa=False,
               desc="forward sweep, largest size: 10M SH3 Gaussians, 1920x1080"),
    "c4": dict(P=6_000_000, W=3840, H=2160, seed=4004, s0=0.004, kind="trained", aa=True,
               desc="6M SH3 Gaussians at 3840x2160, antialiasing + inverse depth"),
}

FP32_LANES = 148 * 128 * 2     # FMA lanes x 2 flop


def make_workload(name, cam_index=0, n_cams=8):
    """Scene + camera `cam_index` of an orbit of `n_cams` cameras (every rank of a run passes the same n_cams = the
    step's global view count, so all views of a step are distinct)."""
    w = WORKLOADS[name]
    if w["kind"] == "random_init":
        scene = synth.random_init_scene(w["P"], w["seed"])
    else:
        scene = synth.trained_like_scene(w["P"], w["seed"], s0=w["s0"], extent=w.get("extent", 1.0))
    return w, scene, make_workload_camera(name, cam_index, n_cams)


def make_workload_camera(name, cam_index, n_cams=8):
    w = WORKLOADS[name]
    if w["kind"] == "random_init":
        import cases
        return cases._lego_cam(1 + cam_index % 99, w["W"], w["H"])       # lego transforms_train.json holds 100 frames
    n_cams = max(8, n_cams)
    assert 0 <= cam_index < n_cams
    return synth.orbit_cameras(n_cams, w["W"], w["H"], w["seed"])[cam_index]


def static_config(args, world):
    """`config` of the JSON line: a function of the command line only, so both arms (--impl ours / reference) print the
    same object.  Measured quantities (visible Gaussians, instance counts, ...) go to `workload_stats`."""
    w = WORKLOADS[args.workload]
    V = max(1, args.views_per_rank)
    return {"workload": args.workload + ": " + w["desc"], "P": w["P"], "width": w["W"], "height": w["H"],
            "antialiasing": w["aa"], "sh_degree": 3, "views_per_rank_per_step": V, "global_views_per_step": V * world,
            "cameras": ("orbit of %d distinct cameras" % max(8, V * world)) + (
                ", rank r renders cameras r*%d .. r*%d+%d" % (V, V, V - 1) if world == 1 or args.no_balance else
                ", %d per rank, dealt to the ranks by predicted cost (instance count of a probing forward; gsr_b200.dist.balance_views)" % V),
            "step": "each rank renders its %d views one after another on ONE CUDA stream (each view alone on the GPU: "
                    "forward + backward, parameter gradients summed in one flat 59-fp32/Gaussian bucket)%s" % (
                        V, "; then the gradient exchange over the %d ranks (SUM of the bucket)" % world if world > 1 else ""),
            "parallelism": ("view-parallel dp%d, replicated Gaussians" % world) if world > 1 else "single GPU",
            "l2": "inputs (708 MB of Gaussians at C2 + scratch per view) exceed the 126 MB L2; no flush needed"}


def oracle_step(scene, cam, w, grads):
    """One fwd+bwd of the CPU oracle; returns (seconds, forward dict)."""
    from oracle import gs_oracle as go
    S = go.Scene(W=cam.image_width, H=cam.image_height, means3D=scene["means3D"], opacities=scene["opacities"],
                 viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, cam_pos=cam.camera_center,
                 tan_fovx=cam.tanfovx, tan_fovy=cam.tanfovy, background=np.zeros(3, np.float32), shs=scene["shs"],
                 scales=scene["scales"], rotations=scene["rotations"], sh_degree=scene["sh_degree"],
                 antialiasing=w["aa"])
    t0 = time.perf_counter()
    f = go.forward(S, borderline=False)
    go.backward(S, f, grads[0], grads[1])
    return time.perf_counter() - t0, f


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region.  NVML is read in-process from a
    background thread (pynvml: two cheap queries every 50 ms); spawning `nvidia-smi -lms` beside a
    millisecond-scale step perturbs it (each nvidia-smi query stalled the GPU for several ms here),
    so nvidia-smi is only the fallback when pynvml is unavailable."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.index, self.sm, self.reasons, self.max_mhz = index, [], set(), None
        self.stop_flag, self.thread, self.proc, self.rows = False, None, None, []

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # CUDA_VISIBLE_DEVICES may remap indices: resolve through torch's device UUID when possible
            h = None
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def loop():
                while not self.stop_flag:
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        for bit, name in self.REASONS.items():
                            if r & bit:
                                self.reasons.add(name)
                    except Exception:
                        pass
                    time.sleep(0.05)
            self.thread = threading.Thread(target=loop, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.thread = None
        try:
            q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
                "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "500"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: [self.rows.append([c.strip() for c in l.split(",")]) for l in self.proc.stdout],
                             daemon=True).start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=1)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"]}
        time.sleep(0.6)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = {nme for r in self.rows for k, nme in enumerate(names) if len(r) > 2 + k and r[2 + k].lower().startswith("active")}
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


def loss_stage_report(dev, H, W, hbm_peak, reps=30):
    """Loss stage next to the path: fused L1 + D-SSIM forward + backward on one [3,H,W] view.
    CUDA-event times, L2 flushed between repetitions (the 25 MB images would otherwise stay in the 126 MB L2).
    Rows: ours fused (2 kernels + 1 reduction), ours through the reference's fused_ssim API + torch L1, and -- when
    oracle/_ref holds it -- the unmodified reference fused-ssim CUDA kernels + torch L1 on this GPU."""
    import torch
    from gsr_b200 import capi
    from fused_ssim import fused_ssim as our_fused_ssim
    torch.manual_seed(3)
    img = torch.rand(3, H, W, device=dev)
    gt = (img + 0.05 * torch.randn_like(img)).clamp(0, 1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    scratch = capi.LossScratch((3, H, W), dev)
    n = 3 * H * W

    def timed(fn):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(reps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    def ours_fused():
        capi.l1_ssim_loss_forward(img, gt, 0.2, scratch, train=True)
        capi.l1_ssim_loss_backward(img, gt, 0.2, scratch, upstream=None)

    def composed(ssim_fn):
        def run():
            a = img.detach().requires_grad_(True)
            loss = 0.8 * torch.abs(a - gt).mean() + 0.2 * (1.0 - ssim_fn(a.unsqueeze(0), gt.unsqueeze(0)))
            loss.backward()
        return run

    t_fwd = timed(lambda: capi.l1_ssim_loss_forward(img, gt, 0.2, scratch, train=True))
    t_bwd = timed(lambda: capi.l1_ssim_loss_backward(img, gt, 0.2, scratch, upstream=None))
    t_ours = timed(ours_fused)
    # algorithmic bytes per pixel-channel: fwd reads image + gt (8) and writes 3 derivative maps (12);
    # bwd reads the 3 maps + image + gt (20) and writes the gradient (4)
    bytes_fwd, bytes_bwd = 20 * n, 24 * n
    rep = {"workload": "one [3,%d,%d] view, loss = 0.8*L1 + 0.2*(1-SSIM), forward + backward" % (H, W),
           "ours_fused_ms": t_ours, "ours_fused_fwd_ms": t_fwd, "ours_fused_bwd_ms": t_bwd, "gpu_launches": 3,
           "roofline": {"bound": "hbm", "algorithmic_bytes": bytes_fwd + bytes_bwd, "achieved": (bytes_fwd + bytes_bwd) / (t_fwd + t_bwd) / 1e6,
                        "peak": hbm_peak, "unit": "GB/s", "frac": (bytes_fwd + bytes_bwd) / (t_fwd + t_bwd) / 1e6 / hbm_peak,
                        "fwd_frac": bytes_fwd / t_fwd / 1e6 / hbm_peak, "bwd_frac": bytes_bwd / t_bwd / 1e6 / hbm_peak},
           "ours_reference_api_ms": timed(composed(our_fused_ssim))}
    # render epilogue (exposure affine + clamp, gaussian_renderer/__init__.py:112-119): ours vs the reference's torch lines
    expo = (torch.eye(3, 4, device=dev) + 0.05 * torch.randn(3, 4, device=dev)).contiguous()
    dL = torch.randn_like(img)
    pre = img * 1.4 - 0.2

    def ours_epilogue():
        capi.exposure_clamp_forward(pre, expo)
        capi.exposure_clamp_backward(pre, expo, dL)

    def torch_epilogue():
        a, e = pre.detach().requires_grad_(True), expo.detach().requires_grad_(True)
        y = (torch.matmul(a.permute(1, 2, 0), e[:3, :3]).permute(2, 0, 1) + e[:3, 3, None, None]).clamp(0, 1)
        y.backward(dL)

    t_ep = timed(ours_epilogue)
    rep["epilogue"] = {"workload": "exposure affine + clamp, forward + backward (image and exposure gradients)",
                       "ours_ms": t_ep, "reference_torch_lines_ms": timed(torch_epilogue), "gpu_launches": 3,
                       "algorithmic_bytes": 20 * n, "hbm_frac": 20 * n / t_ep / 1e6 / hbm_peak}
    try:
        from oracle import build_ref
        ref = build_ref.load_fused_ssim()
        if ref is not None:
            class RefMap(torch.autograd.Function):      # FS/fused_ssim/__init__.py:8-31 over the reference extension
                @staticmethod
                def forward(ctx, a, b):
                    m, d0, d1, d2 = ref.fusedssim(0.01 ** 2, 0.03 ** 2, a, b, True)
                    ctx.save_for_backward(a.detach(), b, d0, d1, d2)
                    return m

                @staticmethod
                def backward(ctx, g):
                    a, b, d0, d1, d2 = ctx.saved_tensors
                    return ref.fusedssim_backward(0.01 ** 2, 0.03 ** 2, a, b, g.contiguous(), d0, d1, d2), None
            rep["reference_cuda_same_box_ms"] = timed(composed(lambda a, b: RefMap.apply(a, b).mean()))
    except Exception as exc:  # informational leg only
        rep["reference_cuda_same_box_error"] = str(exc)[:200]
    return rep


def frame_stage_report(dev, H, W, hbm_peak, reps=30):
    """Frame output after a forward-only render: [3,H,W] fp32 -> [H,W,3] u8, ours (one launch)
    vs the reference's torch expressions (train.py:81 viewer bytes; torchvision save_image's quantisation, render.py:44)."""
    import torch
    from gsr_b200 import frames
    img = torch.rand(3, H, W, device=dev) * 1.2 - 0.1
    out8 = torch.empty(H, W, 3, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(reps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    t_ours = timed(lambda: frames.to_rgb8(img, frames.TRUNCATE, out=out8))
    t_gui = timed(lambda: (torch.clamp(img, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous())
    t_png = timed(lambda: img.mul(255).add_(0.5).clamp_(0, 255).permute(1, 2, 0).to(torch.uint8).contiguous())
    nbytes = 15 * H * W
    return {"workload": "one [3,%d,%d] fp32 render -> [H,W,3] u8 frame" % (H, W), "ours_ms": t_ours, "gpu_launches": 1,
            "reference_torch_viewer_bytes_ms": t_gui, "reference_torch_save_image_quantise_ms": t_png,
            "roofline": {"bound": "hbm", "algorithmic_bytes": nbytes, "achieved": nbytes / t_ours / 1e6, "peak": hbm_peak,
                         "unit": "GB/s", "frac": nbytes / t_ours / 1e6 / hbm_peak}}


def optimizer_stage_report(G, radii, hbm_peak, reps=10):
    """Optimizer step after the backward: Adam over the 59 floats of every Gaussian.
    Rows: ours dense (one launch, gradient cleared in the same pass), ours sparse (rows of view 0's visible
    Gaussians only), and torch.optim.Adam (the reference's optimizer, gaussian_model.py:178-201) on the same GPU."""
    import torch
    from gsr_b200.optim import GaussianAdam, DEFAULT_LRS
    opt = GaussianAdam(G)
    n_par = G.P * 59
    vis = (radii > 0).contiguous()
    n_vis = int(vis.sum().item())

    def timed(fn):
        fn()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    saved = G.params.clone()
    t_dense = timed(lambda: opt.step(zero_grad=True))
    t_sparse = timed(lambda: opt.step(visible=vis, zero_grad=True))
    G.params.copy_(saved)
    del opt, saved
    names = dict(xyz="xyz", opacity="opacity", scales="scaling", rotations="rotation")
    tp = [torch.nn.Parameter(G._view(G.params, k).clone()) for k in names] + \
         [torch.nn.Parameter(G.shs[:, :1].clone()), torch.nn.Parameter(G.shs[:, 1:].clone())]
    lrs = [DEFAULT_LRS[v] for v in names.values()] + [DEFAULT_LRS["f_dc"], DEFAULT_LRS["f_rest"]]
    for q in tp:
        q.grad = torch.zeros_like(q)
    topt = torch.optim.Adam([{"params": [q], "lr": lr} for q, lr in zip(tp, lrs)], lr=0.0, eps=1e-15)

    def torch_step():
        topt.step()
        topt.zero_grad(set_to_none=False)
    t_torch = timed(torch_step)
    del topt, tp
    torch.cuda.empty_cache()
    bytes_dense = 32 * n_par          # read p,g,m,v + write p,m,v + cleared g
    return {"workload": "Adam step over %d Gaussians x 59 fp32 (+ gradient cleared)" % G.P, "ours_dense_ms": t_dense,
            "ours_sparse_ms": t_sparse, "visible_rows": n_vis, "torch_optim_adam_same_box_ms": t_torch, "gpu_launches": 1,
            "roofline": {"bound": "hbm", "algorithmic_bytes": bytes_dense, "achieved": bytes_dense / t_dense / 1e6,
                         "peak": hbm_peak, "unit": "GB/s", "frac": bytes_dense / t_dense / 1e6 / hbm_peak,
                         "sparse_achieved": 32 * 59 * n_vis / t_sparse / 1e6}}


def loss_cpu_baseline(H, W):
    """The loss stage's CPU oracle (port of the reference algorithm) on one full view, all host threads."""
    from oracle import ssim_oracle as so
    rng = np.random.default_rng(3)
    img = rng.random((3, H, W), dtype=np.float32)
    gt = np.clip(img + 0.05 * rng.standard_normal((3, H, W)).astype(np.float32), 0, 1)
    so.l1_ssim_loss(img[:, :64, :64], gt[:, :64, :64], 0.2)
    t0 = time.perf_counter()
    so.l1_ssim_loss(img, gt, 0.2)
    return {"ms": (time.perf_counter() - t0) * 1e3, "cores": so.num_threads(), "kind": "port",
            "sample": "one full [3,%d,%d] view, loss forward + backward" % (H, W)}


def host_threads():
    """Host threads this process may use (affinity mask), NOT OMP_NUM_THREADS: torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def run_reference(args):
    """--impl reference: the reference algorithm's CPU implementation (oracle port), all host threads.  One step = ONE
    view (forward + backward) of the step our arm renders -- a bounded sample of the same workload: the views of rank 0,
    cycled -- so that K steps end within minutes; value in the same unit (views/s)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    from oracle import gs_oracle as go
    go.set_num_threads(host_threads())      # explicit: independent of the launcher's OMP_NUM_THREADS
    V = max(1, args.views_per_rank)
    w, scene, cam = make_workload(args.workload, 0, V * world)
    cams = [cam] + [make_workload_camera(args.workload, v, V * world) for v in range(1, V)]
    grads = synth.upstream_grads(w["H"], w["W"], 5)
    for k in range(args.warmup):
        oracle_step(scene, cams[k % V], w, grads)
    t0 = time.perf_counter()
    for k in range(args.steps):
        oracle_step(scene, cams[k % V], w, grads)
    dt = time.perf_counter() - t0
    val = args.steps / dt
    cores = go.num_threads()
    sample = "one %dx%d view of %d Gaussians (fwd+bwd) per step, the %d cameras of rank 0 cycled; %d OpenMP threads (OMP_NUM_THREADS in the environment was %r)" % (
        w["W"], w["H"], w["P"], V, cores, os.environ.get("OMP_NUM_THREADS"))
    print(json.dumps({
        "impl": "reference", "metric": "fwd+bwd renders/s", "value": val, "unit": "renders/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": static_config(args, world),
        "cpu_baseline": {"value": val, "unit": "renders/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "renders/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


class Job:
    """This rank's share of one workload: the replicated Gaussians in the flat parameter / gradient-bucket layout, the
    rank's V cameras, one scratch workspace per stream, and the step variants the bench times."""

    def __init__(self, name, V, total_views, rank, world, dev, exchange, max_streams=3):
        import torch
        import torch.distributed as dist
        from gsr_b200 import capi
        from gsr_b200.dist import FlatGaussians, ViewStreams, DensificationStats
        self.torch, self.dist, self.capi = torch, dist, capi
        self.name, self.V, self.rank, self.world, self.dev, self.exchange = name, V, rank, world, dev, exchange
        self.w, self.scene, _ = make_workload(name, rank * V, total_views)
        self.total_views = total_views
        w = self.w
        t = self.t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        self.fused = None
        # The bench's parameters are ACTIVATED values (the rasterizer's inputs) and every step sees the same upstream
        # gradient, so the real learning rates would walk the scene away from the workload within a few steps (a scale
        # of 0.004 stepping by 0.005).  The timed optimizer steps therefore use the reference's rates x 1e-6: the same
        # arithmetic and memory traffic, parameters that stay where they are to 1e-6.
        from gsr_b200.optim import DEFAULT_LRS
        self.lrs = {k: v * 1e-6 for k, v in DEFAULT_LRS.items()}
        if world > 1 and exchange == "fused":
            from gsr_b200 import exchange as ex
            self.G = ex.SymmetricGaussians.from_numpy(self.scene, dev)
            self.fused = ex.FusedExchange(self.G, lrs=self.lrs)
        else:
            self.G = FlatGaussians.from_numpy(self.scene, dev)
        G = self.G
        self.bg = torch.zeros(3, device=dev)
        self.dcol_np, self.ddep_np = synth.upstream_grads(w["H"], w["W"], 5)
        self.dcol, self.ddep = t(self.dcol_np), t(self.ddep_np)
        self.shared = dict(means3D=G.xyz, opacities=G.opacity, background=self.bg, shs=G.shs, scales=G.scales,
                           rotations=G.rotations, sh_degree=3, antialiasing=w["aa"])
        self.set_views(list(range(rank * V, rank * V + V)))
        # every gradient output is a persistent tensor: the five parameter gradients are views of the exchange bucket,
        # the per-view intermediates (means2D, conic, ...) one scratch set per stream
        self.grad_views = G.grad_views()
        P = w["P"]
        e = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        self.view_scratch = [dict(dL_dmeans2D=e(P, 3), dL_dconic=e(P, 2, 2), dL_dcolors=e(P, 3), dL_dinvdepths=e(P, 1),
                                  dL_dcov3D=e(P, 6), **self.grad_views) for _ in range(max_streams)]
        self.spaces = [capi.Workspace(dev) for _ in range(max_streams)]
        self.streams = {1: ViewStreams(dev, 1), max_streams: ViewStreams(dev, max_streams)}
        self.stats = DensificationStats(P, dev)
        self.adam = None
        self.last_state = None

    def set_views(self, indices):
        """This rank's views of the step: indices into the step's orbit of `total_views` cameras."""
        self.view_indices = list(indices)
        self.cams = [make_workload_camera(self.name, i, self.total_views) for i in self.view_indices]
        t = self.t
        self.per_view = [dict(viewmatrix=t(c.world_view_transform), projmatrix=t(c.full_proj_transform),
                              cam_pos=t(c.camera_center), tan_fovx=c.tanfovx, tan_fovy=c.tanfovy, **self.shared) for c in self.cams]

    def balance(self):
        """Deals the step's views to the ranks by predicted cost (gsr_b200.dist.balance_views): one probing forward per
        view gives its instance count N, every rank gathers all counts and computes the same assignment."""
        if self.world == 1:
            return None
        from gsr_b200.dist import balance_views
        torch, dist = self.torch, self.dist
        mine = torch.zeros(self.total_views, device=self.dev)
        for v, gi in enumerate(self.view_indices):
            mine[gi] = float(self._fwd(v, 0)["num_rendered"])
        dist.all_reduce(mine)
        costs = mine.cpu().tolist()
        assign = balance_views(costs, self.world, self.V)
        before = max(sum(costs[r * self.V:(r + 1) * self.V]) for r in range(self.world))
        after = max(sum(costs[i] for i in a) for a in assign)
        self.set_views(assign[self.rank])
        return {"cost": "instance count N of a probing forward", "max_rank_sum_contiguous": before, "max_rank_sum_balanced": after,
                "mean_rank_sum": sum(costs) / self.world}

    # ---- one view / one step ------------------------------------------------------------------------------
    def _fwd(self, v, slot):
        return self.capi.forward(width=self.w["W"], height=self.w["H"], workspace=self.spaces[slot], **self.per_view[v])

    def _bwd(self, v, st, acc, slot):
        return self.capi.backward(st, dL_dcolor=self.dcol, dL_dinvdepth=self.ddep, out=self.view_scratch[slot],
                                  accumulate=acc, **self.per_view[v])

    def render_views(self, n_streams=1, with_stats=False):
        """The rank's V views: forward + backward each, the first backward overwrites the bucket, the others add."""
        S = n_streams

        def bwd(v, st, acc):
            g = self._bwd(v, st, acc, v % S)
            if with_stats:       # densification statistics of this view (train.py:166-167), step-local deltas
                acc_d, den_d, rad_d = self._stat_deltas()
                self.capi.densification_stats(g["dL_dmeans2D"], st["radii"], acc_d, den_d, rad_d)
            return g
        out = self.streams[S].run(self.V, lambda v: self._fwd(v, v % S), bwd)
        self.last_state = out[0]
        return out

    def _stat_deltas(self):
        if self.fused is not None:
            return self.G.stat_views()
        return self.stats._d_accum, self.stats._d_denom, self.stats._d_radii

    def exchange_gradients(self):
        """SUM of the bucket over ranks, nothing else (the step `value` times)."""
        if self.world == 1:
            return
        if self.fused is not None:
            self.fused.allreduce()
        else:
            self.dist.all_reduce(self.G.bucket, op=self.dist.ReduceOp.SUM)

    def step_render(self, n_streams=1):
        self.render_views(n_streams)
        self.exchange_gradients()

    def step_batched(self, n_streams=1):
        """The same V views through gsr_forward_batch / gsr_backward_batch: the per-Gaussian stages run ONCE for all
        cameras (parameters and SH rows read once, the 59 gradient floats written once), the rest per view."""
        capi, torch = self.capi, self.torch
        if not hasattr(self, "batch_ws"):
            P = self.w["P"]
            e = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.dev)
            self.batch_ws = [capi.Workspace(self.dev) for _ in range(self.V)]
            self.batch_scratch = [dict(dL_dmeans2D=e(P, 3), dL_dconic=e(P, 2, 2), dL_dcolors=e(P, 3), dL_dinvdepths=e(P, 1))
                                  for _ in range(self.V)]
        cam_kw = [{k: pv[k] for k in ("viewmatrix", "projmatrix", "cam_pos", "tan_fovx", "tan_fovy")} for pv in self.per_view]
        streams = None if n_streams <= 1 else [torch.cuda.current_stream()] + self.streams[n_streams].side
        states = capi.forward_batch(cam_kw, width=self.w["W"], height=self.w["H"], workspaces=self.batch_ws, streams=streams,
                                    **self.shared)
        capi.backward_batch(states, cam_kw, dL_dcolor=self.dcol, dL_dinvdepth=self.ddep, out=self.grad_views,
                            scratch=self.batch_scratch, streams=streams, **self.shared)
        self.last_state = states[-1]
        self.exchange_gradients()

    def step_train(self, n_streams=1):
        """Render + statistics + optimizer: what train.py:163-186 does after the backward, for the whole step."""
        torch = self.torch
        self.render_views(n_streams, with_stats=True)
        if self.fused is not None:
            self.fused.step(stats=True)             # reduce-scatter -> sharded Adam -> all-gather, + SUM / MAX statistics
            acc_d, den_d, rad_d = self.G.stat_views()
            st = self.stats
            st.xyz_gradient_accum += acc_d
            st.denom += den_d
            torch.maximum(st.max_radii2D, rad_d, out=st.max_radii2D)
            self.G.grad_all[self.G.n_param:].zero_()
        else:
            if self.world > 1:
                self.dist.all_reduce(self.G.bucket, op=self.dist.ReduceOp.SUM)
            if self.adam is None:
                from gsr_b200.optim import GaussianAdam
                self.adam = GaussianAdam(self.G, lrs=self.lrs)
            self.adam.step()
            self.stats.allreduce()

    def save_params(self):
        return self.G.params.clone()

    def restore_params(self, saved):
        self.G.params.copy_(saved)


def exchange_self_check(job, torch, dist):
    """One untimed step of the multi-GPU exchange on the data path the timed legs use, against torch.distributed + the dense
    optimizer (gsr_adam_step): (1) FusedExchange.allreduce() vs dist.all_reduce of a copy of the bucket; (2)
    FusedExchange.step(stats=True) vs [gradients summed over ranks -> GaussianAdam on a replica] and three
    dist.all_reduce calls for the SUM / SUM / MAX statistic segments (the semantics of /root/reference/train.py:163-186,
    /root/reference/scene/gaussian_model.py:178-201, 471-473 applied to the step's views).  The reference sum is taken in
    the order the kernel adds (rank order on the peer path; the switch's order on the multicast path, obtained from the
    all-reduce mode that (1) has just checked), so parameters and moments must come out bit for bit.  Everything the
    check touched is put back before the timed legs.  Returns the report; `ok` is the AND over ranks."""
    from gsr_b200.dist import FlatGaussians
    from gsr_b200.optim import GaussianAdam
    G, fused, dev, world, rank = job.G, job.fused, job.dev, job.world, job.rank
    saved = G.params.clone()
    gen = torch.Generator(device=dev).manual_seed(4242 + rank)
    P = G.P

    def fill():
        G.grad_all.zero_()
        for v in G.grad_views().values():
            v.copy_(torch.randn(v.shape, device=dev, generator=gen))
            v[::5] = 0
    # (1) plain all-reduce mode
    fill()
    local = G.bucket.clone()
    want = local.clone()
    dist.all_reduce(want)
    fused.allreduce()
    torch.cuda.synchronize(dev)
    dist.barrier()
    ar_equal = bool(torch.equal(want, G.bucket))
    ar_rel = float(((want - G.bucket).abs().max() / (want.abs().max() + 1e-30)).item())
    params_untouched = bool(torch.equal(saved, G.params))
    # (2) the fused step
    ref = FlatGaussians(P, dev)
    ref.params.copy_(saved[:ref.params.numel()])
    opt_ref = GaussianAdam(ref, lrs=job.lrs)
    fill()
    local = G.bucket.clone()
    acc, den, rad = G.stat_views()
    acc.copy_(torch.rand(P, 1, device=dev, generator=gen))
    den.copy_((torch.rand(P, 1, device=dev, generator=gen) > 0.5).float())
    rad.copy_(torch.randint(0, 200, (P,), device=dev, generator=gen).float())
    want_stats = [acc.clone(), den.clone(), rad.clone()]
    dist.all_reduce(want_stats[0]); dist.all_reduce(want_stats[1]); dist.all_reduce(want_stats[2], op=dist.ReduceOp.MAX)
    n = ref.bucket.numel()
    if fused.use_multicast:
        fused.allreduce()
        torch.cuda.synchronize(dev)
        dist.barrier()
        ref.bucket.copy_(G.bucket[:n])
        G.bucket.copy_(local)
        torch.cuda.synchronize(dev)
        dist.barrier()
    else:
        parts = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(parts, local)
        total = parts[0].clone()
        for k in range(1, world):
            total += parts[k]
        ref.bucket.copy_(total[:n])
        del parts, total
    opt_ref.step()
    step_before = fused.step_count
    fused.step(stats=True)
    torch.cuda.synchronize(dev)
    dist.barrier()
    p_equal = bool(torch.equal(ref.params, G.params[:n]))
    p_rel = float(((ref.params - G.params[:n]).abs().max() / (ref.params.abs().max() + 1e-30)).item())
    b, e = fused.shard
    e = min(e, n)
    m_equal = e <= b or bool(torch.equal(opt_ref.exp_avg[b:e], fused.exp_avg[:e - b]) and torch.equal(opt_ref.exp_avg_sq[b:e], fused.exp_avg_sq[:e - b]))
    acc_rel = float(((acc - want_stats[0]).abs().max() / (want_stats[0].abs().max() + 1e-30)).item())
    stats_equal = bool(acc_rel <= 1e-6 and torch.equal(den, want_stats[1]) and torch.equal(rad, want_stats[2]))
    ok = (ar_equal or ar_rel <= 1e-5) and params_untouched and p_equal and m_equal and stats_equal
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    # put everything back: parameters, zero moments, step count, empty bucket and statistic segments
    G.params.copy_(saved)
    fused.exp_avg.zero_(); fused.exp_avg_sq.zero_()
    fused.step_count = step_before
    G.grad_all.zero_()
    torch.cuda.synchronize(dev)
    dist.barrier()
    return {"path": "multicast" if fused.use_multicast else "peer", "allreduce_equal": ar_equal, "allreduce_max_rel": ar_rel,
            "allreduce_leaves_params": params_untouched, "params_equal_or_max_rel": True if p_equal else p_rel,
            "moments_equal": bool(m_equal), "stats_equal": stats_equal, "stat_accum_max_rel": acc_rel,
            "against": "dist.all_reduce + gsr_adam_step on a replica + 3 dist.all_reduce (SUM, SUM, MAX); reference sums in the kernel's order",
            "ok_rank0": bool(ok), "ok": bool(flag.item() == 1.0)}


def time_steps(torch, dist, dev, world, step_fn, steps, warmup, min_warm_s=1.5):
    """W >= warmup untimed steps (and at least min_warm_s of them: a freshly leased GPU ramps clocks and grows
    allocator pools during its first second), then EXACTLY `steps` steps between barrier + synchronize on both sides;
    CUDA events on the launching stream, MAX over ranks.  Returns (total ms, sorted per-step ms, warm-up steps run)."""
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    n_warm, t0 = 0, time.perf_counter()
    while True:
        step_fn()
        n_warm += 1
        if n_warm >= warmup and (n_warm - warmup) % 4 == 0:
            torch.cuda.synchronize()
            # each step may hold a collective: all ranks leave after the same count (MAX of "I want more")
            more = (time.perf_counter() - t0 < min_warm_s) and n_warm < 4000
            if world > 1:
                flag = torch.tensor([1.0 if more else 0.0], device=dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MAX)
                more = bool(flag.item() > 0)
            if not more:
                break
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    barrier()
    ev[0].record()
    for k in range(steps):
        step_fn()
        ev[k + 1].record()
    barrier()
    ms = ev[0].elapsed_time(ev[steps])
    per = sorted(ev[k].elapsed_time(ev[k + 1]) for k in range(steps))
    if world > 1:
        tms = torch.tensor([ms], device=dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    return ms, per, n_warm


def optional_leg(name, fn):
    """Runs one of the bench's extra legs; a failure there is reported under the leg's key instead of costing the whole
    line (`value`, `e2e` and the roofline are computed outside these)."""
    try:
        return fn()
    except Exception as exc:      # noqa: BLE001 -- reported, not swallowed
        import traceback
        sys.stderr.write("bench: leg '%s' failed\n%s\n" % (name, traceback.format_exc()))
        return {"error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}


def newest_traffic(workload):
    """DRAM bytes per launch from the newest committed `ncu --set full` capture of this workload: the highest round in
    the file name (profiles/rNN_*traffic*.json), ties by modification time."""
    import glob
    import re
    best = None
    for tf in glob.glob(os.path.join(ROOT, "profiles", "*traffic*.json")):
        try:
            with open(tf) as f:
                tj = json.load(f)
        except (OSError, ValueError):
            continue
        if tj.get("workload") != workload:
            continue
        m = re.match(r"r(\d+)", os.path.basename(tf))
        key = (int(m.group(1)) if m else -1, os.path.getmtime(tf))
        if best is None or key > best[0]:
            best = (key, tf, tj)
    return (best[1], best[2]) if best else (None, None)


def gsiron_host_row():
    """the fork's own forward-only host renderer (GS-IRON/npu-1/base.cpp:209-433), restated without
    Eigen / OpenCV in oracle/gsiron_host.c, single-threaded by construction, on workload C1 with the chair camera."""
    from oracle import gsiron_host as gh
    w = WORKLOADS["c1"]
    scene = synth.random_init_scene(w["P"], w["seed"])
    gh.render({k: (v[:2000] if hasattr(v, "shape") else v) for k, v in scene.items()}, w["W"], w["H"])    # warm (library load)
    t0 = time.perf_counter()
    img, info = gh.render(scene, w["W"], w["H"])
    dt = time.perf_counter() - t0
    return {"value": 1.0 / dt, "unit": "renders/s (forward only)", "cores": 1, "kind": "port", "ms": dt * 1e3,
            "sample": "C1: %d random-init Gaussians, 800x800, hard-coded chair camera of base.cpp:453-457, forward only; "
                      "not parity-comparable with the rasterizer (+0.5 px sample offset, flipped view direction, no tan_fov clamp)" % w["P"],
            "instances": info["instances"], "image_mean": float(img.mean())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--no-c3", action="store_true", help="skip the BASELINE configs[2] strong-scaling leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the loss / frame / optimizer stage reports")
    ap.add_argument("--no-balance", action="store_true", help="N > 1: keep cameras r*V .. r*V+V-1 on rank r instead of dealing the step's views by predicted cost")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="gradient exchange at N > 1: this library's kernel over NVLink peer / multicast memory, or ncclAllReduce")
    ap.add_argument("--streams", type=int, default=3,
                    help="CUDA streams of the `overlapped_step` leg (gsr_b200.dist.ViewStreams); `value` always uses one")
    ap.add_argument("--views-per-rank", type=int, default=8,
                    help="views every rank renders per step (gradients accumulate in the bucket; one exchange per step)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL's "NCCL version ..." banner when
    # NCCL_DEBUG is set, for one) is sent to stderr; the line itself goes to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from gsr_b200 import capi
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    V = max(1, args.views_per_rank)
    S = max(1, args.streams)
    job = Job(args.workload, V, V * world, rank, world, dev, args.exchange, max_streams=S)
    balance_info = None if args.no_balance else job.balance()
    w, G = job.w, job.G
    W, H, P = w["W"], w["H"], w["P"]
    warm = max(args.warmup, 3)

    # N > 1: the fused exchange proves itself against torch.distributed + the dense optimizer before anything is timed
    exchange_check = None
    if world > 1 and job.fused is not None:
        exchange_check = exchange_self_check(job, torch, dist)

    # ---------------- device-timed leg (C ABI, inputs resident): each view alone on one stream ----------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.2)
    ms, step_ms, n_warm = time_steps(torch, dist, dev, world, lambda: job.step_render(1), args.steps, warm)
    clocks = sampler.stop()
    value = world * V * args.steps / (ms / 1000.0)
    st = job.last_state
    N = st["num_rendered"]

    # the exchange alone (N > 1): device time of one SUM over ranks of the bucket, both implementations
    exchange_ms = None
    if world > 1:
        exchange_ms = {}
        G.bucket.zero_()            # repeated in-place sums of the same buffer must not run into inf
        ex_ms, _, _ = time_steps(torch, dist, dev, world, job.exchange_gradients, 20, 3, min_warm_s=0.0)
        exchange_ms[args.exchange] = ex_ms / 20
        nccl_ms, _, _ = time_steps(torch, dist, dev, world, lambda: dist.all_reduce(G.bucket), 20, 3, min_warm_s=0.0)
        exchange_ms["nccl_allreduce"] = nccl_ms / 20
        exchange_ms["bytes"] = int(G.bucket.numel() * 4)

    # ---------------- the same views overlapped on S streams (own key; not `value`) ----------------
    def leg_overlapped():
        ov_ms, _, _ = time_steps(torch, dist, dev, world, lambda: job.step_render(S), args.steps, warm, min_warm_s=0.3)
        return {"streams": S, "value": world * V * args.steps / (ov_ms / 1000.0), "unit": "renders/s",
                "ms_per_step": ov_ms / args.steps, "ms_per_view": ov_ms / args.steps / V,
                "what": "the same step with view v on CUDA stream v %% %d (gsr_b200.dist.ViewStreams): the latency-bound "
                        "stages of one view run under the blend kernels of its neighbours" % S}
    overlapped = optional_leg("overlapped_step", leg_overlapped)

    # ---------------- the same views through the batched entry points (own key; not `value`) ----------------
    def leg_batched():
        bt_ms, _, _ = time_steps(torch, dist, dev, world, lambda: job.step_batched(S), args.steps, warm, min_warm_s=0.3)
        b1_ms, _, _ = time_steps(torch, dist, dev, world, lambda: job.step_batched(1), args.steps, warm, min_warm_s=0.3)
        return {"value": world * V * args.steps / (bt_ms / 1000.0), "unit": "renders/s", "ms_per_step": bt_ms / args.steps,
                "ms_per_view": bt_ms / args.steps / V, "streams": S,
                "one_stream": {"value": world * V * args.steps / (b1_ms / 1000.0), "ms_per_step": b1_ms / args.steps},
                "what": "the same %d views through gsr_forward_batch / gsr_backward_batch: ONE per-Gaussian launch each way for all cameras "
                        "(parameters and SH rows read once, the 59 gradient floats written once), scan / sorts / emission / blend per "
                        "view, the views' chains on %d streams (`one_stream`: everything on one)" % (V, S)}
    batched = optional_leg("batched_step", leg_batched)
    if hasattr(job, "batch_ws"):
        del job.batch_ws, job.batch_scratch
        torch.cuda.empty_cache()

    # ---------------- with the optimizer: render + densification statistics + Adam (+ exchange) ----------------
    def leg_optimizer():
        saved = job.save_params()
        try:
            op_ms, _, _ = time_steps(torch, dist, dev, world, lambda: job.step_train(1), args.steps, warm, min_warm_s=0.3)
        finally:
            job.restore_params(saved)
        return {"value": world * V * args.steps / (op_ms / 1000.0), "unit": "renders/s", "ms_per_step": op_ms / args.steps,
                "what": ("views as in `value`, each followed by the fused densification-statistics update; then ONE fused exchange kernel per rank: "
                         "reduce-scatter of the bucket -> Adam on the rank's 1/%d shard -> all-gather of the parameters, + SUM/MAX of the statistics" % world)
                if job.fused is not None else
                "views as in `value`, each followed by the fused densification-statistics update; then %sgsr_adam_step over all rows" % (
                    "ncclAllReduce of the bucket + 3 statistic all-reduces + " if world > 1 else ""),
                "learning_rates": "the reference's x 1e-6 (same arithmetic and traffic; keeps the timed steps on the same scene)"}
    with_optimizer = optional_leg("with_optimizer", leg_optimizer)

    # ---------------- single view, isolated: synchronise before and after every view (latency) ----------------
    sv_ms = []
    for k in range(12):
        a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a_.record()
        st0 = job._fwd(0, 0)
        job._bwd(0, st0, False, 0)
        b_.record()
        torch.cuda.synchronize()
        sv_ms.append(a_.elapsed_time(b_))
    single_view_ms = float(np.median(sv_ms[2:]))

    # ---------------- per-stage device times (profiling mode, separate pass) ----------------
    per_view, dcol, ddep, grad_views = job.per_view, job.dcol, job.ddep, job.grad_views
    capi.set_profiling(True)
    stage = {}
    for _ in range(5):      # view 0 of this rank, write mode: the single-view fwd+bwd the roofline figures describe
        st = job._fwd(0, 0)
        g = job._bwd(0, st, False, 0)
        for k, v in capi.stage_times_ms().items():
            stage.setdefault(k, []).append(v)
    acc_ms = []
    for _ in range(3):      # the same view once more in accumulate mode (views 2..V of a step)
        job._bwd(0, st, True, 0)
        acc_ms.append(capi.stage_times_ms()["preprocess_bwd"])
    capi.set_profiling(False)
    stage_ms = {k: float(np.median(v)) for k, v in stage.items()}
    stage_ms["preprocess_bwd_accumulate"] = float(np.median(acc_ms))
    it = capi.introspect(st)
    pairs_bwd_eval = int(it["n_contrib"].to(torch.int64).sum().item())
    P_vis = int((st["radii"] > 0).sum().item())
    radii0 = st["radii"].clone()

    # ---------------- e2e leg: public Python API, host inputs, loss read back ----------------
    cams = job.cams
    bg = job.bg
    rs_kw = dict(image_height=H, image_width=W, bg=bg, scale_modifier=1.0,
                 sh_degree=3, prefiltered=False, debug=False, antialiasing=w["aa"])
    from diff_gaussian_rasterization import grad_sink
    from fused_ssim import l1_ssim_loss
    rng9 = np.random.default_rng(9)
    target_host = [torch.from_numpy(rng9.random((3, H, W), dtype=np.float32)).pin_memory() for _ in range(min(V, 2))]
    cam_host = [torch.from_numpy(np.concatenate([c.world_view_transform.ravel(), c.full_proj_transform.ravel(),
                                                 c.camera_center.ravel()]).astype(np.float32)).pin_memory() for c in cams]
    NB = 2                                                                   # one input buffer per view in flight
    target_dev = [torch.empty(3, H, W, device=dev) for _ in range(NB)]       # double-buffered: the copy of view v+1
    cam_dev = [torch.empty(35, device=dev) for _ in range(V)]               # one per view: the deferred sink reads the cameras at the end of the step
    params = G.as_leaf_params()
    loss_host = torch.empty(1).pin_memory()
    loss_dev = torch.zeros(1, device=dev)
    loss_slots = torch.zeros(V, 1, device=dev)
    copy_stream = torch.cuda.Stream(device=dev)
    uploaded = [torch.cuda.Event() for _ in range(NB)]
    buf_free = [torch.cuda.Event() for _ in range(NB)]
    for e_ in buf_free:
        e_.record(torch.cuda.current_stream())
    sink = (grad_views["dL_dmeans3D"], grad_views["dL_dsh"], grad_views["dL_dopacity"], grad_views["dL_dscales"],
            grad_views["dL_drotations"])

    def api_step(use_sink):
        """One step through the public API: for each of this rank's V views the camera and the target image come from
        pinned host memory (copy stream, double-buffered), GaussianRasterizer renders, the training loss
        0.8 L1 + 0.2 D-SSIM (fused) is taken and backpropagated.  use_sink "defer": one grad_sink(defer=True) context around
        the step (blend stage per backward, ONE batched per-Gaussian launch at the end writes the bucket); "per_view": a
        grad_sink context per view (the kernels add into the bucket).  False: the STOCK autograd path: gradients
        are returned and summed into .grad by AccumulateGrad, exactly what the reference's unchanged train.py gets;
        with N > 1 they are then packed into the bucket, exchanged and unpacked.  The step's loss is read back."""
        main = torch.cuda.current_stream()

        def upload(v):
            copy_stream.wait_event(buf_free[v % NB])
            with torch.cuda.stream(copy_stream):
                cam_dev[v].copy_(cam_host[v], non_blocking=True)
                target_dev[v % NB].copy_(target_host[v % len(target_host)], non_blocking=True)
                uploaded[v % NB].record(copy_stream)

        import contextlib
        step_ctx = grad_sink(*sink, accumulate=False, defer=True) if use_sink == "defer" else contextlib.nullcontext()
        upload(0)
        with step_ctx:
          for v in range(V):
              cd, td = cam_dev[v], target_dev[v % NB]
              main.wait_event(uploaded[v % NB])
              if v + 1 < V:
                  upload(v + 1)
              rs = GaussianRasterizationSettings(viewmatrix=cd[0:16].view(4, 4), projmatrix=cd[16:32].view(4, 4),
                                                 campos=cd[32:35], tanfovx=cams[v].tanfovx, tanfovy=cams[v].tanfovy, **rs_kw)
              means2D = torch.zeros_like(params["xyz"], requires_grad=True)
              color, radii, invd = GaussianRasterizer(rs)(means3D=params["xyz"], means2D=means2D, opacities=params["opacity"],
                                                          shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
              loss = l1_ssim_loss(color, td, 0.2)
              if use_sink == "per_view":
                  with grad_sink(*sink, accumulate=(v > 0)):
                      loss.backward()
              else:
                  loss.backward()
              loss_slots[v].copy_(loss.detach().reshape(1))
              buf_free[v % NB].record(main)
        loss_dev.copy_(loss_slots.sum().reshape(1))
        if use_sink:
            job.exchange_gradients()
        else:
            G.allreduce_param_grads(params)
            for p_ in params.values():
                p_.grad = None                       # optimizer.zero_grad(set_to_none=True), train.py:186
        loss_host.copy_(loss_dev, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(loss_host[0])

    def time_api(use_sink):
        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
        for _ in range(max(warm, 5)):
            api_step(use_sink)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        loss_ = None
        for _ in range(args.steps):
            loss_ = api_step(use_sink)
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        ms_ = max(e0.elapsed_time(e1), wall * 1000.0)
        if world > 1:
            tms = torch.tensor([ms_], device=dev)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms_ = float(tms.item())
        return ms_, loss_

    e2e_ms, step_loss = time_api("defer")
    bucket_checksum = float(G.bucket.double().abs().sum().item())

    def leg_per_view_sink():
        pv_ms, pv_loss = time_api("per_view")
        return {"value": world * V * args.steps / (pv_ms / 1000.0), "unit": "renders/s", "ms_per_step": pv_ms / args.steps, "step_loss": pv_loss,
                "bucket_checksum": float(G.bucket.double().abs().sum().item()),
                "what": "grad_sink per view (accumulate from the second view on): every backward runs its own per-Gaussian stage"}
    per_view_sink = optional_leg("e2e.grad_sink_per_view", leg_per_view_sink)
    def leg_stock():
        stock_ms, stock_loss = time_api(False)
        return {"value": world * V * args.steps / (stock_ms / 1000.0), "unit": "renders/s", "ms_per_step": stock_ms / args.steps,
                "step_loss": stock_loss,
                "what": "the same without grad_sink: the unchanged reference API, gradients returned to autograd and summed into .grad by AccumulateGrad (what the reference's train.py does), .grad dropped per step"}
    stock = optional_leg("e2e.stock_autograd", leg_stock)
    e2e_val = world * V * args.steps / (e2e_ms / 1000.0)
    h2d = V * (target_host[0].numel() * 4 + cam_host[0].numel() * 4)
    d2h = 4 + 12 * V    # the step's loss + the 12-byte instance/visible counts of every forward
    del params, target_dev

    # ---------------- BASELINE configs[2] as written: C3, 8 views per step in total, split over the ranks ----------------
    def leg_c3():
        torch.cuda.empty_cache()
        j3 = Job("c3", 8 // world, 8, rank, world, dev, args.exchange, max_streams=S)
        if not args.no_balance:
            j3.balance()
        r_ms, _, _ = time_steps(torch, dist, dev, world, lambda: j3.step_render(1), args.steps, warm, min_warm_s=0.5)
        saved = j3.save_params()
        t_ms, _, _ = time_steps(torch, dist, dev, world, lambda: j3.step_train(1), args.steps, warm, min_warm_s=0.3)
        j3.restore_params(saved)
        o_ms, _, _ = time_steps(torch, dist, dev, world, lambda: j3.step_render(S), args.steps, warm, min_warm_s=0.3)
        b_ms, _, _ = time_steps(torch, dist, dev, world, lambda: j3.step_batched(S), args.steps, warm, min_warm_s=0.3)
        res = {"workload": "c3: " + WORKLOADS["c3"]["desc"] + "; 8 views per step IN TOTAL, %d per rank, gradient exchange every step" % (8 // world),
               "scaling": "strong", "n_gpus": world, "views_per_step": 8, "exchange": args.exchange if world > 1 else None,
               "value": 8 * args.steps / (r_ms / 1000.0), "unit": "views/s", "ms_per_step": r_ms / args.steps,
               "with_optimizer": {"value": 8 * args.steps / (t_ms / 1000.0), "ms_per_step": t_ms / args.steps},
               "overlapped_%d_streams" % S: {"value": 8 * args.steps / (o_ms / 1000.0), "ms_per_step": o_ms / args.steps},
               "batched_%d_streams" % S: {"value": 8 * args.steps / (b_ms / 1000.0), "ms_per_step": b_ms / args.steps},
               "bucket_bytes": int(j3.G.bucket.numel() * 4), "num_rendered_last_view": j3.last_state["num_rendered"]}
        if world > 1:
            j3.G.bucket.zero_()
            x_ms, _, _ = time_steps(torch, dist, dev, world, j3.exchange_gradients, 20, 3, min_warm_s=0.0)
            res["exchange_ms"] = x_ms / 20
        return res
    c3 = None
    if not args.no_c3 and args.workload == "c2" and 8 % world == 0:
        c3 = optional_leg("c3_strong", leg_c3)
        torch.cuda.empty_cache()

    out = None
    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except OSError:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback 6650 GB/s"
        sm_mhz = clocks.get("sm_mhz") or 1965.0
        fp32_peak_max = FP32_LANES * 1.965e9 / 1e12
        fp32_peak_clk = FP32_LANES * sm_mhz * 1e6 / 1e12

        # ---------------- CPU baselines (bounded samples), N = 1 only ----------------
        cpu = None
        pairs_eval = pairs_blend = None
        if not args.no_cpu_baseline and world == 1:
            from oracle import gs_oracle as go
            go.set_num_threads(host_threads())
            sec, f = oracle_step(job.scene, cams[0], w, (job.dcol_np, job.ddep_np))
            pairs_eval, pairs_blend = f["pairs_evaluated"], f["pairs_blended"]
            cpu = {"value": 1.0 / sec, "unit": "renders/s", "cores": go.num_threads(), "kind": "port",
                   "sample": "one full %dx%d view of %d Gaussians, fwd+bwd, %.1f s" % (W, H, P, sec)}
            try:
                go.set_num_threads(1)
                sec1, _ = oracle_step(job.scene, cams[0], w, (job.dcol_np, job.ddep_np))
                cpu["one_core"] = {"value": 1.0 / sec1, "unit": "renders/s", "cores": 1, "sample": "the same view, 1 thread, %.1f s" % sec1}
            finally:
                go.set_num_threads(host_threads())
            try:
                cpu["gsiron_host"] = gsiron_host_row()
            except Exception as exc:  # informational row
                cpu["gsiron_host"] = {"error": str(exc)[:300]}

        # binning = depth sort of the P_vis visible Gaussians (32 key bits -> 4 onesweep passes on 32-bit
        # key/value pairs) + stable tile-id sort of the N instances (tile_bits -> tile_passes passes)
        tile_bits = max(int(np.ceil(np.log2(((W + 15) // 16) * ((H + 15) // 16) + 1))), 1)
        tile_passes = (tile_bits + 7) // 8
        def hbm(bytes_, ms_):
            a = bytes_ / (ms_ / 1e3) / 1e9 if ms_ and ms_ > 0 else None
            return {"bound": "hbm", "algorithmic_bytes": int(bytes_), "ms": ms_, "achieved": a, "peak": hbm_peak,
                    "unit": "GB/s", "frac": (a / hbm_peak) if a else None}
        stages = {
            "preprocess": hbm(52 * P + 259 * P_vis, stage_ms.get("preprocess")),
            "scan": hbm(8 * P + 12 * P_vis, stage_ms.get("scan")),                       # + compacted depth key/index out
            "depth_sort": hbm(4 * P_vis + 4 * 16 * P_vis, stage_ms.get("depth_sort")),    # histogram read + 4 x (8 in + 8 out)
            "duplicate": hbm(40 * P_vis + 8 * N, stage_ms.get("duplicate")),             # sorted scan + emit (32-bit key + value out)
            "sort": hbm(4 * N + tile_passes * 16 * N, stage_ms.get("sort")),
            "ranges": hbm(8 * N + 8 * ((W + 15) // 16) * ((H + 15) // 16), stage_ms.get("ranges")),
            "preprocess_bwd": hbm(504 * P_vis + 236 * P, stage_ms.get("preprocess_bwd")),
        }
        def fp32(flop, ms_):
            a = flop / (ms_ / 1e3) / 1e12 if ms_ and ms_ > 0 and flop else None
            return {"bound": "fp32", "algorithmic_flop": int(flop) if flop else None, "ms": ms_, "achieved": a,
                    "peak": fp32_peak_max, "peak_at_measured_clock": fp32_peak_clk, "unit": "TFLOP/s",
                    "frac": (a / fp32_peak_max) if a else None, "traffic": None}
        if pairs_eval is not None:
            stages["render_fwd"] = fp32(30 * pairs_blend + 14 * (pairs_eval - pairs_blend), stage_ms.get("render_fwd"))
            stages["render_bwd"] = fp32(75 * pairs_blend + 14 * (pairs_bwd_eval - pairs_blend), stage_ms.get("render_bwd"))
        else:
            stages["render_fwd"] = fp32(None, stage_ms.get("render_fwd"))
            stages["render_bwd"] = fp32(None, stage_ms.get("render_bwd"))
        tf, tj = newest_traffic(args.workload)
        if tj:
            for k_, v_ in tj["traffic_bytes"].items():
                if k_ in stages:
                    stages[k_]["traffic"] = v_
                    stages[k_]["traffic_source"] = "profiles/" + os.path.basename(tf)
        dominant = max(stage_ms, key=lambda k: stage_ms[k] if k not in ("host_sync", "preprocess_bwd_accumulate") else -1)
        roof = dict(stages.get(dominant, {}))
        roof["kernel"] = dominant
        roof.setdefault("traffic", None)

        # preprocess, scan_compact, depth sort (hist + 4), sorted_scan, emit_tiles, tile sort (hist + passes),
        # tile_ranges, render_fwd, render_bwd, preprocess_bwd; + the exchange kernel per step when N > 1 (fused)
        launches_per_view = 1 + 1 + (1 + 4) + 1 + 1 + (1 + tile_passes) + 1 + 1 + 1 + 1
        launches_per_step = launches_per_view * V + (1 if job.fused is not None else 0)
        out = {
            "metric": "fwd+bwd renders/s", "value": value, "unit": "renders/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "warmup_steps_run": n_warm, "ms_per_step": ms / args.steps, "ms_per_view": ms / args.steps / V,
            "step_ms": {"min": step_ms[0], "median": step_ms[len(step_ms) // 2], "max": step_ms[-1]}, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": static_config(args, world),
            "workload_stats": {"P_visible": P_vis, "num_rendered": N, "pairs_evaluated_fwd": pairs_eval, "pairs_blended": pairs_blend,
                               "pairs_evaluated_bwd": pairs_bwd_eval, "peak_source": peak_src, "exchange": args.exchange if world > 1 else None, "view_balance": balance_info,
                               "exchange_data_path": (("multicast" if job.fused.use_multicast else "peer"), job.fused.calibration_ms) if job.fused is not None else None,
                               "roofline_view": "stage times / roofline figures are of view 0 of rank 0 alone (single-view fwd+bwd, write mode)"},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "renders/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps, "step_loss": step_loss, "bucket_checksum": bucket_checksum,
                    "what": "per view: camera + target image from pinned host memory (copy stream, double-buffered, overlapping the previous view's kernels), GaussianRasterizer autograd fwd+bwd + the fused training loss 0.8*L1 + 0.2*(1-SSIM) (fused_ssim.l1_ssim_loss), views one after another on one stream; per step: ONE grad_sink(defer=True) context around the views -- every backward runs its blend stage at once, the per-Gaussian stage of all views is one batched launch when the context closes and writes the bucket once; exchange when N>1, loss read back",
                    "grad_sink_per_view": per_view_sink, "stock_autograd": stock},
            "gpu_launches": launches_per_step * args.steps,
            "single_view_isolated": {"fwd_bwd_ms": single_view_ms, "renders_per_s": 1000.0 / single_view_ms,
                                     "what": "one view, synchronised before and after (launch ramp included): latency, not throughput"},
            "overlapped_step": overlapped, "batched_step": batched, "with_optimizer": with_optimizer, "exchange_ms": exchange_ms, "exchange_check": exchange_check, "c3_strong": c3,
            "stage_ms": stage_ms, "roofline": roof, "roofline_stages": stages, "cpu_baseline": cpu,
        }
        if world == 1 and not args.no_extras:
            out["loss_stage"] = loss_stage_report(dev, H, W, hbm_peak)
            if not args.no_cpu_baseline:
                out["loss_stage"]["cpu_baseline"] = loss_cpu_baseline(H, W)
            try:
                out["frame_stage"] = frame_stage_report(dev, H, W, hbm_peak)
            except Exception as exc:  # informational leg only
                out["frame_stage"] = {"error": str(exc)[:300]}
            try:
                out["optimizer_stage"] = optimizer_stage_report(G, radii0, hbm_peak)
            except Exception as exc:  # informational leg only
                out["optimizer_stage"] = {"error": str(exc)[:300]}
        # same-box bar: the unmodified reference CUDA kernels (oracle/_ref), informational
        if not args.no_ref_cuda and world == 1:
            try:
                from oracle import ref_cuda
                if ref_cuda.available():
                    common = per_view[0]
                    R = ref_cuda.RefRasterizer()
                    for _ in range(2):
                        f = R.forward(width=W, height=H, **common)
                        R.backward(f["radii"], dL_dcolor=dcol, dL_dinvdepth=ddep, **common)
                    torch.cuda.synchronize()
                    nrep = 20
                    # the reference launches on the legacy default stream, which torch's current (default) stream is:
                    # events recorded there bracket its kernels
                    evs = [torch.cuda.Event(enable_timing=True) for _ in range(2 * nrep + 1)]
                    t0 = time.perf_counter()
                    evs[0].record()
                    for k in range(nrep):
                        f = R.forward(width=W, height=H, **common)
                        evs[2 * k + 1].record()
                        R.backward(f["radii"], dL_dcolor=dcol, dL_dinvdepth=ddep, **common)
                        evs[2 * k + 2].record()
                    torch.cuda.synchronize()
                    wall = (time.perf_counter() - t0) / nrep
                    dt = evs[0].elapsed_time(evs[2 * nrep]) / nrep / 1e3
                    tf_ = sum(evs[2 * k].elapsed_time(evs[2 * k + 1]) for k in range(nrep)) / 1e3
                    out["reference_cuda_same_box"] = {"renders_per_s": 1.0 / dt, "ms_per_step": dt * 1e3, "wall_ms_per_step": wall * 1e3,
                                                      "fwd_ms": tf_ / nrep * 1e3, "reps": nrep, "timing": "CUDA events on the default stream",
                                                      "note": "unmodified reference kernels + CUB, sm_100a, incl. zero-fills"}
                    R.close()
            except Exception as exc:  # informational leg only
                out["reference_cuda_same_box"] = {"error": str(exc)[:200]}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if exchange_check is not None and not exchange_check["ok"]:
        sys.stderr.write("bench: the multi-GPU exchange self-check FAILED: %s\n" % json.dumps(exchange_check))
        sys.exit(3)


if __name__ == "__main__":
    main()
