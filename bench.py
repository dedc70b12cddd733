#!/usr/bin/env python
"""bench.py -- fwd+bwd renders/s of the differentiable Gaussian-splatting rasterizer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c1|c3|c4]

Workload at N=1 is BASELINE.json configs[1] ("c2"): 3M SH-3 Gaussians, one 1920x1080 view, forward
+ backward.  One "step" = one view's forward + backward (+ the NCCL gradient all-reduce when N>1:
every rank renders its own camera of the same replicated scene, weak scaling, views/s aggregate).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for what each key means.

  value        device-timed (CUDA events around the K steps) through the C ABI, inputs resident in HBM
  e2e          the same step through the public Python API (diff_gaussian_rasterization autograd
               Function) with the per-step inputs (camera + target image) copied from pinned host
               memory and the loss read back to the host inside the timed region
  roofline     the dominant kernel (backward blend, FP32-pipe bound) + roofline_stages for every stage
  cpu_baseline the CPU oracle (a port of the reference algorithm) on the same workload, host threads
  --impl reference : the reference algorithm's CPU implementation (the oracle port) timed alone.
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


def make_workload(name, cam_index=0):
    w = WORKLOADS[name]
    if w["kind"] == "random_init":
        scene = synth.random_init_scene(w["P"], w["seed"])
        import cases
        cam = cases._lego_cam(1, w["W"], w["H"])
    else:
        scene = synth.trained_like_scene(w["P"], w["seed"], s0=w["s0"], extent=w.get("extent", 1.0))
        cam = synth.orbit_cameras(8, w["W"], w["H"], w["seed"])[cam_index % 8]
    return w, scene, cam


def make_workload_camera(name, cam_index):
    w = WORKLOADS[name]
    if w["kind"] == "random_init":
        import cases
        return cases._lego_cam(1 + cam_index % 8, w["W"], w["H"])
    return synth.orbit_cameras(8, w["W"], w["H"], w["seed"])[cam_index % 8]


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


def run_reference(args):
    """--impl reference: the reference algorithm's CPU implementation (oracle port), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import gs_oracle as go
    w, scene, cam = make_workload(args.workload)
    grads = synth.upstream_grads(w["H"], w["W"], 5)
    for _ in range(args.warmup):
        oracle_step(scene, cam, w, grads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle_step(scene, cam, w, grads)
    dt = time.perf_counter() - t0
    val = args.steps / dt
    cores = go.num_threads()
    sample = "full workload: one %dx%d view of %d Gaussians, fwd+bwd per step" % (w["W"], w["H"], w["P"])
    print(json.dumps({
        "impl": "reference", "metric": "fwd+bwd renders/s", "value": val, "unit": "renders/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload + ": " + w["desc"]},
        "cpu_baseline": {"value": val, "unit": "renders/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "renders/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--streams", type=int, default=3,
                    help="CUDA streams the views of a step are spread over (gsr_b200.dist.ViewStreams)")
    ap.add_argument("--views-per-rank", type=int, default=8,
                    help="views every rank renders per step (gradients accumulate in the bucket; one all-reduce per step)")
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
    # this rank's slice of the step's camera batch: V views, global view index rank*V + v
    w, scene, cam = make_workload(args.workload, cam_index=rank * V)
    cams = [cam] + [make_workload_camera(args.workload, rank * V + v) for v in range(1, V)]
    W, H, P = w["W"], w["H"], w["P"]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    # replicated Gaussians, resident in HBM; parameters live in one flat [P,59] gradient-bucket layout
    from gsr_b200.dist import FlatGaussians
    G = FlatGaussians.from_numpy(scene, dev)
    bg = torch.zeros(3, device=dev)
    dcol_np, ddep_np = synth.upstream_grads(H, W, 5)
    dcol, ddep = t(dcol_np), t(ddep_np)
    shared = dict(means3D=G.xyz, opacities=G.opacity, background=bg, shs=G.shs, scales=G.scales, rotations=G.rotations,
                  sh_degree=3, antialiasing=w["aa"])
    per_view = [dict(viewmatrix=t(c.world_view_transform), projmatrix=t(c.full_proj_transform), cam_pos=t(c.camera_center),
                     tan_fovx=c.tanfovx, tan_fovy=c.tanfovy, **shared) for c in cams]
    common = per_view[0]

    grad_views = G.grad_views()     # the backward kernels write straight into the all-reduce bucket

    from gsr_b200.dist import ViewStreams
    view_streams = ViewStreams(dev, args.streams)
    # one growth-only scratch set per stream (a view's backward follows its forward on the same stream, so one live
    # state per stream): the binning buffer's size changes with every view, and per-call sizing lets the caching
    # allocator split / merge / occasionally cudaMalloc inside the timed region
    spaces = [capi.Workspace(dev) for _ in range(max(1, args.streams))]

    def abi_step():
        # one step = this rank's V views: the first backward overwrites the bucket, the others add into it
        return view_streams.run(
            V, lambda v: capi.forward(width=W, height=H, workspace=spaces[v % len(spaces)], **per_view[v]),
            lambda v, st, acc: capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=grad_views, accumulate=acc, **per_view[v]))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-timed leg (C ABI, inputs resident) ----------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.2)
    # warm-up: at least W (>= 3) steps AND at least ~1.5 s of work -- on a freshly leased box the first
    # second after context creation (clock/power-state ramp, allocator growth, lazy module loads) ran
    # up to 2x slower than steady state, which a 3-step warm-up of a 3 ms step does not cover
    n_warm, t_w0 = 0, time.perf_counter()
    while True:
        st, g = abi_step()
        if world > 1:
            G.allreduce_from(g)
        n_warm += 1
        if n_warm % 8 == 0 or n_warm == max(args.warmup, 3):
            torch.cuda.synchronize()
            # every rank must leave the loop after the same number of steps (each step holds a collective):
            # the decision is rank-agreed through a MAX all-reduce of "I still want to warm up"
            more = n_warm < max(args.warmup, 3) or (time.perf_counter() - t_w0 < 1.5 and n_warm < 2000)
            if world > 1:
                flag = torch.tensor([1.0 if more else 0.0], device=dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MAX)
                more = bool(flag.item() > 0)
            if not more:
                break
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    step_ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    barrier()
    e0.record()
    step_ev[0].record()
    for k in range(args.steps):
        st, g = abi_step()
        if world > 1:
            G.allreduce_from(g)
        step_ev[k + 1].record()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    step_ms = sorted(step_ev[k].elapsed_time(step_ev[k + 1]) for k in range(args.steps))
    clocks = sampler.stop()
    if world > 1:
        tms = torch.tensor([ms], device=dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    value = world * V * args.steps / (ms / 1000.0)
    N = st["num_rendered"]

    # ---------------- single-view latency: view 0 alone, one stream, write mode (BASELINE configs[1]) ----------------
    sv_ms = []
    for _ in range(12):
        a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a_.record()
        st = capi.forward(width=W, height=H, **per_view[0])
        capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=grad_views, **per_view[0])
        b_.record()
        torch.cuda.synchronize()
        sv_ms.append(a_.elapsed_time(b_))
    single_view_ms = float(np.median(sv_ms[2:]))

    # ---------------- per-stage device times (profiling mode, separate pass) ----------------
    capi.set_profiling(True)
    stage = {}
    for _ in range(3):      # view 0 of this rank, write mode: the single-view fwd+bwd the roofline figures describe
        st = capi.forward(width=W, height=H, **per_view[0])
        fwd_t = capi.stage_times_ms()
        g = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=grad_views, **per_view[0])
        for k, v in capi.stage_times_ms().items():
            stage.setdefault(k, []).append(v)
    acc_ms = []
    for _ in range(3):      # the same view once more in accumulate mode (views 2..V of a step)
        capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, out=grad_views, accumulate=True, **per_view[0])
        acc_ms.append(capi.stage_times_ms()["preprocess_bwd"])
    capi.set_profiling(False)
    stage_ms = {k: float(np.median(v)) for k, v in stage.items()}
    stage_ms["preprocess_bwd_accumulate"] = float(np.median(acc_ms))
    it = capi.introspect(st)
    pairs_bwd_eval = int(it["n_contrib"].to(torch.int64).sum().item())
    P_vis = int((st["radii"] > 0).sum().item())

    # ---------------- e2e leg: public Python API, host inputs, loss read back ----------------
    rs_kw = dict(image_height=H, image_width=W, bg=bg, scale_modifier=1.0,
                 sh_degree=3, prefiltered=False, debug=False, antialiasing=w["aa"])
    from diff_gaussian_rasterization import grad_sink
    from fused_ssim import l1_ssim_loss
    rng9 = np.random.default_rng(9)
    target_host = [torch.from_numpy(rng9.random((3, H, W), dtype=np.float32)).pin_memory() for _ in range(min(V, 2))]
    cam_host = [torch.from_numpy(np.concatenate([c.world_view_transform.ravel(), c.full_proj_transform.ravel(),
                                                 c.camera_center.ravel()]).astype(np.float32)).pin_memory() for c in cams]
    NB = max(2, args.streams)                                                # one input buffer per view in flight
    target_dev = [torch.empty(3, H, W, device=dev) for _ in range(NB)]       # double-buffered: the copy of view v+1
    cam_dev = [torch.empty(35, device=dev) for _ in range(NB)]              # may run under view v's kernels
    params = G.as_leaf_params()
    loss_host = torch.empty(1).pin_memory()
    loss_dev = torch.zeros(1, device=dev)
    loss_slots = torch.zeros(V, 1, device=dev)              # one slot per view: views on different streams never share a word
    copy_stream = torch.cuda.Stream(device=dev)
    uploaded = [torch.cuda.Event() for _ in range(NB)]
    buf_free = [torch.cuda.Event() for _ in range(NB)]
    for e_ in buf_free:
        e_.record(torch.cuda.current_stream())
    sink = (grad_views["dL_dmeans3D"], grad_views["dL_dsh"], grad_views["dL_dopacity"], grad_views["dL_dscales"],
            grad_views["dL_drotations"])

    def api_step():
        """One training-step's worth of rendering through the public API: for each of this rank's V views the camera
        and the target image come from pinned host memory, GaussianRasterizer renders, the training loss (0.8 L1 + 0.2 D-SSIM, fused) is taken and
        backpropagated; parameter gradients of the V views add up in the flat bucket (grad_sink), which is
        all-reduced over ranks; the step's loss is read back to the host."""
        main = torch.cuda.current_stream()

        def upload(v):      # H2D of view v's camera + target on the copy stream, into buffer v % NB
            copy_stream.wait_event(buf_free[v % NB])          # the view that last used this buffer is done with it
            with torch.cuda.stream(copy_stream):
                cam_dev[v % NB].copy_(cam_host[v], non_blocking=True)
                target_dev[v % NB].copy_(target_host[v % len(target_host)], non_blocking=True)
                uploaded[v % NB].record(copy_stream)

        upload(0)
        streams = [main] + view_streams.side                 # view v runs on stream v % S (gsr_b200.dist.ViewStreams)
        if len(streams) > 1:
            start = torch.cuda.Event()
            start.record(main)
            for s_ in streams[1:]:
                s_.wait_event(start)
        first_done = None
        for v in range(V):
            cd, td = cam_dev[v % NB], target_dev[v % NB]
            sv = streams[v % len(streams)]
            sv.wait_event(uploaded[v % NB])
            if v + 1 < V:
                upload(v + 1)                                 # travels while this view renders
            with torch.cuda.stream(sv):
                rs = GaussianRasterizationSettings(viewmatrix=cd[0:16].view(4, 4), projmatrix=cd[16:32].view(4, 4),
                                                   campos=cd[32:35], tanfovx=cams[v].tanfovx, tanfovy=cams[v].tanfovy, **rs_kw)
                means2D = torch.zeros_like(params["xyz"], requires_grad=True)
                color, radii, invd = GaussianRasterizer(rs)(means3D=params["xyz"], means2D=means2D, opacities=params["opacity"],
                                                            shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
                loss = l1_ssim_loss(color, td, 0.2)           # train.py:118-126, two fused launches
                if v > 0 and sv is not main and first_done is not None:
                    sv.wait_event(first_done)                 # view 0's backward overwrites the bucket: add only after it
                with grad_sink(*sink, accumulate=(v > 0)):
                    loss.backward()
                if v == 0 and len(streams) > 1:
                    first_done = torch.cuda.Event()
                    first_done.record(sv)
                loss_slots[v].copy_(loss.detach().reshape(1))
                buf_free[v % NB].record(sv)
        for s_ in streams[1:]:
            done = torch.cuda.Event()
            done.record(s_)
            main.wait_event(done)
        loss_dev.copy_(loss_slots.sum().reshape(1))
        if world > 1:
            dist.all_reduce(G.bucket, op=dist.ReduceOp.SUM)
        loss_host.copy_(loss_dev, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(loss_host[0])

    for _ in range(max(args.warmup, 5)):                    # the first steps of this leg grow the autograd path's buffers
        api_step()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    step_loss = None
    for _ in range(args.steps):
        step_loss = api_step()
    e1.record()
    barrier()
    wall = time.perf_counter() - t0
    e2e_ms = max(e0.elapsed_time(e1), wall * 1000.0)
    if world > 1:
        tms = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        e2e_ms = float(tms.item())
    e2e_val = world * V * args.steps / (e2e_ms / 1000.0)
    h2d = V * (target_host[0].numel() * 4 + cam_host[0].numel() * 4)
    d2h = 4 + 12 * V    # the step's loss + the 12-byte instance/visible counts of every forward

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

        # ---------------- CPU baseline (oracle port), bounded sample = one full view ----------------
        cpu = None
        pairs_eval = pairs_blend = None
        if not args.no_cpu_baseline and world == 1:
            from oracle import gs_oracle as go
            sec, f = oracle_step(scene, cam, w, (dcol_np, ddep_np))
            pairs_eval, pairs_blend = f["pairs_evaluated"], f["pairs_blended"]
            cpu = {"value": 1.0 / sec, "unit": "renders/s", "cores": go.num_threads(), "kind": "port",
                   "sample": "one full %dx%d view of %d Gaussians, fwd+bwd, %.1f s" % (W, H, P, sec)}

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
            "preprocess_bwd": hbm(504 * P_vis + 304 * (P - P_vis) * 0 + 236 * P, stage_ms.get("preprocess_bwd")),
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
        # measured DRAM traffic per launch from the newest committed `ncu --set full` capture of this workload
        import glob
        for tf in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")), reverse=True):
            with open(tf) as f:
                tj = json.load(f)
            if tj.get("workload") == args.workload:
                for k_, v_ in tj["traffic_bytes"].items():
                    if k_ in stages:
                        stages[k_]["traffic"] = v_
                        stages[k_]["traffic_source"] = "profiles/" + os.path.basename(tf)
                break
        dominant = max(stage_ms, key=lambda k: stage_ms[k] if k not in ("host_sync", "preprocess_bwd_accumulate") else -1)
        roof = dict(stages.get(dominant, {}))
        roof["kernel"] = dominant
        roof.setdefault("traffic", None)

        # preprocess, scan_compact, depth sort (hist + 4), sorted_scan, emit_tiles, tile sort (hist + passes),
        # tile_ranges, render_fwd, render_bwd, preprocess_bwd
        launches_per_step = 1 + 1 + (1 + 4) + 1 + 1 + (1 + tile_passes) + 1 + 1 + 1 + 1
        out = {
            "metric": "fwd+bwd renders/s", "value": value, "unit": "renders/s", "n_gpus": world, "steps": args.steps,
            "warmup": n_warm, "ms_per_step": ms / args.steps, "ms_per_view": ms / args.steps / V,
            "step_ms": {"min": step_ms[0], "median": step_ms[len(step_ms) // 2], "max": step_ms[-1]}, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload + ": " + w["desc"], "P": P, "P_visible": P_vis, "num_rendered": N,
                       "width": W, "height": H, "pairs_evaluated_fwd": pairs_eval, "pairs_blended": pairs_blend,
                       "pairs_evaluated_bwd": pairs_bwd_eval, "l2": "inputs (708 MB of Gaussians + scratch) exceed the 126 MB L2",
                       "views_per_rank_per_step": V, "global_views_per_step": V * world, "streams": args.streams,
                       "step": "each rank renders its %d views (fwd+bwd, parameter gradients accumulated in one flat bucket), view v on CUDA stream v %% %d%s" % (
                           V, max(1, args.streams), ", then ONE NCCL all-reduce of the 59 fp32/Gaussian bucket" if world > 1 else ""),
                       "roofline_view": "stage times / roofline figures are of view 0 alone on one stream (single-view fwd+bwd, write mode)",
                       "parallelism": "view-parallel dp%d, replicated Gaussians, NCCL all-reduce of 59 fp32/Gaussian" % world
                       if world > 1 else "single GPU", "peak_source": peak_src},
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": "renders/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / args.steps, "step_loss": step_loss,
                    "bucket_checksum": float(G.bucket.double().abs().sum().item()),
                    "what": "per view: camera + target image from pinned host memory (copy stream, double-buffered, overlapping the previous view's kernels), GaussianRasterizer autograd fwd+bwd + the fused training loss 0.8*L1 + 0.2*(1-SSIM) (fused_ssim.l1_ssim_loss); per step: gradients of the rank's views summed in the bucket (grad_sink), all-reduce when N>1, loss read back"},
            "gpu_launches": launches_per_step * V * args.steps,
            "single_view": {"fwd_bwd_ms": single_view_ms, "renders_per_s": 1000.0 / single_view_ms,
                            "what": "one view alone on one stream, forward + backward through the C ABI, inputs resident (latency; `value` is the throughput of the 8-view step)"},
            "stage_ms": stage_ms, "roofline": roof, "roofline_stages": stages, "cpu_baseline": cpu,
        }
        if world == 1:
            out["loss_stage"] = loss_stage_report(dev, H, W, hbm_peak)
            if not args.no_cpu_baseline:
                out["loss_stage"]["cpu_baseline"] = loss_cpu_baseline(H, W)
            try:
                out["frame_stage"] = frame_stage_report(dev, H, W, hbm_peak)
            except Exception as exc:  # informational leg only
                out["frame_stage"] = {"error": str(exc)[:300]}
            try:
                out["optimizer_stage"] = optimizer_stage_report(G, st["radii"], hbm_peak)
            except Exception as exc:  # informational leg only
                out["optimizer_stage"] = {"error": str(exc)[:300]}
        # same-box bar: the unmodified reference CUDA kernels (oracle/_ref), informational
        if not args.no_ref_cuda and world == 1:
            try:
                from oracle import ref_cuda
                if ref_cuda.available():
                    R = ref_cuda.RefRasterizer()
                    for _ in range(2):
                        f = R.forward(width=W, height=H, **common)
                        R.backward(f["radii"], dL_dcolor=dcol, dL_dinvdepth=ddep, **common)
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    nrep = 5
                    tf = 0.0
                    for _ in range(nrep):
                        t1 = time.perf_counter()
                        f = R.forward(width=W, height=H, **common)
                        tf += time.perf_counter() - t1
                        R.backward(f["radii"], dL_dcolor=dcol, dL_dinvdepth=ddep, **common)
                    torch.cuda.synchronize()
                    dt = (time.perf_counter() - t0) / nrep
                    out["reference_cuda_same_box"] = {"renders_per_s": 1.0 / dt, "ms_per_step": dt * 1e3,
                                                      "fwd_ms": tf / nrep * 1e3, "note": "unmodified reference kernels + CUB, sm_100a, incl. zero-fills"}
                    R.close()
            except Exception as exc:  # informational leg only
                out["reference_cuda_same_box"] = {"error": str(exc)[:200]}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
