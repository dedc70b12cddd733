"""A compact pass over every kernel of libgsr_b200.so for compute-sanitizer (memcheck / racecheck / initcheck / synccheck).

    compute-sanitizer --tool memcheck  --error-exitcode 9 python tools/sanitize.py > gpurun_out/san_memcheck.log 2>&1
    compute-sanitizer --tool racecheck --error-exitcode 9 python tools/sanitize.py > gpurun_out/san_racecheck.log 2>&1
    compute-sanitizer --tool synccheck --error-exitcode 9 python tools/sanitize.py > gpurun_out/san_synccheck.log 2>&1
    compute-sanitizer --tool initcheck --error-exitcode 9 python tools/sanitize.py > gpurun_out/san_initcheck.log 2>&1

Sizes are small (the tools slow kernels down 10-100x) but cross every structural boundary: more than one sort tile and
several look-back partitions, ranges longer than one staging batch in the blends, two views through the batch entry
points, a ragged image (W, H not multiples of 16), the allocation-free sort, the loss, the optimizer and frame output.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from gsr_b200 import capi, synth  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    P, W, H = 30_000, 331, 203
    scene = synth.trained_like_scene(P, 3, s0=0.03)
    cams = synth.orbit_cameras(8, W, H, 3)[:2]
    dcol, ddep = (t(x) for x in synth.upstream_grads(H, W, 5))
    shared = dict(means3D=t(scene["means3D"]), opacities=t(scene["opacities"]), background=torch.zeros(3, device=dev),
                  shs=t(scene["shs"]), scales=t(scene["scales"]), rotations=t(scene["rotations"]), sh_degree=3)
    cam_kw = [dict(viewmatrix=t(c.world_view_transform), projmatrix=t(c.full_proj_transform), cam_pos=t(c.camera_center),
                   tan_fovx=c.tanfovx, tan_fovy=c.tanfovy) for c in cams]
    for aa in (False, True):
        st = capi.forward(width=W, height=H, antialiasing=aa, **cam_kw[0], **shared)
        g = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, antialiasing=aa, **cam_kw[0], **shared)
        g2 = capi.backward(st, dL_dcolor=dcol, dL_dinvdepth=ddep, antialiasing=aa, accumulate=True,
                           out={k: g[k] for k in ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations")},
                           **cam_kw[0], **shared)
        torch.cuda.synchronize()
        print("single view aa=%d: N=%d" % (aa, st["num_rendered"]), flush=True)
    capi.mark_visible(shared["means3D"], cam_kw[0]["viewmatrix"], cam_kw[0]["projmatrix"])
    states = capi.forward_batch(cam_kw, width=W, height=H, **shared)
    gb = capi.backward_batch(states, cam_kw, dL_dcolor=dcol, dL_dinvdepth=ddep, **shared)
    torch.cuda.synchronize()
    print("batch of 2 views: N=%s" % [s["num_rendered"] for s in states], flush=True)

    n = 70_001
    gen = torch.Generator(device=dev).manual_seed(1)
    keys = torch.randint(0, 1 << 45, (n,), generator=gen, device=dev)
    vals = torch.arange(n, dtype=torch.int32, device=dev)
    ko, vo = capi.sort_pairs(keys.clone(), vals.clone(), 45)
    ko2, vo2 = torch.empty_like(keys), torch.empty_like(vals)
    capi.sort_pairs_static(keys.clone(), vals.clone(), ko2, vo2, 45)
    torch.cuda.synchronize()
    assert torch.equal(ko, ko2) and torch.equal(vo, vo2)
    print("sorts ok", flush=True)

    from fused_ssim import fused_ssim, l1_ssim_loss
    img = torch.rand(3, H, W, device=dev, requires_grad=True)
    gt = torch.rand(3, H, W, device=dev)
    l1_ssim_loss(img, gt, 0.2).backward()
    fused_ssim(img.unsqueeze(0), gt.unsqueeze(0)).backward()
    torch.cuda.synchronize()
    print("loss ok", flush=True)

    from gsr_b200.dist import FlatGaussians
    from gsr_b200.optim import GaussianAdam
    flat = FlatGaussians.from_numpy(scene, dev)
    opt = GaussianAdam(flat)
    for v in flat.grad_views().values():
        v.copy_(torch.randn(v.shape, device=dev))
    opt.step()
    opt.step(visible=(st["radii"] > 0), zero_grad=True)
    torch.cuda.synchronize()
    print("optimizer ok", flush=True)

    from gsr_b200 import frames
    rgb = frames.to_rgb8(st["out_color"]) if hasattr(frames, "to_rgb8") else None
    torch.cuda.synchronize()
    print("frame output %s" % ("ok" if rgb is not None else "skipped"), flush=True)
    print("sanitize pass complete", flush=True)


if __name__ == "__main__":
    main()
