"""Diagnostic: run-to-run noise of the fp32 atomics in the reference's own backward vs our difference to it (C4)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-npu_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import bench, cases
wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
w, scene, cam = bench.make_workload(wl)
opts = dict(bg=(0.0, 0.0, 0.0), antialiasing=w["aa"], mode="sh")
grads = cases.grads_for(cam)
r1 = cases.run_ref(scene, cam, opts, grads)
r2 = cases.run_ref(scene, cam, opts, grads)
c1 = cases.run_cuda(scene, cam, opts, grads)
c2 = cases.run_cuda(scene, cam, opts, grads)
for k in cases.BWD_KEYS:
    if r1.get(k) is None or np.asarray(r1[k]).size == 0:
        continue
    y = np.asarray(r1[k], np.float64); s = np.abs(y).max()
    f = lambda a, b: np.abs(np.asarray(a[k], np.float64) - np.asarray(b[k], np.float64)).max() / s
    d = np.abs(np.asarray(c1[k], np.float64) - y).reshape(y.shape[0], -1).max(axis=1)
    i = int(d.argmax())
    print("%-14s ref-vs-ref %.2e  ours-vs-ours %.2e  ours-vs-ref %.2e  (worst row %d: ref %s ours %s ref2 %s)" % (
        k, f(r2, r1), f(c2, c1), f(c1, r1), i, np.asarray(r1[k]).reshape(y.shape[0], -1)[i][:4],
        np.asarray(c1[k]).reshape(y.shape[0], -1)[i][:4], np.asarray(r2[k]).reshape(y.shape[0], -1)[i][:4]))
# fp64-accumulating CPU oracle as the arbiter on the worst dL_dscales row
o = cases.run_oracle(scene, cam, opts, grads)
y = np.asarray(r1["dL_dscales"], np.float64)
i = int(np.abs(np.asarray(c1["dL_dscales"], np.float64) - y).max(axis=1).argmax())
for k in ("dL_dconic", "dL_dcov3D", "dL_dscales"):
    tr = np.asarray(o[k], np.float64).reshape(y.shape[0], -1)[i]
    e = lambda a: np.abs(np.asarray(a[k], np.float64).reshape(y.shape[0], -1)[i] - tr).max() / (np.abs(tr).max() + 1e-30)
    print("row %d %-10s oracle %s | rel err: ref run1 %.2e  ref run2 %.2e  ours %.2e" % (i, k, tr[:4], e(r1), e(r2), e(c1)))
for k in ("dL_dscales", "dL_dcov3D", "dL_drotations"):
    tr = np.asarray(o[k], np.float64); s = np.abs(tr).max()
    print("%-14s whole tensor vs oracle: ref %.2e  ours %.2e" % (k, np.abs(np.asarray(r1[k], np.float64) - tr).max() / s,
                                                                  np.abs(np.asarray(c1[k], np.float64) - tr).max() / s))
