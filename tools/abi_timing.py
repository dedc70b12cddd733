import sys, time, os
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); [sys.path.insert(0,p) for p in (ROOT, ROOT+'/gaussian-splatting-npu_b200', ROOT+'/tests')]
import numpy as np, torch
import bench
from gsr_b200 import capi, synth
from gsr_b200.dist import FlatGaussians
w, scene, cam = bench.make_workload('c2')
dev=torch.device('cuda:0')
G=FlatGaussians.from_numpy(scene, dev)
t=lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
bg=torch.zeros(3,device=dev); view,proj,campos=t(cam.world_view_transform),t(cam.full_proj_transform),t(cam.camera_center)
dcol_np,ddep_np=synth.upstream_grads(w['H'],w['W'],5); dcol,ddep=t(dcol_np),t(ddep_np)
common=dict(means3D=G.xyz,opacities=G.opacity,viewmatrix=view,projmatrix=proj,cam_pos=campos,background=bg,tan_fovx=cam.tanfovx,tan_fovy=cam.tanfovy,shs=G.shs,scales=G.scales,rotations=G.rotations,sh_degree=3,antialiasing=False)
gv=G.grad_views()
def step(out):
    t0=time.perf_counter(); st=capi.forward(width=w['W'],height=w['H'],**common); t1=time.perf_counter()
    g=capi.backward(st,dL_dcolor=dcol,dL_dinvdepth=ddep,out=out,**common); t2=time.perf_counter()
    return st,g,t1-t0,t2-t1
for label,out in (('bucket-out',gv),('fresh-out',None)):
    for _ in range(3): st,g,_,_=step(out)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    tf=tb=0; t0=time.perf_counter(); e0.record()
    for _ in range(20):
        st,g,a,b=step(out); tf+=a; tb+=b
    e1.record(); torch.cuda.synchronize(); wall=time.perf_counter()-t0
    print(label,'event ms/step',e0.elapsed_time(e1)/20,'wall',wall/20*1e3,'fwd call',tf/20*1e3,'bwd call',tb/20*1e3)
# torch binding path without autograd
from diff_gaussian_rasterization import _C
empty=torch.Tensor([])
def step2():
    r=_C.rasterize_gaussians(bg,G.xyz,empty,G.opacity,G.scales,G.rotations,1.0,empty,view,proj,cam.tanfovx,cam.tanfovy,w['H'],w['W'],G.shs,3,campos,False,False,False)
    n,color,radii,gb,bb,ib,invd=r
    return _C.rasterize_gaussians_backward(bg,G.xyz,radii,empty,G.opacity,G.scales,G.rotations,1.0,empty,view,proj,cam.tanfovx,cam.tanfovy,dcol,ddep,G.shs,3,campos,gb,n,bb,ib,False,False)
for _ in range(3): step2()
torch.cuda.synchronize(); e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
t0=time.perf_counter(); e0.record()
for _ in range(20): step2()
e1.record(); torch.cuda.synchronize(); print('_C path event ms/step', e0.elapsed_time(e1)/20, 'wall', (time.perf_counter()-t0)/20*1e3)
