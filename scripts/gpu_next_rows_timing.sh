#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python - <<'PY' 2>&1 | tee gpurun_out/next_rows_timing.log
import torch, time, numpy as np, ctypes as C
from simubox_b200 import VoronoiCell, synth, _lib, splat_particles
from simubox_b200._device import ptr, stream_ptr
def timed(fn, n=6):
    ts=[]
    for i in range(n):
        torch.cuda.synchronize(); e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return float(np.median(ts[1:]))
shape=(256,256,256)
field=synth.fields_torch(1,1,shape,cells=4)[0]
_,pn=_lib.i64(shape)
cap=4096
centres=torch.empty((cap,3),dtype=torch.float64,device='cuda'); vols=torch.empty(cap,dtype=torch.int64,device='cuda'); comp=torch.empty(shape,dtype=torch.int32,device='cuda')
n=C.c_int64()
def seeds():
    _lib.check(_lib.lib().sbx_domain_centres(0,3,pn,ptr(field),0.5,1,ptr(comp),cap,ptr(centres),ptr(vols),C.byref(n),stream_ptr(0)))
ms=timed(seeds); print(f"domain_centres 256^3 BCC (fg 22%): {ms:.3f} ms, {n.value} components, {256**3/ms/1e6:.1f} G voxels/s")
noisy=torch.rand(shape,dtype=torch.float64,device='cuda')
def seeds2():
    _lib.check(_lib.lib().sbx_domain_centres(0,3,pn,ptr(noisy),0.7,1,ptr(comp),cap,ptr(centres),ptr(vols),C.byref(n),stream_ptr(0)))
ms=timed(seeds2); print(f"domain_centres 256^3 random fg 30%: {ms:.3f} ms, {n.value} components")
# splat: 1M and 4M particles on 128^3, L=128 (h=1), r_cut=2 -> 216-point stencil
for P in (1_000_000, 4_000_000):
    L=(128.0,)*3; N=(128,128,128)
    xyz=(torch.rand((P,3),dtype=torch.float64,device='cuda')-0.5)*128.0
    typ=torch.randint(0,3,(P,),dtype=torch.int32,device='cuda')
    ms=timed(lambda: splat_particles(xyz,typ,3,N,L,2.0,return_tensor=True))
    print(f"splat {P} particles -> 3 x 128^3, r_cut=2 (216-point stencil): {ms:.3f} ms, {P/ms/1e3:.1f} M particles/s")
PY
