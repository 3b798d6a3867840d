#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_cellstats.py -m gpu -x -q > $O/pytest_cellstats.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_cellstats.log
timeout 300 python - <<'PY'
import torch, time, numpy as np
from simubox_b200 import VoronoiCell, synth, _lib
from simubox_b200.voronoi import ptr, stream_ptr
shape=(256,256,256)
pts=synth.seeds_lattice(shape,16,4.0,99)
field=synth.fields_torch(1,1,shape,cells=4)[0]
lab=VoronoiCell.label_grid(pts,0,shape=shape,box=(256.0,)*3,return_tensor=True)
S=4096; words=S//32
vol=torch.empty(S,dtype=torch.int64,device='cuda'); mass=torch.empty(S,dtype=torch.float64,device='cuda')
adj=torch.empty((S,words),dtype=torch.int32,device='cuda'); deg=torch.empty(S,dtype=torch.int32,device='cuda')
_,pn=_lib.i64(shape)
flush=torch.empty(1<<30,dtype=torch.uint8,device='cuda')
for name,args in [("volume only",(None,None,None,None)),("volume+mass",(field,mass,None,None)),("all",(field,mass,adj,deg))]:
    f,m,a,d=args
    ts=[]
    for i in range(8):
        flush.zero_()
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(_lib.lib().sbx_cell_stats(0,3,pn,ptr(lab),ptr(f),S,1,ptr(vol),ptr(m),ptr(a),ptr(d),stream_ptr(0)))
        e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ms=float(np.median(ts[2:])); byt=256**3*(4+(8 if f is not None else 0))
    print(f"cell_stats 256^3 S=4096 {name}: {ms:.4f} ms, {byt/ms/1e6:.0f} GB/s algorithmic")
PY
