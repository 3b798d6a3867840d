#!/bin/bash
set -u
mkdir -p gpurun_out
cat > gpurun_out/rows_prof.py <<'PY'
import torch, numpy as np, ctypes as C
from simubox_b200 import VoronoiCell, synth, _lib
from simubox_b200._device import ptr, stream_ptr
shape=(256,256,256)
field=synth.fields_torch(1,1,shape,cells=4)[0]
_,pn=_lib.i64(shape)
cap=4096
centres=torch.empty((cap,3),dtype=torch.float64,device='cuda'); vols=torch.empty(cap,dtype=torch.int64,device='cuda'); comp=torch.empty(shape,dtype=torch.int32,device='cuda')
n=C.c_int64()
for i in range(3):
    _lib.check(_lib.lib().sbx_domain_centres(0,3,pn,ptr(field),0.5,1,ptr(comp),cap,ptr(centres),ptr(vols),C.byref(n),stream_ptr(0)))
pts=synth.seeds_lattice(shape,16,4.0,99)
lab=VoronoiCell.label_grid(pts,0,shape=shape,box=(256.0,)*3,return_tensor=True)
for i in range(3):
    st=VoronoiCell.cell_statistics(lab, field, n_cells=4096)
torch.cuda.synchronize()
print("ok", n.value)
PY
PYTHONPATH=. python gpurun_out/rows_prof.py > gpurun_out/rows_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"seeds_|cell_stats|degree" -c 60 --csv --log-file gpurun_out/launches_rows.csv env PYTHONPATH=. python gpurun_out/rows_prof.py > gpurun_out/ncu_rows.log 2>&1
echo "ncu rc=$?"
