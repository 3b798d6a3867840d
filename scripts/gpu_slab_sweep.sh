#!/bin/bash
# r02: slab FFT mode with every column exchanged, fused push with a persistent grid of C CTAs vs the NCCL exchange (N = $1)
N=${1:-2}
mkdir -p gpurun_out/r02
O=gpurun_out/r02/slab_push_sweep_n$N.jsonl; : > $O
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --workload slab --grid 1024 --cells 8 --steps 10 --slab-variants $2 2>/dev/null | tail -1 >> $O; }
for C in 32 48 96; do SBX_SLAB_PUSH_CTAS=$C run 2960$((C % 10)) fft_allcols; done
SBX_SLAB_CHUNKS=8 run 29611 fft_allcols
run 29612 fft_allcols_nccl
python - <<PY
import json
for i,l in enumerate(open("$O")):
    d=json.loads(l)
    for m,r in d["modes"].items(): print(i, m, r.get("exchange"), "ms", round(r["ms_per_step"],3), {k:round(v,3) for k,v in r["stage_ms"].items()})
PY
