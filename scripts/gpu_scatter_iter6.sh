#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_scatter.py -m gpu -x -q > $O/pytest_scatter.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_scatter.log
run() { env $1 python bench.py --cpu-fields 2 --e2e-steps 1 $2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1 $2', d['roofline']['kernel'], round(d['value']), d['ms_per_step'], d['roofline']['frac'], d['parity_vs_cpu_sample'])"; }
run "SBX_FUSED_TMA=1" "--max-batch 128"
run "SBX_FUSED_TMA=1" "--max-batch 256"
run "SBX_FUSED_TMA=2 SBX_RING_WARPS=8" "--max-batch 256"
run "SBX_FUSED_TMA=2 SBX_RING_WARPS=8 SBX_FUSED_XC=8" "--max-batch 256"
run "SBX_FUSED_TMA=2 SBX_RING_WARPS=16 SBX_FUSED_XC=8" "--max-batch 256"
run "SBX_NO_FUSED=1" "--max-batch 128"
python bench.py --grid 256 --cells 4 --fields 32 --max-batch 32 --host-batch 4 --cpu-fields 1 --e2e-steps 1 2>$O/b256.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('256^3', d['config']['path'], d['roofline']['kernel'], round(d['value']), d['ms_per_step'], d['roofline']['frac'], d['parity_vs_cpu_sample'])"
