#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_scatter.py -m gpu -x -q > $O/pytest_scatter.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_scatter.log
bash scripts/gpu_scatter_ab.sh SBX_FUSED_TMA=0 SBX_FUSED_TMA=1 "SBX_FUSED_TMA=1 SBX_FUSED_XC=8" "SBX_FUSED_TMA=1 SBX_FUSED_XC=2" "SBX_FUSED_TMA=1 SBX_RING_SLOTS=12"
python bench.py --grid 256 --cells 4 --fields 32 --max-batch 32 --host-batch 4 --cpu-fields 1 --e2e-steps 1 2>$O/b256.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('256^3', d['config']['path'], round(d['value']), d['ms_per_step'], d['roofline']['frac'], d['parity_vs_cpu_sample'])"; tail -2 $O/b256.err
