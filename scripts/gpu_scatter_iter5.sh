#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_scatter.py -m gpu -x -q > $O/pytest_scatter.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_scatter.log
SBX_FUSED_TMA=2 timeout 1200 python -m pytest tests/test_gpu_scatter.py -m gpu -x -q > $O/pytest_scatter_ring.log 2>&1; echo "pytest(ring forced) rc=$?"; tail -3 $O/pytest_scatter_ring.log
bash scripts/gpu_scatter_ab.sh "SBX_FUSED_TMA=2 SBX_RING_WARPS=8" "SBX_FUSED_TMA=2 SBX_RING_WARPS=8 SBX_FUSED_XC=8" "SBX_FUSED_TMA=2 SBX_RING_WARPS=16" SBX_FUSED_TMA=0 SBX_NO_FUSED=1
for v in "SBX_FUSED_TMA=1" "SBX_FUSED_TMA=0"; do
env $v python bench.py --grid 256 --cells 4 --fields 32 --max-batch 32 --host-batch 4 --cpu-fields 1 --e2e-steps 1 2>$O/b256.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('256^3 $v', d['config']['path'], d['roofline']['kernel'], round(d['value']), d['ms_per_step'], d['roofline']['frac'], d['parity_vs_cpu_sample'])"; done
P="env SBX_FUSED_TMA=2 SBX_RING_WARPS=8 python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-fields 2"
$P > $O/plain_scatter2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pruned_ring -s 4 -c 1 -f -o $O/prof_ring8 $P > $O/ncu_ring.log 2>&1
echo "ncu full rc=$?"
