#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests/test_gpu_scatter.py tests/test_gpu_slab.py -m gpu -x -q > $O/pytest_blocked.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_blocked.log
for G in 256 512 1024; do
  timeout 600 python bench.py --workload slab --grid $G --cells 8 --steps 5 --warmup 3 2> $O/bench_slab_${G}_n1.err | grep '^{' | tee $O/bench_slab_${G}_n1.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('slab $G', d['config']['path'], d['ms_per_step'], d['roofline']['frac'], d['stage_ms'])"
done
bash scripts/gpu_scatter_ab.sh SBX_FUSED_BLOCKED=0 SBX_FUSED_BLOCKED=1
