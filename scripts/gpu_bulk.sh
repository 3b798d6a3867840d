#!/bin/bash
# bulk-staged blocked kernel: parity tests, then one 512^3 / 1024^3 field with and without it
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_scatter.py tests/test_gpu_slab.py -m gpu -x -q > $O/pytest_bulk.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_bulk.log
for G in 512 1024; do
for B in 0 1; do
  SBX_BULK=$B timeout 600 python bench.py --workload slab --grid $G --cells 8 --steps 5 --warmup 3 2> $O/bench_slab_${G}_bulk$B.err | grep '^{' | tee $O/bench_slab_${G}_bulk$B.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('slab $G bulk=$B', d['config']['path'], d['ms_per_step'], d['roofline']['frac'], d['stage_ms'])"
done; done

