#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests/test_gpu_scatter.py tests/test_gpu_slab.py -m gpu -x -q > $O/pytest_blocked.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_blocked.log
for G in 512 1024; do
  timeout 600 python bench.py --workload slab --grid $G --cells 8 --steps 5 --warmup 3 2> $O/bench_slab_${G}_n1.err | tee $O/bench_slab_${G}_n1.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('slab $G', d['config']['path'], d['ms_per_step'], d['roofline']['frac'], d['stage_ms'])"
done
python bench.py --grid 512 --cells 8 --fields 4 --max-batch 4 --host-batch 1 --cpu-fields 1 --e2e-steps 1 --steps 5 2>$O/b512.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('512^3 x4', d['config']['path'], d['roofline']['kernel'], round(d['value'],1), d['ms_per_step'], d['roofline']['frac'], d['parity_vs_cpu_sample'])"
