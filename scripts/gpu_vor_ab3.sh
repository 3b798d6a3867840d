#!/bin/bash
set -u
mkdir -p gpurun_out
for v in "$@"; do
for rep in 1 2; do
  env $v python bench.py --workload voronoi --cpu-voxels 64 --e2e-steps 1 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['roofline']['evaluations_per_live_voxel'], d['parity_vs_cpu_sample'])"
done; done
timeout 900 python -m pytest tests/test_gpu_voronoi.py -m gpu -x -q > gpurun_out/pytest_vor.log 2>&1; echo "pytest(default) rc=$?"; tail -3 gpurun_out/pytest_vor.log
