#!/bin/bash
# A/B: label kernel time under env variants (ncu launch times, same command)
set -u
mkdir -p gpurun_out
O=gpurun_out
V="python bench.py --workload voronoi --steps 2 --warmup 3 --e2e-steps 1 --cpu-voxels 256"
for variant in "$@"; do
  env $variant $V > $O/plain_ab.log 2>&1 && env $variant ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"label_kernel" -c 20 --csv --log-file $O/launches_ab.csv $V > $O/ncu_ab.log 2>&1
  grep -v "^==" $O/launches_ab.csv | python -c "
import csv,sys
v=[float(r['Metric Value'].replace(',',''))*(1e-3 if r['Metric Unit']=='ns' else 1) for r in csv.DictReader(sys.stdin)]
print('$variant', len(v), sum(v)/len(v), min(v))
"
  env $variant python bench.py --workload voronoi --cpu-voxels 256 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  bench ms', d['ms_per_step'])"
done
