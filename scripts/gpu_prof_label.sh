#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
V="python bench.py --workload voronoi --steps 2 --warmup 3 --e2e-steps 1 --cpu-voxels 256"
$V > $O/plain_vor2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"label_kernel" -s 3 -c 1 -f -o $O/prof_label $V > $O/ncu_label.log 2>&1
echo "ncu full label rc=$?"
