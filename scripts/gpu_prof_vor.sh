#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
V="python bench.py --workload voronoi --steps 2 --warmup 3 --e2e-steps 1 --cpu-voxels 256"
$V > $O/plain_vor.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"coarse_kernel|mid_kernel|label_kernel" -c 60 --csv --log-file $O/launches_voronoi.csv $V > $O/ncu_vor.log 2>&1
echo "ncu launches vor rc=$?"
$V > $O/plain_vor2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"coarse_kernel|mid_kernel|label_kernel" -s 9 -c 3 -f -o $O/prof_voronoi $V > $O/ncu_vorfull.log 2>&1
echo "ncu full vor rc=$?"
