#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
P="python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-fields 2"
$P > $O/plain_scatter2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pruned_fused -s 4 -c 1 -f -o $O/prof_tma $P > $O/ncu_tma.log 2>&1
echo "ncu full rc=$?"
