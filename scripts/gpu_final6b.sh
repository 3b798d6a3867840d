#!/bin/bash
# v6b: after the L2 prefetch in the fused LDG kernel -- GPU suite, default bench line, launch list and full capture of that kernel
set -u
mkdir -p gpurun_out/final6b
O=gpurun_out/final6b
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_gpu.log
python bench.py > $O/bench_scatter.json 2> $O/bench_scatter.err; echo "bench scatter rc=$?"; cat $O/bench_scatter.json
P="python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-fields 2"
$P > $O/plain_scatter.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"pruned|shell|peaks|slab" -c 400 --csv --log-file $O/launches_scatter.csv $P > $O/ncu_scatter.log 2>&1
echo "ncu launches scatter rc=$?"
$P > $O/plain_scatter2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pruned_fused -s 4 -c 1 -f -o $O/prof_fused $P > $O/ncu_fused.log 2>&1
echo "ncu full fused rc=$?"
