#!/bin/bash
# Voronoi iteration: parity tests, bench, per-kernel launch times
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_voronoi.py -m gpu -x -q > $O/pytest_vor.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_vor.log
python bench.py --workload voronoi > $O/bench_voronoi.json 2> $O/bench_voronoi.err; echo "bench voronoi rc=$?"; cut -c1-260 $O/bench_voronoi.json; grep -o "evaluations_per_live_voxel[^,]*" $O/bench_voronoi.json
SBX_LABEL_CTAS=4 python bench.py --workload voronoi > $O/bench_voronoi4.json 2>&1; cut -c1-260 $O/bench_voronoi4.json
V="python bench.py --workload voronoi --steps 2 --warmup 3 --e2e-steps 1 --cpu-voxels 256"
$V > $O/plain_vor.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"coarse_kernel|mid_kernel|label_kernel" -c 60 --csv --log-file $O/launches_voronoi.csv $V > $O/ncu_vor.log 2>&1
echo "ncu launches vor rc=$?"
grep -v "^==" $O/launches_voronoi.csv | python -c "
import csv,sys,collections
agg=collections.defaultdict(list)
for row in csv.DictReader(sys.stdin):
    agg[row['Kernel Name'].split('(')[0]].append(float(row['Metric Value'].replace(',',''))*(1e-3 if row['Metric Unit']=='ns' else 1))
for k,v in agg.items(): print(k, len(v), sum(v)/len(v))
"
export SBX_LABEL_CTAS=4
$V > $O/plain_vor4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"label_kernel" -c 20 --csv --log-file $O/launches_voronoi4.csv $V > $O/ncu_vor4.log 2>&1
grep -v "^==" $O/launches_voronoi4.csv | python -c "
import csv,sys
v=[float(r['Metric Value'].replace(',',''))*(1e-3 if r['Metric Unit']=='ns' else 1) for r in csv.DictReader(sys.stdin)]
print('label CTAS=4', len(v), sum(v)/len(v))
"
