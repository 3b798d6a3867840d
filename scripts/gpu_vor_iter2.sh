#!/bin/bash
# Voronoi iteration: timing x2, parity tests, then the per-kernel launch list under ncu
set -u
mkdir -p gpurun_out
O=gpurun_out
bash scripts/gpu_vor_ab3.sh "SBX_VOR_PDL=1"
V="python bench.py --workload voronoi --steps 2 --warmup 3 --e2e-steps 1 --cpu-voxels 64"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"coarse_kernel|mid_kernel|label_kernel" -c 30 --csv --log-file $O/launches_voronoi.csv $V > $O/ncu_vor.log 2>&1
echo "ncu launches vor rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/launches_voronoi.csv")) if len(r) > 5]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value")
d = collections.defaultdict(list)
for r in rows[1:]:
    try: d[r[ki].split("(")[0][:40]].append(float(r[vi].replace(",", "")))
    except Exception: pass
for k, v in d.items(): print(k, len(v), "median", sorted(v)[len(v)//2])
PY
