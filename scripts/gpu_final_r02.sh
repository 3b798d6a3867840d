#!/bin/bash
# r02 final evidence on one B200: smoke, GPU parity suite, default bench line + reference arm, ncu launch list and
# --set full capture of the same bench command, two A/B lines of the headline kernel
mkdir -p gpurun_out/r02
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02/smoke_final.txt 2>&1; tail -2 gpurun_out/r02/smoke_final.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r02/pytest_gpu_final.txt; cat gpurun_out/r02/pytest_gpu_final.txt
python bench.py > gpurun_out/r02/bench_final_n1.json 2> gpurun_out/r02/bench_final_n1.err; echo "bench rc $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02/bench_reference_n1.json 2>/dev/null; echo "ref rc $?"
SBX_FUSED_BLOCKED=1 python bench.py --sub none --steps 10 > gpurun_out/r02/bench_ab_blocked128.json 2>/dev/null; echo "ab1 rc $?"
SBX_FUSED_TMA=2 python bench.py --sub none --steps 10 > gpurun_out/r02/bench_ab_ring128.json 2>/dev/null; echo "ab2 rc $?"
CMD="python bench.py --steps 2 --warmup 3 --sub voronoi --cpu-fields 8 --cpu-voxels 512 --e2e-steps 1"
$CMD > gpurun_out/r02/plain_final.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02/launches_final.csv $CMD > gpurun_out/r02/ncu_launches.log 2>&1
$CMD > gpurun_out/r02/plain_final2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'pruned_fused_kernel|label_kernel|coarse_kernel|mid_kernel' -s 6 -c 6 -o gpurun_out/r02/final_full $CMD > gpurun_out/r02/ncu_full.log 2>&1
python - <<'PY'
import json
for f in ("bench_final_n1","bench_reference_n1","bench_ab_blocked128","bench_ab_ring128"):
    try:
        d=json.loads(open(f"gpurun_out/r02/{f}.json").read().strip().splitlines()[-1])
        r=d.get("roofline",{})
        print(f, round(d["value"],1), d.get("ms_per_step"), r.get("kernel"), r.get("frac"), "e2e", d.get("e2e",{}).get("value"), "cpu", d.get("cpu_baseline",{}).get("value"), d.get("clocks"))
        for k in ("voronoi","sweep"):
            if k in d: print("  ",k, {kk:d[k].get(kk) for kk in ("value","ms_per_step","error","leg_seconds","parity_vs_cpu_sample")}, d[k].get("roofline",{}).get("frac"))
    except Exception as e: print(f, "ERR", e)
PY
