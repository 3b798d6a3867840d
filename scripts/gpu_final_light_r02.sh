#!/bin/bash
# r02 final state on one B200 without the ncu passes: smoke, GPU parity suite, default bench line, reference arm
mkdir -p gpurun_out/r02
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02/smoke_final.txt 2>&1; tail -1 gpurun_out/r02/smoke_final.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r02/pytest_gpu_final.txt; cat gpurun_out/r02/pytest_gpu_final.txt
python bench.py > gpurun_out/r02/bench_final_n1.json 2> gpurun_out/r02/bench_final_n1.err; echo "bench rc $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02/bench_reference_n1.json 2>/dev/null; echo "ref rc $?"
python - <<'PY'
import json
for f in ("bench_final_n1","bench_reference_n1"):
    d=json.loads(open(f"gpurun_out/r02/{f}.json").read().strip().splitlines()[-1]); r=d.get("roofline",{})
    print(f, round(d["value"],1), d.get("ms_per_step"), r.get("kernel"), r.get("frac"), "e2e", d.get("e2e",{}).get("value"), d.get("clocks"))
    for k in ("voronoi","sweep"):
        if k in d: print("  ",k, {kk:d[k].get(kk) for kk in ("value","ms_per_step","error","leg_seconds","parity_vs_cpu_sample")})
PY
