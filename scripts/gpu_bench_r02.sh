#!/bin/bash
# r02: GPU parity suite, the default bench line (with sub-records), the dense (cuFFT) path and its ncu capture
mkdir -p gpurun_out/r02
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r02/pytest_gpu.txt; cat gpurun_out/r02/pytest_gpu.txt
python bench.py > gpurun_out/r02/bench_default.json 2> gpurun_out/r02/bench_default.err; echo "bench rc $?"; tail -c 1500 gpurun_out/r02/bench_default.err
python bench.py --path dense --sub none --steps 10 > gpurun_out/r02/bench_dense.json 2> gpurun_out/r02/bench_dense.err; echo "dense rc $?"
python bench.py --path dense --sub none --steps 3 --warmup 3 --cpu-fields 8 > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:shell_histogram -s 8 -c 2 -o gpurun_out/r02/hist_dense python bench.py --path dense --sub none --steps 3 --warmup 3 --cpu-fields 8 > gpurun_out/r02/ncu_hist.log 2>&1
python - <<'PY'
import json
for f in ("bench_default","bench_dense"):
    try:
        d=json.loads(open(f"gpurun_out/r02/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline"].get("step_split_ms"), "e2e", d["e2e"]["value"])
        for k in ("voronoi","sweep"):
            if k in d: print("  ",k, {kk:d[k].get(kk) for kk in ("value","ms_per_step","error","leg_seconds","parity_vs_cpu_sample")})
    except Exception as e: print(f, "ERR", e)
PY
