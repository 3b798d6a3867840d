#!/bin/bash
# r02: N-GPU run (N = $1): NCCL slab tests, then the driver-style bench line with slab / sharded / sweep sub-records
N=${1:-2}
mkdir -p gpurun_out/r02
if [ -z "$SKIP_TESTS" ]; then python -m pytest tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r02/pytest_slab_n$N.txt; cat gpurun_out/r02/pytest_slab_n$N.txt; fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r02/bench_default_n$N.json 2> gpurun_out/r02/bench_default_n$N.err; echo "bench rc $?"; tail -c 800 gpurun_out/r02/bench_default_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02/bench_default_n$N.json").read().strip().splitlines()[-1])
print("scatter", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"].get("h2d_GBps"))
for k in ("voronoi","voronoi_sharded","slab","sweep"):
    v=d.get(k)
    if not v: print(k, None); continue
    if "error" in v: print(k, "ERROR", v["error"]); continue
    if k=="slab":
        print(k, "parity", v.get("slab_parity"), v.get("slab_parity_worst_rel"), "leg", v.get("leg_seconds"))
        for m,r in v["modes"].items(): print("   ", m, r.get("exchange"), round(r["ms_per_step"],3), {k:round(x,3) for k,x in r["stage_ms"].items()}, "a2a GB/s", r.get("all_to_all_GBps_per_gpu"), "hidden", r.get("all_to_all_hidden_frac"), "frac", round(r["roofline"]["frac"],3))
    elif k=="voronoi_sharded":
        for g in v["grids"]: print(k, g)
    else: print(k, v.get("value"), v.get("ms_per_step"), "leg", v.get("leg_seconds"))
PY
