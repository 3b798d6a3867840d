#!/bin/bash
# r02, 8-GPU box: concurrent H2D matrix at 1/2/4/8 ranks, then the driver-style bench line at N = 8 and N = 4
mkdir -p gpurun_out/r02
nvidia-smi topo -m > gpurun_out/r02/topo.txt 2>&1; lscpu | grep -E "NUMA|Socket|^CPU\(s\)|Model name" > gpurun_out/r02/lscpu.txt; free -g | head -2 >> gpurun_out/r02/lscpu.txt
: > gpurun_out/r02/h2d_matrix.jsonl
for N in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N scripts/h2d_matrix.py >> gpurun_out/r02/h2d_matrix.jsonl 2> gpurun_out/r02/h2d_$N.err
done
cat gpurun_out/r02/h2d_matrix.jsonl
bash scripts/gpu_multi_r02.sh 8 2>&1 | grep -v "^\.\|OMP_NUM\|\*\*\*\*"
for N in 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r02/bench_default_n$N.json 2> gpurun_out/r02/bench_default_n$N.err; echo "n$N rc $?"
done
python - <<'PY'
import json
for N in (4,8):
    try:
        d=json.loads(open(f"gpurun_out/r02/bench_default_n{N}.json").read().strip().splitlines()[-1])
        print(N, "scatter", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["e2e"].get("h2d_GBps"), d["e2e"].get("numa_binding_rank0"))
    except Exception as e: print(N, "ERR", e)
PY
