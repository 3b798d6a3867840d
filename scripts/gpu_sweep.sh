#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_seeds.py -m gpu -x -q 2>&1 | tail -2
bash scripts/gpu_next_rows_timing.sh 2>&1 | grep domain_centres
timeout 900 python bench.py --workload sweep --fields 64 --steps 5 --warmup 3 > gpurun_out/bench_sweep_n1.json 2> gpurun_out/bench_sweep_n1.err; echo "sweep rc=$?"; cat gpurun_out/bench_sweep_n1.json; tail -3 gpurun_out/bench_sweep_n1.err
