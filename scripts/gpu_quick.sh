#!/bin/bash
# quick GPU check: parity suite + voronoi bench (+ optional extra command in $1)
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_gpu.log
python bench.py --workload voronoi > $O/bench_voronoi.json 2> $O/bench_voronoi.err; echo "bench voronoi rc=$?"; cat $O/bench_voronoi.json; tail -3 $O/bench_voronoi.err
