#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_seeds.py -m gpu -x -q 2>&1 | tail -5
bash scripts/gpu_next_rows_timing.sh 2>&1 | grep domain_centres
