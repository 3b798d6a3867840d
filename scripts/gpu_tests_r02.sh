#!/bin/bash
# r02: full GPU parity suite + the f2/f3 relation probe
mkdir -p gpurun_out/r02
python scripts/probe_f2f3.py > gpurun_out/r02/probe_f2f3.txt 2>&1; tail -30 gpurun_out/r02/probe_f2f3.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r02/pytest_gpu.txt; cat gpurun_out/r02/pytest_gpu.txt
