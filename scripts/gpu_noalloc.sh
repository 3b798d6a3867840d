#!/bin/bash
# the allocation-free entry point: its test alone in a fresh process, then at the end of the whole GPU suite
mkdir -p gpurun_out/r02
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r02/pytest_noalloc_alone.txt; cat gpurun_out/r02/pytest_noalloc_alone.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r02/pytest_gpu.txt; cat gpurun_out/r02/pytest_gpu.txt
