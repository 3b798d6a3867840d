#!/bin/bash
# usage: gpu_pytest.sh <pytest args...>   (log in gpurun_out/pytest_last.log)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest "$@" -m gpu -x -q > gpurun_out/pytest_last.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_last.log
