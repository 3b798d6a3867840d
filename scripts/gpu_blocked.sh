#!/bin/bash
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests/test_gpu_scatter.py tests/test_gpu_slab.py -m gpu -x -q > $O/pytest_blocked.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_blocked.log
for G in 512 1024; do
  timeout 600 python bench.py --workload slab --grid $G --cells 8 --steps 5 --warmup 3 > $O/bench_slab_${G}_n1.json 2> $O/bench_slab_${G}_n1.err; echo "slab $G rc=$?"; cat $O/bench_slab_${G}_n1.json; tail -2 $O/bench_slab_${G}_n1.err
done
SBX_SLAB_PATH=fft timeout 600 python bench.py --workload slab --grid 1024 --cells 8 --steps 3 --warmup 3 > $O/bench_slab_1024_fft_n1.json 2> $O/bench_slab_1024_fft_n1.err; echo "slab 1024 fft rc=$?"; cat $O/bench_slab_1024_fft_n1.json; tail -2 $O/bench_slab_1024_fft_n1.err
python bench.py --cpu-fields 2 --e2e-steps 1 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('default 128^3', d['config']['path'], d['roofline']['kernel'], round(d['value']), d['ms_per_step'], d['parity_vs_cpu_sample'])"
