#!/bin/bash
# --gpus N (2, 4 or 8), r01 v6: NCCL slab parity test, one 1024^3 field in pruned slab mode, sweep, sharded Voronoi and scatter
set -u
N=${1:-8}
mkdir -p gpurun_out/multi6
O=gpurun_out/multi6
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_gpu_slab.py -m gpu -x -q -k nccl > $O/pytest_slab_n$N.log 2>&1; echo "pytest slab rc=$?"; tail -2 $O/pytest_slab_n$N.log
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
SBX_SLAB_PATH=pruned timeout 600 bash -c "$(declare -f tr); N=$N; tr 29611 --workload slab --grid 1024 --cells 8 --steps 5 --warmup 3" 2> $O/bench_slab_1024_pruned_n$N.err | grep '^{' | tee $O/bench_slab_1024_pruned_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('slab 1024 pruned n$N', d['ms_per_step'], d['stage_ms'], d['roofline']['frac'], d['roofline']['kernel'])"
timeout 600 bash -c "$(declare -f tr); N=$N; tr 29631 --workload sweep --fields 32 --steps 3 --warmup 3" 2> $O/bench_sweep_n$N.err | grep '^{' | tee $O/bench_sweep_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('sweep n$N', round(d['value']), d['ms_per_step'])"
timeout 600 bash -c "$(declare -f tr); N=$N; tr 29641 --workload voronoi --steps 10 --warmup 3 --e2e-steps 1 --cpu-voxels 64" 2> $O/bench_voronoi_n$N.err | grep '^{' | tee $O/bench_voronoi_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('voronoi n$N', d['value'], d['ms_per_step'])"
timeout 600 bash -c "$(declare -f tr); N=$N; tr 29601 --steps 10 --warmup 3 --e2e-steps 1 --cpu-fields 2" 2> $O/bench_scatter_n$N.err | grep '^{' | tee $O/bench_scatter_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('scatter n$N', round(d['value']), d['ms_per_step'], round(d['e2e']['value']))"
