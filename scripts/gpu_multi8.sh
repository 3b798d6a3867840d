#!/bin/bash
# --gpus 8 (or 4): sharded scatter bench, sweep, slab field in both modes
set -u
N=${1:-8}
mkdir -p gpurun_out/multi
O=gpurun_out/multi
nvidia-smi -L | wc -l
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
timeout 600 bash -c "$(declare -f tr); N=$N; tr 29601 --steps 10 --warmup 3 --e2e-steps 1 --cpu-fields 2" 2> $O/bench_scatter_n$N.err | grep '^{' | tee $O/bench_scatter_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('scatter n$N', round(d['value']), d['ms_per_step'], round(d['e2e']['value']))"
for P in pruned fft; do
SBX_SLAB_PATH=$P timeout 600 bash -c "$(declare -f tr); N=$N; tr 2961${#P} --workload slab --grid 1024 --cells 8 --steps 5 --warmup 3" 2> $O/bench_slab_1024_${P}_n$N.err | grep '^{' | tee $O/bench_slab_1024_${P}_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('slab 1024 $P n$N', d['ms_per_step'], d['stage_ms'])"
done
timeout 600 bash -c "$(declare -f tr); N=$N; tr 29631 --workload sweep --fields 32 --steps 3 --warmup 3" 2> $O/bench_sweep_n$N.err | grep '^{' | tee $O/bench_sweep_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('sweep n$N', round(d['value']), d['ms_per_step'])"
timeout 600 bash -c "$(declare -f tr); N=$N; tr 29641 --workload voronoi --steps 10 --warmup 3 --e2e-steps 1 --cpu-voxels 64" 2> $O/bench_voronoi_n$N.err | grep '^{' | tee $O/bench_voronoi_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('voronoi n$N', d['value'], d['ms_per_step'])"
