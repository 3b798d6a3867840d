#!/bin/bash
# --gpus 2, r01 v6: one 1024^3 and one 512^3 field in pruned slab mode (bulk-staged kernel per rank)
set -u
mkdir -p gpurun_out/multi6
O=gpurun_out/multi6
for G in 1024 512; do
SBX_SLAB_PATH=pruned timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2971${G:0:1} bench.py --gpus 2 --workload slab --grid $G --cells 8 --steps 5 --warmup 3 2> $O/bench_slab_${G}_pruned_n2.err | grep '^{' | tee $O/bench_slab_${G}_pruned_n2.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('slab $G pruned n2', d['ms_per_step'], d['stage_ms'], d['roofline']['frac'], d['roofline']['kernel'])"
done
