#!/bin/bash
# --gpus 2: NCCL slab tests (both modes), slab benches, sweep at N=2
set -u
mkdir -p gpurun_out/multi
O=gpurun_out/multi
timeout 900 python -m pytest tests/test_gpu_slab.py -m gpu -x -q -k nccl > $O/pytest_slab_nccl.log 2>&1; echo "pytest nccl rc=$?"; tail -4 $O/pytest_slab_nccl.log
i=0
for G in 512 1024; do for P in pruned fft; do
i=$((i+1))
SBX_SLAB_PATH=$P timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2956$i bench.py --gpus 2 --workload slab --grid $G --cells 8 --steps 5 --warmup 3 2> $O/bench_slab_${G}_${P}_n2.err | grep '^{' | tee $O/bench_slab_${G}_${P}_n2.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('slab $G $P n2', d['config']['path'], d['ms_per_step'], d['roofline']['frac'], d['stage_ms'])"
done; done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --workload sweep --fields 32 --steps 3 --warmup 3 2> $O/bench_sweep_n2.err | grep '^{' | tee $O/bench_sweep_n2.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('sweep n2', d['value'], d['ms_per_step'])"
