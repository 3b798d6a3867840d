#!/bin/bash
# --gpus 2: real NCCL slab path, sharded bench at N=2, slab bench
set -u
mkdir -p gpurun_out/multi
O=gpurun_out/multi
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_slab.py -m gpu -x -q > $O/pytest_slab.log 2>&1; echo "pytest slab rc=$?"; tail -5 $O/pytest_slab.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 --e2e-steps 1 --cpu-fields 2 > $O/bench_scatter_n2.json 2> $O/bench_scatter_n2.err; echo "bench n2 rc=$?"; cat $O/bench_scatter_n2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --workload voronoi --steps 10 --warmup 3 --e2e-steps 1 --cpu-voxels 256 > $O/bench_voronoi_n2.json 2> $O/bench_voronoi_n2.err; echo "bench vor n2 rc=$?"; cat $O/bench_voronoi_n2.json
for G in 512 1024; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2954${G:0:1} bench.py --gpus 2 --workload slab --grid $G --cells 4 --steps 5 --warmup 3 > $O/bench_slab_${G}_n2.json 2> $O/bench_slab_${G}_n2.err; echo "bench slab $G n2 rc=$?"; cat $O/bench_slab_${G}_n2.json; tail -3 $O/bench_slab_${G}_n2.err
done
timeout 600 python bench.py --workload slab --grid 512 --cells 4 --steps 5 --warmup 3 > $O/bench_slab_512_n1.json 2> $O/bench_slab_512_n1.err; echo "bench slab 512 n1 rc=$?"; cat $O/bench_slab_512_n1.json; tail -3 $O/bench_slab_512_n1.err
