#!/bin/bash
# One gpurun call (v6): smoke, GPU parity suite, every bench line + reference arm, ncu launch lists and one
# full capture per changed top kernel (each only after the identical command exited 0 without ncu).
set -u
mkdir -p gpurun_out/final6
O=gpurun_out/final6
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python bench.py > $O/bench_scatter.json 2> $O/bench_scatter.err; echo "bench scatter rc=$?"; cat $O/bench_scatter.json
python bench.py --workload voronoi > $O/bench_voronoi.json 2> $O/bench_voronoi.err; echo "bench voronoi rc=$?"; cat $O/bench_voronoi.json
python bench.py --grid 256 --cells 4 --fields 32 --max-batch 32 --host-batch 4 --cpu-fields 2 --e2e-steps 1 > $O/bench_scatter_256.json 2> $O/bench_scatter_256.err; echo "bench 256 rc=$?"; cat $O/bench_scatter_256.json
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench ref rc=$?"; cat $O/bench_reference.json
for G in 512 1024; do
  timeout 600 python bench.py --workload slab --grid $G --cells 8 --steps 5 --warmup 3 2> $O/bench_slab_${G}_n1.err | grep '^{' > $O/bench_slab_${G}_n1.json; echo "slab $G rc=$?"; cat $O/bench_slab_${G}_n1.json
done
timeout 900 python bench.py --workload sweep --fields 64 --steps 5 --warmup 3 > $O/bench_sweep_n1.json 2> $O/bench_sweep_n1.err; echo "sweep rc=$?"; cat $O/bench_sweep_n1.json
P="python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-fields 2"
$P > $O/plain_scatter.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"pruned|shell|peaks|slab" -c 400 --csv --log-file $O/launches_scatter.csv $P > $O/ncu_scatter.log 2>&1
echo "ncu launches scatter rc=$?"
V="python bench.py --workload voronoi --steps 2 --warmup 3 --e2e-steps 1 --cpu-voxels 256"
$V > $O/plain_vor.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"coarse_kernel|mid_kernel|label_kernel" -c 200 --csv --log-file $O/launches_voronoi.csv $V > $O/ncu_vor.log 2>&1
echo "ncu launches vor rc=$?"
$V > $O/plain_vor2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"coarse_kernel|mid_kernel|label_kernel" -s 9 -c 3 -f -o $O/prof_voronoi $V > $O/ncu_vorfull.log 2>&1
echo "ncu full vor rc=$?"
S="python bench.py --workload slab --grid 1024 --cells 8 --steps 2 --warmup 3"
$S > $O/plain_slab.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"pruned|fused|shell|peaks|slab" -c 100 --csv --log-file $O/launches_slab_1024.csv $S > $O/ncu_slab.log 2>&1
echo "ncu launches slab rc=$?"
$S > $O/plain_slab2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pruned_bulk -s 3 -c 1 -f -o $O/prof_bulk $S > $O/ncu_bulkfull.log 2>&1
echo "ncu full bulk rc=$?"
