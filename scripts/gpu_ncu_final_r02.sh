#!/bin/bash
# r02: ncu launch list (own kernels only) and --set full capture of the Voronoi kernels for the default bench command
mkdir -p gpurun_out/r02
CMD="python bench.py --steps 2 --warmup 3 --sub voronoi --cpu-fields 8 --cpu-voxels 512 --e2e-steps 1"
K='regex:pruned_|label_|coarse_kernel|mid_kernel|cell_|shell_|peaks_kernel|regular_fft|vector_fft'
$CMD > gpurun_out/r02/plain_final.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/r02/launches_final.csv $CMD > gpurun_out/r02/ncu_launches.log 2>&1
$CMD > gpurun_out/r02/plain_final2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:label_kernel|coarse_kernel|mid_kernel' -s 9 -c 3 -o gpurun_out/r02/final_voronoi $CMD > gpurun_out/r02/ncu_full_vor.log 2>&1
tail -3 gpurun_out/r02/ncu_full_vor.log; wc -l gpurun_out/r02/launches_final.csv
