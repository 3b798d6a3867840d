#!/bin/bash
# r02: launch list + full capture of the cell-list labelling kernels (env passed through)
set -x
mkdir -p gpurun_out/r02
python scripts/vor_probe.py --steps 3 > gpurun_out/r02/plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'cell|label' -s 9 -c 9 --csv --log-file gpurun_out/r02/launches_vor_cells.csv python scripts/vor_probe.py --steps 3 > gpurun_out/r02/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'label_cells|cell_count|cell_scatter' -s 9 -c 3 -o gpurun_out/r02/vor_cells python scripts/vor_probe.py --steps 3 > gpurun_out/r02/ncu2.log 2>&1
tail -3 gpurun_out/r02/ncu2.log
cat gpurun_out/r02/launches_vor_cells.csv | tail -12
