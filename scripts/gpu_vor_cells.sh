#!/bin/bash
# r02: cell-list labelling path: parity suite, timing of the knob variants, launch list + full ncu capture
set -x
mkdir -p gpurun_out/r02
python -m pytest tests/test_gpu_voronoi.py -x -q 2>&1 | tail -15 > gpurun_out/r02/vor_pytest.txt
cat gpurun_out/r02/vor_pytest.txt
P=gpurun_out/r02/vor_probe.jsonl
: > $P
python scripts/vor_probe.py --tag default >> $P
SBX_VOR_MINB=4 python scripts/vor_probe.py --tag minb4 >> $P
SBX_VOR_MINB=5 python scripts/vor_probe.py --tag minb5 >> $P
SBX_VOR_MINB=8 python scripts/vor_probe.py --tag minb8 >> $P
SBX_VOR_CELL=3 python scripts/vor_probe.py --tag cell8 >> $P
SBX_VOR_R0=100 python scripts/vor_probe.py --tag r0-100 >> $P
python scripts/vor_probe.py --nomask --tag nomask >> $P
python scripts/vor_probe.py --grid 512 --per-axis 32 --tag 512 >> $P
cat $P
python scripts/vor_probe.py --steps 3 > gpurun_out/r02/plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'cell|label' -s 9 -c 9 --csv --log-file gpurun_out/r02/launches_vor_cells.csv python scripts/vor_probe.py --steps 3 > gpurun_out/r02/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'label_cells|cell_count|cell_scatter' -s 9 -c 3 -o gpurun_out/r02/vor_cells python scripts/vor_probe.py --steps 3 > gpurun_out/r02/ncu2.log 2>&1
grep -E "cell|label" gpurun_out/r02/launches_vor_cells.csv | awk -F'","' '{print $5, $NF}' | tail -3
