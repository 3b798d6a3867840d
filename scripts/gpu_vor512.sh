#!/bin/bash
# r02: whole-grid labelling, tile path vs cell path over seed densities and grid sizes (basis of the automatic path choice)
mkdir -p gpurun_out/r02
python -m pytest tests/test_gpu_voronoi.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4
P=gpurun_out/r02/voronoi_paths.jsonl; : > $P
for cfg in "256 8" "256 16" "256 24" "256 32" "512 32" "512 16" "128 16"; do
  set -- $cfg
  SBX_VOR_PATH=tiles python scripts/vor_probe.py --grid $1 --per-axis $2 --steps 10 --tag tiles >> $P
  SBX_VOR_PATH=cells python scripts/vor_probe.py --grid $1 --per-axis $2 --steps 10 --tag cells >> $P
done
python - <<'PY'
import json
for l in open("gpurun_out/r02/voronoi_paths.jsonl"):
    d=json.loads(l); print(d["tag"], d["grid"], d["S"], "ms", round(d["ms_median"],4), "evals", round(d["evals_per_live_voxel"],2), "over", d["overflow"], d["sha1"])
PY
