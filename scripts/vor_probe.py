"""Timing / checksum probe of one Voronoi labelling configuration (development tool; env knobs SBX_VOR_* are read
by the library at first use, so every variant runs in its own process):  python scripts/vor_probe.py [--grid 256]"""
import argparse, hashlib, json, os, sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from simubox_b200 import VoronoiCell, synth

ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=256)
ap.add_argument("--per-axis", type=int, default=16)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--nomask", action="store_true")
ap.add_argument("--tag", default="")
a = ap.parse_args()
n = a.grid
shape, box = (n, n, n), (float(n),) * 3
dev = torch.device("cuda", 0)
pts = torch.from_numpy(synth.seeds_lattice(shape, a.per_axis, 4.0, 1234)).to(dev)
mask = None if a.nomask else (synth.fields_torch(1, 1, shape, cells=4, device=dev)[0] > 0.5).to(torch.uint8).contiguous()
flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
for _ in range(3):
    flush.zero_()
    lab = VoronoiCell.label_grid(pts, 0, shape=shape, box=box, mask=mask, return_tensor=True)
pairs = []
for _ in range(a.steps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    lab = VoronoiCell.label_grid(pts, 0, shape=shape, box=box, mask=mask, return_tensor=True)
    e1.record()
    pairs.append((e0, e1))
torch.cuda.synchronize()
t = sorted(x.elapsed_time(y) for x, y in pairs)
lab, st = VoronoiCell.label_grid(pts, 0, shape=shape, box=box, mask=mask, return_tensor=True, stats=True)
live = int(mask.sum().item()) if mask is not None else n ** 3
h = hashlib.sha1(lab.cpu().numpy().tobytes()).hexdigest()[:16]
env = {k: v for k, v in os.environ.items() if k.startswith("SBX_")}
print(json.dumps({"tag": a.tag, "env": env, "grid": n, "S": a.per_axis ** 3, "ms_mean": sum(t) / len(t), "ms_median": t[len(t) // 2], "ms_min": t[0],
                  "evals_per_live_voxel": st["evaluations"] / max(live, 1), "tiles": st["tiles"], "overflow": st["overflow_tiles"], "sha1": h}))
