"""Relation between the GPU seed finder / cell statistics and the reference's OpenCV centres / triangle coord_dict on
demos/datasets/voronoi (input and reference outputs from tests/golden).  Prints the statistics the f2 / f3 parity tests encode."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np

from simubox_b200 import VoronoiCell

g = np.load(Path(__file__).resolve().parent.parent / "tests/golden/analyze_voronoi_demo.npz")
tiled, ref_c = g["tiled"], g["centers"]          # (192, 384) field, 67 x (x = column, y = row) float32-precision centres
print("tiled", tiled.shape, "centres", ref_c.shape, ref_c[:3])
for thr in (127.0 / 255.0, 128.0 / 255.0, 0.5):
    dc = VoronoiCell.domain_centres(tiled, threshold=thr, periodic=False)
    c = dc["centres"][:, ::-1]  # (row, col) -> (x = col, y = row)
    d = np.sqrt(((ref_c[:, None, :] - c[None, :, :]) ** 2).sum(-1))
    print(f"thr {thr:.6f}: components {len(c)}; per reference centre nearest component distance: max {d.min(1).max():.3f} median {np.median(d.min(1)):.3f}"
          f"; components without a reference centre within 1 px: {(d.min(0) > 1).sum()}  volumes of those: {np.sort(dc['volumes'][d.min(0) > 1])[:10]}")
dc = VoronoiCell.domain_centres(tiled, threshold=127.0 / 255.0, periodic=False)
c = dc["centres"][:, ::-1]
d = np.sqrt(((ref_c[:, None, :] - c[None, :, :]) ** 2).sum(-1))
order = d.argmin(1)
print("matching unique:", len(set(order.tolist())) == len(order))
# worst offenders
bad = np.argsort(-d.min(1))[:8]
for b in bad:
    print("  ref", ref_c[b], "nearest comp", c[order[b]], "vol", dc["volumes"][order[b]], "dist", d.min(1)[b])
# f3: label the tiled image with the reference centres (x, y) -> grid axes (row, col) and compare adjacency with triangle coord_dict
pts = np.around(ref_c).astype(int)[:, ::-1].astype(float)  # (row, col) pixel units, rounded like Analyze does
lab = VoronoiCell.label_grid(pts, 0, shape=tiled.shape)
st = VoronoiCell.cell_statistics(lab, n_cells=len(pts), periodic=False)
tri = {}
for a, b in g["coord_pairs"]:
    tri.setdefault(int(a), set()).add(int(b))
grid = st["coord_dict"]
sub = sum(1 for k, v in tri.items() if v <= grid.get(k, set()))
eq = sum(1 for k, v in tri.items() if v == grid.get(k, set()))
extra = sum(len(grid.get(k, set()) - v) for k, v in tri.items())
missing = sum(len(v - grid.get(k, set())) for k, v in tri.items())
print(f"triangle coord_dict: {len(tri)} centres; subset of grid adjacency for {sub}; equal for {eq}; grid-only pairs {extra}; triangle-only pairs {missing}")
from scipy.spatial import Delaunay
dl = Delaunay(g["centers_lxly"])
full = {}
for t in dl.simplices:
    for i in range(3):
        full.setdefault(int(t[i]), set()).update(int(x) for x in t if x != t[i])
eq2 = sum(1 for k, v in full.items() if v == grid.get(k, set()))
miss2 = sum(len(v - grid.get(k, set())) for k, v in full.items())
extra2 = sum(len(grid.get(k, set()) - v) for k, v in full.items())
print(f"unfiltered Delaunay (physical coords): {len(full)} centres; equal to grid adjacency for {eq2}; Delaunay-only pairs {miss2}; grid-only pairs {extra2}")
dl2 = Delaunay(pts)
full2 = {}
for t in dl2.simplices:
    for i in range(3):
        full2.setdefault(int(t[i]), set()).update(int(x) for x in t if x != t[i])
eq3 = sum(1 for k, v in full2.items() if v == grid.get(k, set()))
miss3 = sum(len(v - grid.get(k, set())) for k, v in full2.items())
extra3 = sum(len(grid.get(k, set()) - v) for k, v in full2.items())
print(f"unfiltered Delaunay (pixel coords): equal for {eq3}; Delaunay-only pairs {miss3}; grid-only pairs {extra3}")
