"""Small exercise of the kernels written in round 2, for `compute-sanitizer --tool memcheck` (one tool, small cases):
cell-list labelling (2D / 3D, periodic / not, ragged grids, masks, plane ranges, overflow box scan), the allocation-free
entry point, the row-wise shell histogram (dense path, batched) and the chunked slab emulation."""
import ctypes as C, os, sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from simubox_b200 import Scatter, VoronoiCell, _lib, distributed as D, synth

rng = np.random.default_rng(3)
for shape, S, per in [((40, 24, 48), 30, True), ((33, 17, 29), 200, False), ((70, 90), 25, True), ((16, 16, 16), 900, True), ((9, 5, 3), 2, True)]:
    pts = rng.uniform(-0.1, 1.1, (S, len(shape))) * np.array(shape)
    box = tuple(float(n) for n in shape) if per else None
    mask = rng.random(shape) > 0.5
    for pl in ((0, shape[0]), (1, shape[0] - 1), (0, max(1, shape[0] // 2))):
        VoronoiCell.label_grid(pts, 0, shape=shape, box=box, mask=mask, planes=pl)
pts = np.concatenate([np.tile([[8.0, 8.0, 8.0]], (200, 1)), rng.uniform(0, 16, (20, 3))])
VoronoiCell.label_grid(pts, 0, shape=(16, 16, 16), planes=(0, 15))            # staging overflow -> box scan
d_pts = torch.from_numpy(rng.uniform(0, 32, (50, 3))).cuda()
lab = torch.empty((32, 32, 32), dtype=torch.int32, device="cuda")
_, pn = _lib.i64((32, 32, 32)); _k, pb = _lib.f64((32.0, 32.0, 32.0))
for _ in range(2):
    _lib.check(_lib.lib().sbx_voronoi_label_noalloc(0, 3, pn, C.c_void_p(d_pts.data_ptr()), 50, pb, None, C.c_void_p(lab.data_ptr()), 0, 32, None))
mats = np.stack([synth.field_numpy(i, (32, 32, 32)) for i in range(5)])
Scatter.sacttering_peak_batch(torch.from_numpy(mats).cuda(), synth.box_for(2), peaks=True, path=1, max_batch=5)
Scatter.sacttering_peak_batch(torch.from_numpy(rng.random((2, 20, 12, 14))).cuda(), (3.0, 2.0, 2.5), cutoff=40, path=1, max_batch=2)
for env in ({"SBX_SLAB_CHUNKS": "3"}, {"SBX_SLAB_CHUNKS": "8", "SBX_SLAB_ALLCOLS": "1"}):
    os.environ.update(env, SBX_SLAB_PATH="fft")
    D.slab_emulate(rng.random((24, 16, 12)), (2.0, 1.0, 1.5), 4, cutoff=30)
    D.slab_emulate(rng.random((32, 32, 32)), (8.0, 8.0, 8.0), 2, cutoff=400)
torch.cuda.synchronize()
print("sanitize exercise done")
