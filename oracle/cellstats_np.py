"""TEST INFRASTRUCTURE ONLY (imported by tests/ and __graft_entry__.smoke(); never by simubox_b200/).

numpy restatement of the per-cell statistics kernel (csrc/cellstats.cu).  The reference has no label
array -- it builds coord_dict from Delaunay triangles (SimuBox/calculator/voronoi.py:150-213) and
in-cell densities with a point-in-polygon loop (demos/scripts.ipynb cells 1-2) -- so this oracle
defines the grid semantics itself: parity unpinned against the reference.
"""
import numpy as np


def cell_statistics(labels: np.ndarray, field=None, n_cells=None, periodic: bool = True) -> dict:
    labels = np.asarray(labels)
    S = int(labels.max()) + 1 if n_cells is None else int(n_cells)
    ok = (labels >= 0) & (labels < S)
    volume = np.bincount(labels[ok], minlength=S).astype(np.int64)
    out = dict(volume=volume)
    if field is not None:
        out["mass"] = np.bincount(labels[ok], weights=np.asarray(field, dtype=np.float64)[ok], minlength=S)
    adj = np.zeros((S, S), dtype=bool)
    for ax in range(labels.ndim):
        n = labels.shape[ax]
        if n < 2:
            continue
        a = labels if periodic else np.take(labels, np.arange(n - 1), axis=ax)
        b = np.roll(labels, -1, axis=ax) if periodic else np.take(labels, np.arange(1, n), axis=ax)
        m = (a != b) & (a >= 0) & (b >= 0) & (a < S) & (b < S)
        adj[a[m], b[m]] = True
        adj[b[m], a[m]] = True
    out["adjacency"] = adj
    out["degree"] = adj.sum(axis=1).astype(np.int32)
    return out
