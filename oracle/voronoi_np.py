"""CPU restatement of the reference grid-Voronoi labelling -- TEST INFRASTRUCTURE ONLY.

Follows ``/root/reference/SimuBox/calculator/voronoi.py:244-292``: per-pixel
``additive`` / ``power`` distance + ``np.argmin`` (first minimum = lowest seed index).
The reference has no minimum-image code; the periodic oracle is the one SURVEY.md 8(c)
defines: every seed replicated into its 3^d images ``p + m*L`` seed-major, reference
arithmetic, ``label = argmin // 3^d``.
"""
from __future__ import annotations

from itertools import product

import numpy as np


def additive(points, arr, weights):
    """``VoronoiCell.additive`` (voronoi.py:244-246)."""
    return np.linalg.norm(points - arr, axis=1) - weights


def power(points, arr, weights):
    """``VoronoiCell.power`` (voronoi.py:249-255) -- the seed INDEX is added to every
    coordinate (reference behaviour, kept)."""
    return np.linalg.norm(points - arr + np.arange(len(points))[:, np.newaxis], axis=1) ** 2 - weights


def image_offsets(box):
    """3^d periodic image offsets m*L, m in {-1,0,1}^d, in ``itertools.product`` order."""
    box = np.asarray(box, dtype=np.float64)
    return np.array([np.array(m) * box for m in product((-1.0, 0.0, 1.0), repeat=len(box))])


def label_voxels(points, weights, voxels, method="additive", box=None, chunk=4096):
    """Labels for an explicit list of integer voxel coordinates ``voxels (V,d)``.

    Vectorised over voxels but op-for-op the reference expression per (voxel, seed):
    ``sqrt(((p - x)**2 summed x->y->z)) - w`` (np.linalg.norm over a length-d axis is a
    sequential sum), then first-minimum argmin.
    """
    points = np.asarray(points, dtype=np.float64)
    S, d = points.shape
    w = weights
    M = 1
    if box is not None:
        offs = image_offsets(box)
        M = len(offs)
        points = (points[:, None, :] + offs[None, :, :]).reshape(S * M, d)
        if isinstance(w, np.ndarray):
            w = np.repeat(w, M)
    voxels = np.asarray(voxels)
    out = np.empty(len(voxels), dtype=np.int32)
    shift = (np.arange(len(points)) // M).astype(np.float64)[None, :]  # seed index (per image too)
    for s in range(0, len(voxels), chunk):
        x = voxels[s:s + chunk].astype(np.float64)
        acc = None
        for a in range(d):
            diff = points[None, :, a] - x[:, a, None]
            if method == "power":
                diff = diff + shift
            sq = diff * diff
            acc = sq if acc is None else acc + sq
        r = np.sqrt(acc)
        dis = (r * r if method == "power" else r) - w
        out[s:s + chunk] = np.argmin(dis, axis=1) // M
    return out


def label_grid(points, weights, size, method="additive", box=None, mask=None):
    """Label every voxel of ``product(range(size[0]), range(size[1]), ...)`` (voronoi.py:284).
    ``mask`` (uint8, same shape): voxels with mask == 0 get label -1 (drop-in extension,
    phin_with_mask-style last column)."""
    grids = np.meshgrid(*[np.arange(n) for n in size], indexing="ij")
    vox = np.stack([g.reshape(-1) for g in grids], axis=1)
    lab = label_voxels(points, weights, vox, method=method, box=box).reshape(size)
    if mask is not None:
        lab = np.where(np.asarray(mask) != 0, lab, -1).astype(np.int32)
    return lab


def label_grid_loop(points, weights, size, method="additive"):
    """The reference loop verbatim (one numpy call per pixel); small grids only."""
    func = additive if method == "additive" else power
    points = np.asarray(points, dtype=np.float64)
    out = np.empty(size, dtype=np.int32)
    for arr in product(*[range(n) for n in size]):
        out[arr] = np.argmin(func(points, np.array(arr, dtype=int), weights))
    return out
