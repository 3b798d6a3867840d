"""TEST INFRASTRUCTURE ONLY (imported by tests/ and __graft_entry__.smoke(); never by simubox_b200/).

numpy / scipy restatement of the periodic domain-centre finder (csrc/seeds.cu).  The reference's own
seed finder (SimuBox/calculator/voronoi.py:31-68) is 2D-only OpenCV contour moments on a 3 x 3 tiled
uint8 image -- a different estimator (polygon moments, float32) -- so this oracle defines the voxel
semantics itself: parity unpinned against the reference; tests/test_gpu_seeds.py additionally checks
that on well-separated 2D blobs both estimators agree to a fraction of a pixel when cv2 is importable.
"""
import numpy as np


def domain_centres(field: np.ndarray, threshold: float, periodic: bool = True) -> dict:
    """Components of field > threshold under face connectivity (wrap-around when periodic), ordered by
    their smallest C-order voxel index; centroid = root + mean minimum-image integer displacement."""
    import scipy.ndimage as ndi

    fg = np.asarray(field) > threshold
    shape = fg.shape
    lab, n = ndi.label(fg)  # faces only (default structure), non-periodic
    parent = np.arange(n + 1)

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a

    if periodic:
        for ax in range(fg.ndim):
            if shape[ax] < 2:
                continue
            a = np.take(lab, 0, axis=ax).ravel()
            b = np.take(lab, shape[ax] - 1, axis=ax).ravel()
            for u, v in set(zip(a[(a > 0) & (b > 0)].tolist(), b[(a > 0) & (b > 0)].tolist())):
                ru, rv = find(u), find(v)
                if ru != rv:
                    parent[max(ru, rv)] = min(ru, rv)
    rootlab = np.array([find(a) for a in range(n + 1)])
    merged = rootlab[lab]
    flat = merged.ravel()
    idx = np.flatnonzero(flat > 0)
    # order components by their first voxel
    first = {}
    for i, l in zip(idx.tolist(), flat[idx].tolist()):
        if l not in first:
            first[l] = i
    order = sorted(first, key=first.get)
    remap = np.full(n + 1, -1, dtype=np.int64)
    for k, l in enumerate(order):
        remap[l] = k
    comp = np.where(merged > 0, remap[merged], -1).astype(np.int32)
    S = len(order)
    centres = np.zeros((S, fg.ndim))
    volumes = np.zeros(S, dtype=np.int64)
    coords = np.stack(np.unravel_index(idx, shape), axis=1)
    cid = comp.ravel()[idx]
    for k, l in enumerate(order):
        root = np.array(np.unravel_index(first[l], shape))
        d = coords[cid == k] - root
        if periodic:
            N = np.array(shape)
            h = N // 2
            d = (d + h) % N - h
        volumes[k] = len(d)
        c = root.astype(np.float64) + d.sum(axis=0).astype(np.float64) / float(len(d))
        if periodic:
            N = np.array(shape, dtype=np.float64)
            c = np.where(c < 0, c + N, np.where(c >= N, c - N, c))
        centres[k] = c
    return dict(centres=centres, volumes=volumes, labels=comp, n_components=S)
