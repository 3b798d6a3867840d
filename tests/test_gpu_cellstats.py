"""GPU parity suite for the per-cell statistics kernel (csrc/cellstats.cu) against oracle/cellstats_np.py:
volumes and the adjacency graph exact, masses within 1e-12 relative (fp64 atomics, summation order not fixed)."""
import numpy as np
import pytest

from oracle import cellstats_np

pytestmark = pytest.mark.gpu


def check(got, want, S):
    assert np.array_equal(got["volume"], want["volume"])
    if "mass" in want:
        scale = np.abs(want["mass"]).max() + 1.0
        assert np.allclose(got["mass"], want["mass"], rtol=1e-12, atol=1e-12 * scale)
    assert np.array_equal(got["degree"], want["degree"])
    adj = np.zeros((S, S), dtype=bool)
    for s in range(S):
        adj[s, got["indices"][got["indptr"][s]:got["indptr"][s + 1]]] = True
    assert np.array_equal(adj, want["adjacency"])
    assert got["coord_dict"] == {s: set(np.flatnonzero(want["adjacency"][s]).tolist()) for s in range(S) if want["adjacency"][s].any()}


@pytest.mark.parametrize("shape,S,periodic", [((8, 8, 8), 5, True), ((6, 10, 24), 40, False), ((5, 7, 13), 9, True),
                                              ((1, 9, 16), 4, True), ((12, 20), 7, True), ((33, 31), 50, False),
                                              ((16, 16, 64), 300, True), ((2, 2, 2), 3, True)])
def test_random_labels_vs_oracle(sbx_lib, shape, S, periodic):
    """Incoherent labels (every voxel its own run): exercises the in-thread middle segments, ragged rows
    (N2 % 8 != 0, scalar loads), masked voxels (-1) and labels beyond n_cells (ignored)."""
    from simubox_b200 import VoronoiCell

    rng = np.random.default_rng(sum(shape) + S)
    lab = rng.integers(-1, S + 2, size=shape).astype(np.int32)
    field = rng.standard_normal(shape)
    got = VoronoiCell.cell_statistics(lab, field, n_cells=S, periodic=periodic)
    check(got, cellstats_np.cell_statistics(lab, field, n_cells=S, periodic=periodic), S)


@pytest.mark.parametrize("shape,S", [((48, 48, 48), 150), ((96, 80), 60)])
def test_voronoi_labels_vs_oracle(sbx_lib, shape, S):
    from simubox_b200 import VoronoiCell

    rng = np.random.default_rng(S)
    pts = rng.random((S, len(shape))) * np.array(shape)
    mask = rng.random(shape) > 0.1
    lab = VoronoiCell.label_grid(pts, 0, shape=shape, box=tuple(float(n) for n in shape), mask=mask)
    field = rng.random(shape)
    got = VoronoiCell.cell_statistics(lab, field, n_cells=S)
    check(got, cellstats_np.cell_statistics(lab, field, n_cells=S), S)
    no_graph = VoronoiCell.cell_statistics(lab, n_cells=S, neighbours=False)
    assert np.array_equal(no_graph["volume"], got["volume"]) and "degree" not in no_graph and "mass" not in no_graph


def test_properties_at_baseline_size(sbx_lib):
    """256^3 / 4096 seeds (BASELINE config 3): volumes add up to the live voxels, masses to the masked field sum,
    the graph is symmetric and has the
    coordination a 3D Voronoi tessellation must have (mean face count of a jittered lattice: 12..20, at least 4 per cell)."""
    import torch

    from simubox_b200 import VoronoiCell, synth

    shape = (256, 256, 256)
    pts = synth.seeds_lattice(shape, 16, 4.0, 99)
    field = synth.fields_torch(1, 1, shape, cells=4)[0]
    mask = (field > 0.2).to(torch.uint8)
    lab = VoronoiCell.label_grid(pts, 0, shape=shape, box=(256.0,) * 3, mask=mask, return_tensor=True)
    got = VoronoiCell.cell_statistics(lab, field, n_cells=4096)
    live = int(mask.sum().item())
    assert int(got["volume"].sum()) == live
    want_mass = float(field[mask.bool()].sum().item())
    assert abs(got["mass"].sum() - want_mass) <= 1e-10 * abs(want_mass)
    S = 4096
    adj = np.zeros((S, S), dtype=bool)
    rows = np.repeat(np.arange(S), got["degree"])
    adj[rows, got["indices"]] = True
    assert np.array_equal(adj, adj.T) and not adj.diagonal().any()
    full = VoronoiCell.cell_statistics(VoronoiCell.label_grid(pts, 0, shape=shape, box=(256.0,) * 3, return_tensor=True), n_cells=S)
    assert int(full["volume"].sum()) == 256 ** 3
    assert 12 <= full["degree"].mean() <= 20 and full["degree"].min() >= 4


def test_adjacency_vs_delaunay_of_reference_centres(sbx_lib):
    """f3 pin on the notebook example (67 OpenCV centres of demos/datasets/voronoi, tests/golden/analyze_voronoi_demo.npz).
    ``cell_statistics`` returns the FACE adjacency of the labelled cells inside the image; the reference's ``triangle``
    (voronoi.py:163-182) returns Delaunay neighbours of the centres in physical coordinates after dropping obtuse
    triangles.  Voronoi face adjacency is the Delaunay graph of the same points in the same (pixel) metric, clipped to
    the window, so the relation that must hold is: grid adjacency is a subset of the unfiltered pixel-metric Delaunay
    graph, the difference being pairs whose common face lies outside the 192 x 384 window (hull edges).  Measured: 53
    of 67 cells identical, 0 grid-only pairs, 16 Delaunay-only directed pairs.  Against the reference's filtered,
    physical-metric coord_dict it is a different definition (41 of 63 cells identical) -- recorded, not asserted."""
    from conftest import golden
    from scipy.spatial import Delaunay
    from simubox_b200 import VoronoiCell

    g = golden("analyze_voronoi_demo")
    pts = np.around(g["centers"]).astype(int)[:, ::-1].astype(float)   # rounded like Analyze, (row, col)
    lab = VoronoiCell.label_grid(pts, 0, shape=g["tiled"].shape)
    grid = VoronoiCell.cell_statistics(lab, n_cells=len(pts), periodic=False)["coord_dict"]
    dl = {}
    for t in Delaunay(pts).simplices:
        for i in range(3):
            dl.setdefault(int(t[i]), set()).update(int(x) for x in t if x != t[i])
    assert all(grid.get(k, set()) <= dl[k] for k in dl), "two cells share a face but their seeds are not Delaunay neighbours"
    assert sum(1 for k in dl if grid.get(k, set()) == dl[k]) >= 50
    assert sum(len(dl[k] - grid.get(k, set())) for k in dl) <= 20
