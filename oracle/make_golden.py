"""Generate ``tests/golden/*.npz`` by running the LIVE reference in this container.

Test infrastructure only.  Run as ``python -m oracle.make_golden`` from the repo root (needs
``/root/reference``).  Every fixture stores inputs and the reference's outputs so that the
GPU box (no reference there) can check the CUDA path and the numpy restatement against them.
While generating, the numpy restatement (oracle/scatter_np.py, oracle/voronoi_np.py) is
checked against the live outputs -- a mismatch aborts.  Version skew vs the reference's lock
file is recorded in ``versions.json``.
"""
from __future__ import annotations

import json
import sys
import warnings
from pathlib import Path

import numpy as np

from oracle import ref_live, scatter_np, voronoi_np
from simubox_b200 import synth

OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"
REF = Path(ref_live.REF_ROOT)


def pack_qidx(q_idx: dict):
    """dict shell -> list of d-vectors  =>  (shell_ids, vectors) arrays."""
    ids, vecs = [], []
    for s in sorted(q_idx):
        for v in q_idx[s]:
            ids.append(s)
            vecs.append(np.asarray(v, dtype=np.float64))
    return np.asarray(ids, dtype=np.int32), np.stack(vecs)


def check_restatement(mat, L, cutoff, threshold, qI_ref, qidx_ref):
    """numpy restatements vs live reference: q exact, counts exact, I to 1e-9 (rows < cutoff)."""
    for fn in (scatter_np.scattering_peak, scatter_np.scattering_peak_table):
        qI, qidx = fn(mat, L, cutoff, threshold)
        assert qI.shape == qI_ref.shape, (fn.__name__, qI.shape, qI_ref.shape)
        assert np.array_equal(qI[:, 0], qI_ref[:, 0]), fn.__name__
        n = len(qI)
        full = n - 1 if n == cutoff + 1 else n
        tol = 1e-9 * np.abs(qI_ref[:full, 1]) + 1e-12 * np.abs(qI_ref[:, 1]).max()
        assert np.all(np.abs(qI[:full, 1] - qI_ref[:full, 1]) <= tol), fn.__name__
        for s in range(n):
            assert len(qidx[s]) == len(qidx_ref[s]), (fn.__name__, s)


def scatter_case(name, mat, L, cutoff=30, threshold=1e-12, peak_kw=None, store_input=True, extra=None):
    qI, qidx, res = ref_live.ref_scatter(mat, L, cutoff=cutoff, threshold=threshold)
    check_restatement(mat, L, cutoff, threshold, qI, qidx)
    ids, vecs = pack_qidx(qidx)
    data = dict(shape=np.array(mat.shape), L=np.asarray(L, dtype=np.float64), cutoff=cutoff, threshold=threshold,
                q_Intensity=qI, qidx_shell=ids, qidx_vec=vecs)
    if store_input:
        data["mat"] = np.asarray(mat, dtype=np.float64)
    if peak_kw is not None:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            sp = ref_live.ref_show_peak(res, **peak_kw)
        mine = scatter_np.show_peak(qI, **peak_kw)
        assert np.array_equal(mine["plot_y"], sp.plot_y) and np.array_equal(mine["plot_x"], sp.plot_x)
        assert np.array_equal(mine["peaks_location"], sp.peaks_location)
        assert np.array_equal(scatter_np.find_peaks_height(sp.plot_y, peak_kw.get("min_height", 1)), mine["peaks_index"])
        data.update(peak_kw=json.dumps(peak_kw), plot_y=sp.plot_y, plot_x=sp.plot_x, peaks_index=mine["peaks_index"],
                    peaks_location=sp.peaks_location, peak_heights=sp.peaks_height["peak_heights"])
    if extra:
        data.update(extra)
    np.savez_compressed(OUT / f"{name}.npz", **data)
    print(f"  {name}: shape {mat.shape} rows {len(qI)}" + (f" peaks {data['peaks_index']}" if peak_kw else ""))


def gen_scatter():
    sb = ref_live.load()
    rd, rp = sb.toolkit.read_density, sb.toolkit.read_printout
    pk = dict(width=0.5, step=2000, cutoff=300.0, min_height=1)
    # --- the notebook's own scatter demo (demos/scripts.ipynb cell 17): phout_C4, with printout repair
    for tag in ("C4", "C6"):
        d = rd(REF / f"demos/datasets/scatter/phout_{tag}.txt")
        po = rp(REF / f"demos/datasets/scatter/printout_{tag}.txt")
        L1 = d.lxlylz.copy()
        d.repair_from_printout(po)
        for target in (0, 1):
            mat = d.data[target].values.reshape(d.shape)
            scatter_case(f"scatter_phout{tag}_t{target}_repaired", mat, d.lxlylz, cutoff=300, peak_kw=pk,
                         extra=dict(source=f"demos/datasets/scatter/phout_{tag}.txt", target=target))
        mat = d.data[0].values.reshape(d.shape)
        scatter_case(f"scatter_phout{tag}_t0_unitbox", mat, L1, cutoff=30, peak_kw=dict(width=0.5, step=2000, cutoff=20, min_height=1))
    # --- non-cubic 3D demo field (pins quirk Q1): demos/datasets/voronoi 2x64x128
    d = rd(REF / "demos/datasets/voronoi/phout.txt")
    po = rp(REF / "demos/datasets/voronoi/printout.txt")
    d.repair_from_printout(po)
    scatter_case("scatter_voronoi_demo_2x64x128", d.data[0].values.reshape(d.shape), d.lxlylz, cutoff=30,
                 extra=dict(source="demos/datasets/voronoi/phout.txt", target=0))
    # --- masked field (phin_with_mask stand-in)
    d = rd(REF / "work/桥连构型比率计算/BABAB线型五嵌段共聚物/phin_with_mask.txt")
    scatter_case("scatter_phin_with_mask_t0", d.data[0].values.reshape(d.shape), d.lxlylz, cutoff=30,
                 extra=dict(mask=d.data[4].values.reshape(d.shape).astype(np.uint8)))
    # --- random non-cubic / unequal-L grids (Q1), 1D/2D/3D, odd thresholds and cutoffs
    rng = np.random.default_rng(20240611)
    for shape, L, cutoff, thr in [((8, 12, 10), (1.0, 2.0, 3.0), 10, 1e-12), ((12, 8, 6), (2.0, 1.5, 3.0), 30, 1e-12),
                                  ((10, 10, 10), (1.0, 1.3, 1.7), 30, 1e-12), ((12, 16), (1.0, 2.5), 25, 1e-12),
                                  ((16, 16), (3.0, 3.0), 1000, 1e-12), ((32,), (5.0,), 30, 1e-12),
                                  ((16, 16, 16), (4.0, 4.0, 4.0), 5000, 1e-12), ((16, 16, 16), (4.0, 4.0, 4.0), 12, 0.3),
                                  ((16, 16, 16), (4.0, 4.0, 4.0), 0, 1e-12), ((6, 4, 2), (1.0, 1.0, 1.0), 30, 1e-12)]:
        tag = "x".join(map(str, shape)) + f"_c{cutoff}" + ("" if thr == 1e-12 else f"_thr{thr}")
        scatter_case(f"scatter_random_{tag}", rng.random(shape), L, cutoff=cutoff, threshold=thr)
    # --- synthetic gyroid / BCC (SURVEY.md 8(d) config 2 generator), inputs regenerated in the tests
    pk2 = dict(width=0.5, step=2000, cutoff=20, min_height=1)
    for n in (16, 32, 64):
        for idx in (0, 1):
            mat = synth.field_numpy(idx, (n, n, n), cells=2)
            scatter_case(f"scatter_synth_{n}_f{idx}", mat, synth.box_for(2), cutoff=30, peak_kw=pk2,
                         store_input=(n <= 32), extra=dict(field_index=idx, cells=2, checksum=float(mat.sum())))


def gen_scatter_baseline():
    """BASELINE.json configs 2 and 4 at their full grid sizes through the LIVE reference: 128^3 fields 0-7 (cells = 2,
    SURVEY.md 8(d) config 2 'parity subset') and two 256^3 fields (cells = 4).  Inputs are regenerated in the tests
    from the closed-form generator (checksum stored); a 256^3 reference run needs ~5 GB and ~15 s."""
    pk2 = dict(width=0.5, step=2000, cutoff=20, min_height=1)
    for n, cells, fields in ((128, 2, range(8)), (256, 4, range(2))):
        for idx in fields:
            mat = synth.field_numpy(idx, (n, n, n), cells=cells)
            scatter_case(f"scatter_synth_{n}_f{idx}", mat, synth.box_for(cells), cutoff=30, peak_kw=pk2, store_input=False,
                         extra=dict(field_index=idx, cells=cells, checksum=float(mat.sum())))


def demo_density(sb_or_mod, mat, NxNyNz, lxlylz):
    """A ``Density`` holding one column, the way ``read_density`` builds it (reader.py:218-230)."""
    import pandas as pd

    N = np.asarray(NxNyNz)
    return sb_or_mod.Density(path=None, data=pd.DataFrame(np.asarray(mat).reshape(-1, 1)), NxNyNz=N, lxlylz=np.asarray(lxlylz, dtype=np.float64)[N != 1],
                             shape=N[N != 1])


def gen_analyze():
    """``VoronoiCell.OpenCV`` / ``Analyze`` (VORONOI and TRIANGLE modes) of the live reference on the notebook's own
    example (demos/scripts.ipynb cell 28: demos/datasets/voronoi, slices=(0, 0), expand=3): 67 centres, Qhull
    vertices and ridges, filtered Delaunay triangles, coord_dict.  The input field is the one already stored in
    scatter_voronoi_demo_2x64x128.npz."""
    sb = ref_live.load()
    d = sb.toolkit.read_density(REF / "demos/datasets/voronoi/phout.txt")
    d.repair_from_printout(sb.toolkit.read_printout(REF / "demos/datasets/voronoi/printout.txt"))
    stored = np.load(OUT / "scatter_voronoi_demo_2x64x128.npz")
    assert np.array_equal(stored["mat"], d.data[0].values.reshape(d.shape)) and np.array_equal(stored["L"], d.lxlylz)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        tri = sb.calculator.VoronoiCell.Analyze(d, sb.schema.AnalyzeMode.TRIANGLE, slices=(0, 0), expand=3, interactive=False)
        tri60 = sb.calculator.VoronoiCell.Analyze(d, sb.schema.AnalyzeMode.TRIANGLE, slices=(0, 0), expand=3, interactive=False, theta_max=np.pi / 3 * 1.2)
        vor = sb.calculator.VoronoiCell.Analyze(d, sb.schema.AnalyzeMode.VORONOI, slices=(0, 0), expand=3, interactive=False)
    cv = tri.cv_result
    pd_ = cv.parsed_density
    vecs = [np.linspace(0, l, N) for l, N in zip(pd_.lxlylz, pd_.NxNyNz)][::-1]
    X, Y = np.meshgrid(*vecs)
    c = np.around(cv.centers).astype(int)
    centers_lxly = np.stack([X[c[:, 1], c[:, 0]], Y[c[:, 1], c[:, 0]]], axis=1)
    assert np.array_equal(vor.voronoi.points, centers_lxly)
    keys = sorted(int(k) for k in tri.coord_dict)
    pairs = np.array([(k, int(v)) for k in keys for v in sorted(tri.coord_dict[k])], dtype=np.int32)
    np.savez_compressed(OUT / "analyze_voronoi_demo.npz", centers=cv.centers, centers_lxly=centers_lxly, tiled=pd_.mat.squeeze(),
                        tiled_lxlylz=pd_.lxlylz, triangles=np.array(tri.triangle, dtype=np.int32), coord_pairs=pairs,
                        triangles_theta72=np.array(tri60.triangle, dtype=np.int32),
                        vor_vertices=vor.voronoi.vertices, vor_ridge_points=vor.voronoi.ridge_points.astype(np.int32),
                        point_cut=sb.calculator.VoronoiCell.set_point_limits(cv, centers_lxly))
    print(f"  analyze_voronoi_demo: {len(cv.centers)} centres, {len(tri.triangle)} triangles ({len(tri60.triangle)} at 72 deg), "
          f"{len(vor.voronoi.vertices)} Voronoi vertices")


def gen_voronoi():
    rng = np.random.default_rng(77)
    cases = []
    pts = rng.uniform(0, 1, (7, 2)) * [40, 30]
    cases.append(("vor2d_additive_w", pts, rng.uniform(0, 3, 7), (40, 30), "additive", None))
    cases.append(("vor2d_additive_0", pts, 0, (40, 30), "additive", None))
    cases.append(("vor2d_power_w", pts, rng.uniform(0, 50, 7), (40, 30), "power", None))
    cases.append(("vor2d_notebook", np.array([[10., 35.], [10., 10.], [35., 35.], [35., 10.]]), np.array([1e1, 0, 0, 0]),
                  (50, 50), "additive", None))
    pts3 = rng.uniform(0, 1, (9, 3)) * [12, 10, 8]
    cases.append(("vor3d_additive_0", pts3, 0, (12, 10, 8), "additive", None))
    cases.append(("vor3d_additive_w", pts3, rng.uniform(0, 2, 9), (12, 10, 8), "additive", None))
    cases.append(("vor3d_power_0", pts3, 0, (12, 10, 8), "power", None))
    # periodic (oracle = 3^d replicated images through the reference's additive + argmin)
    cases.append(("vor2d_periodic", pts, 0, (40, 30), "additive", (40.0, 30.0)))
    cases.append(("vor3d_periodic", pts3, 0, (12, 10, 8), "additive", (12.0, 10.0, 8.0)))
    cases.append(("vor3d_periodic_w", pts3, rng.uniform(0, 2, 9), (12, 10, 8), "additive", (12.0, 10.0, 8.0)))
    # exact ties: unjittered lattice, seeds mirrored about voxel planes
    lat = synth.seeds_lattice((16, 16, 16), per_axis=2, jitter=0.0)
    cases.append(("vor3d_ties_periodic", lat, 0, (16, 16, 16), "additive", (16.0, 16.0, 16.0)))
    cases.append(("vor3d_ties", lat, 0, (16, 16, 16), "additive", None))
    lat2 = synth.seeds_lattice((24, 24), per_axis=3, jitter=0.0)
    cases.append(("vor2d_ties_periodic", lat2, 0, (24, 24), "additive", (24.0, 24.0)))
    for name, p, w, size, method, box in cases:
        images = voronoi_np.image_offsets(box) if box is not None else None
        lab = ref_live.ref_label(p, w, size, method=method, images=images)
        mine = voronoi_np.label_grid(p, w, size, method=method, box=box)
        assert np.array_equal(lab, mine), name
        if box is None:
            assert np.array_equal(lab, voronoi_np.label_grid_loop(p, w, size, method=method)), name
        np.savez_compressed(OUT / f"{name}.npz", points=p, weights=np.asarray(w, dtype=np.float64), size=np.array(size),
                            method=method, box=np.array(box if box is not None else []), labels=lab)
        print(f"  {name}: {size} seeds {len(p)} labels used {len(np.unique(lab))}")


def gen_reader():
    """Reference read_density / parse_density outputs on small files written here (text + binary)."""
    sb = ref_live.load()
    rd, parse = sb.toolkit.read_density, sb.toolkit.parse_density
    rng = np.random.default_rng(5)
    tmp = OUT / "files"
    tmp.mkdir(exist_ok=True)
    shape = (4, 6, 8)
    table = rng.random((int(np.prod(shape)), 5))
    table[:, 4] = (table[:, 0] > 0.5)
    txt = tmp / "phin_with_mask_4x6x8.txt"
    with open(txt, "w") as fh:
        fh.write("4 6 8\n")
        fh.write("2.500000 3.000000 4.250000\n")
        for row in table:
            fh.write(" ".join(f"{v:.6f}" for v in row[:4]) + f" {row[4]:.6f} \n")
    txt2 = tmp / "phout_2d_12x10.txt"
    t2 = rng.random((120, 3))
    with open(txt2, "w") as fh:
        fh.write("1 12 10\n")
        for row in t2:
            fh.write(" ".join(f"{v:.9e}" for v in row) + "\n")
    binp = tmp / "block_6x4x2.bin"
    tb = rng.random((48, 3))
    with open(binp, "wb") as fh:
        fh.write(np.array([6, 4, 2], dtype=np.int32).tobytes())
        fh.write(np.array([1.5, 2.5, 3.5]).tobytes())
        fh.write(np.asfortranarray(tb).tobytes(order="F"))
    out = {}
    for key, path, kw in [("txt3d", txt, dict(parse_L=True)), ("txt2d", txt2, {}), ("bin", binp, dict(parse_L=True))]:
        d = rd(path, **kw)
        out[f"{key}_data"] = d.data.values
        out[f"{key}_NxNyNz"] = d.NxNyNz
        out[f"{key}_lxlylz"] = d.lxlylz
        out[f"{key}_shape"] = d.shape
        variants = {"plain": {}, "expand3": dict(expand=3), "t1": dict(target=1),
                    "all": dict(target="all"), "expand_vec": dict(expand=[1, 2] if len(d.shape) == 2 else [1, 2, 1.5])}
        if len(d.shape) == 3:
            variants["slice"] = dict(slices=(1, 0))
            variants["permute"] = dict(permute=[2, 0, 1])
        for vname, vkw in variants.items():
            p = parse(d, **vkw)
            for f in ("mat", "NxNyNz", "lxlylz", "raw_NxNyNz", "raw_lxlylz", "expand"):
                out[f"{key}_{vname}_{f}"] = np.asarray(getattr(p, f))
            out[f"{key}_{vname}_target"] = np.asarray(p.target)
    np.savez_compressed(OUT / "reader_cases.npz", **out)
    print("  reader_cases:", len(out), "arrays")
    # printout tail
    po = sb.toolkit.read_printout(REF / "demos/datasets/scatter/printout_C4.txt")
    np.savez_compressed(OUT / "printout_C4.npz", box=po.box, lxlylz=po.lxlylz, step=po.step, freeEnergy=po.freeEnergy,
                        inCompMax=po.inCompMax)


def write_snapshot(path, xyz, types, L):
    """Minimal HOOMD/GALAMOST-style xml the reference's XmlTransformer reads (xmltrans.py:25-61)."""
    body = ["<?xml version=\"1.0\" encoding=\"UTF-8\"?>", "<galamost_xml version=\"1.3\">",
            f"<configuration time_step=\"0\" dimensions=\"3\" natoms=\"{len(xyz)}\" >",
            f"<box lx=\"{float(L[0])!r}\" ly=\"{float(L[1])!r}\" lz=\"{float(L[2])!r}\"/>", f"<position num=\"{len(xyz)}\">"]
    body += [" ".join(repr(float(v)) for v in p) for p in xyz]
    body += ["</position>", f"<type num=\"{len(xyz)}\">"] + list(types) + ["</type>", "</configuration>", "</galamost_xml>"]
    Path(path).write_text("\n".join(body) + "\n")


def gen_splat():
    """XmlTransformer.parse_xml + transform of the live reference on small synthetic snapshots."""
    from oracle import splat_np

    sb = ref_live.load()
    files = OUT / "files"
    rng = np.random.default_rng(2024)
    cases = {"splat_cubic": dict(L=(8.0, 8.0, 8.0), P=300, N=None, r_cut=2.0),
             "splat_noncubic_fine": dict(L=(6.5, 9.0, 5.2), P=200, N=(12, 10, 14), r_cut=1.3),
             "splat_coarse": dict(L=(10.0, 10.0, 12.0), P=150, N=(6, 4, 8), r_cut=2.5)}
    for name, c in cases.items():
        L = np.array(c["L"])
        xyz = (rng.random((c["P"], 3)) - 0.5) * L
        xyz[:6] = [[-L[0] / 2, 0, 0], [L[0] / 2 - 1e-9, 0.1, 0.2], [0, -L[1] / 2, L[2] / 2 - 1e-12], [0.0, 0.0, 0.0],
                   [-L[0] / 2 + 0.5, -L[1] / 2 + 0.5, -L[2] / 2 + 0.5], [L[0] / 4, L[1] / 4, L[2] / 4]]
        # digits in type names only work with merge=False (with merge=True the reference maps the stripped names and
        # leaves "A1" unmapped -> NaN index -> IndexError in transform)
        merge = name != "splat_coarse"
        types = rng.choice(["A", "B", "C"] if merge else ["A", "B", "A1", "C2"], c["P"]).tolist()
        write_snapshot(files / f"{name}.xml", xyz, types, L)
        tr = sb.toolkit.xmltrans.XmlTransformer(files / f"{name}.xml")
        x = tr.parse_xml(NxNyNz=c["N"], merge=merge)
        res = sb.toolkit.xmltrans.XmlTransformer.transform(x, r_cut=c["r_cut"])
        mine = splat_np.transform(x.data[["x", "y", "z"]].values, x.data["index"].values, len(x.atoms_type), x.NxNyNz, x.lxlylz, c["r_cut"])
        assert np.array_equal(mine, res.phi), name
        np.savez_compressed(OUT / f"{name}.npz", phi=res.phi, NxNyNz=x.NxNyNz, lxlylz=x.lxlylz, r_cut=c["r_cut"],
                            index=x.data["index"].values, xyz=x.data[["x", "y", "z"]].values, atoms_type=np.array(x.atoms_type), merge=merge)
        print(f"  {name}: N {x.NxNyNz} types {x.atoms_type} total {res.phi.sum():.0f}")


def main():
    if not ref_live.available():
        sys.exit("live reference not present; goldens can only be generated in the build container")
    OUT.mkdir(parents=True, exist_ok=True)
    import cv2, pandas, pydantic, scipy
    (OUT / "versions.json").write_text(json.dumps(dict(
        numpy=np.__version__, scipy=scipy.__version__, pandas=pandas.__version__, cv2=cv2.__version__,
        pydantic=pydantic.__version__, python=sys.version.split()[0],
        pinned_by_reference=dict(numpy="1.26.4", scipy="1.11.4", pandas="2.1.4", opencv="4.8.1.78")), indent=1))
    only = set(sys.argv[1:])
    if only:  # e.g. python -m oracle.make_golden baseline analyze  (adds fixtures without rewriting the others)
        if "baseline" in only:
            print("scatter, BASELINE grid sizes:"); gen_scatter_baseline()
        if "analyze" in only:
            print("analyze:"); gen_analyze()
        return
    print("scatter:"); gen_scatter()
    print("scatter, BASELINE grid sizes:"); gen_scatter_baseline()
    print("analyze:"); gen_analyze()
    print("voronoi:"); gen_voronoi()
    print("reader:"); gen_reader()
    print("splat:"); gen_splat()


if __name__ == "__main__":
    main()
