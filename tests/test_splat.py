"""Particle -> grid splat (XmlTransformer.parse_xml / transform, reference toolkit/xmltrans.py:63-153):
CPU part pins the numpy oracle and the host-side xml parsing against goldens of the live reference;
the GPU part checks the CUDA kernel bit for bit against those goldens and against the oracle."""
import numpy as np
import pytest

from conftest import GOLDEN, golden
from oracle import splat_np

CASES = ["splat_cubic", "splat_noncubic_fine", "splat_coarse"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_and_parse_xml_match_reference_golden(name):
    from simubox_b200 import XmlTransformer

    g = golden(name)
    phi = splat_np.transform(g["xyz"], g["index"], len(g["atoms_type"]), g["NxNyNz"], g["lxlylz"], float(g["r_cut"]))
    assert np.array_equal(phi, g["phi"])
    grid = None if name == "splat_cubic" else g["NxNyNz"]
    x = XmlTransformer(GOLDEN / "files" / f"{name}.xml").parse_xml(NxNyNz=grid, merge=bool(g["merge"]))
    assert np.array_equal(x.data[["x", "y", "z"]].values, g["xyz"]) and np.array_equal(x.data["index"].values, g["index"])
    assert list(x.atoms_type) == list(g["atoms_type"]) and np.array_equal(x.NxNyNz, g["NxNyNz"]) and np.array_equal(x.lxlylz, g["lxlylz"])
    assert np.array_equal(x.grid_spacing, g["lxlylz"] / g["NxNyNz"])


def test_parse_xml_mapping_errors():
    from simubox_b200 import XmlTransformer

    tr = XmlTransformer(GOLDEN / "files" / "splat_coarse.xml")
    with pytest.raises(AssertionError):
        tr.parse_xml(atoms_mapping={"A": 0, "B": 1}, merge=False)           # A1, C2 unmapped
    with pytest.raises(AssertionError):
        tr.parse_xml(atoms_mapping={"A": 0, "A1": 2, "B": 3, "C2": 4}, merge=False)  # not 0, 1, 2, ...


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_splat_matches_reference_golden(sbx_lib, name):
    from simubox_b200 import XmlTransformer

    g = golden(name)
    grid = None if name == "splat_cubic" else g["NxNyNz"]
    x = XmlTransformer(GOLDEN / "files" / f"{name}.xml").parse_xml(NxNyNz=grid, merge=bool(g["merge"]))
    res = XmlTransformer.transform(x, r_cut=float(g["r_cut"]))
    assert res.phi.dtype == np.float64 and np.array_equal(res.phi, g["phi"])


@pytest.mark.gpu
@pytest.mark.parametrize("N,L,r_cut,P", [((32, 32, 32), (20.0, 20.0, 20.0), 2.0, 6000), ((16, 40, 24), (7.3, 19.9, 12.1), 1.1, 3000),
                                        ((8, 8, 8), (30.0, 30.0, 30.0), 2.0, 500)])
def test_cuda_splat_matches_oracle_random(sbx_lib, N, L, r_cut, P):
    from simubox_b200 import splat_particles

    rng = np.random.default_rng(P)
    xyz = (rng.random((P, 3)) - 0.5) * np.array(L)
    xyz[:3] = [[-L[0] / 2, -L[1] / 2, -L[2] / 2], [L[0] / 2 - 1e-13, 0, 0], [0, 0, 0]]
    idx = rng.integers(0, 3, P).astype(np.int32)
    got = splat_particles(xyz, idx, 3, N, L, r_cut)
    assert np.array_equal(got, splat_np.transform(xyz, idx, 3, N, L, r_cut))
    idx2 = idx.copy()
    idx2[::7] = 5  # type ids outside [0, n_types) are skipped
    keep = idx2 < 3
    assert np.array_equal(splat_particles(xyz, idx2, 3, N, L, r_cut), splat_np.transform(xyz[keep], idx2[keep], 3, N, L, r_cut))


@pytest.mark.gpu
def test_splat_to_scatter_pipeline(sbx_lib, tmp_path):
    """Snapshot -> densities -> phin-style file -> read_density -> scattering: the path the splat feeds."""
    from simubox_b200 import Scatter, XmlTransformer, read_density

    g = golden("splat_cubic")
    x = XmlTransformer(GOLDEN / "files" / "splat_cubic.xml").parse_xml()
    res = XmlTransformer.transform(x)
    out = tmp_path / "phin.txt"
    res.write(path=out)
    d = read_density(out)
    assert list(d.NxNyNz) == [8, 8, 8] and d.data.shape == (512, 3)
    assert np.array_equal(d.data.values, np.stack([p.reshape(-1) for p in g["phi"]], axis=1))
    d.lxlylz = x.lxlylz
    sc = Scatter.sacttering_peak(d, cutoff=10)
    assert sc.q_Intensity.shape == (11, 2) and abs(sc.q_Intensity[0, 1] - g["phi"][0].sum() ** 2 / 512) < 1e-9 * g["phi"][0].sum() ** 2
