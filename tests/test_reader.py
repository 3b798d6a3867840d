"""CPU suite: read_density / parse_density / read_printout against outputs of the live reference
(tests/golden/reader_cases.npz, written by oracle/make_golden.py next to the input files)."""
import numpy as np
import pytest

from conftest import GOLDEN, golden
from simubox_b200 import Density, parse_density, read_density, read_printout

FILES = GOLDEN / "files"
CASES = {"txt3d": ("phin_with_mask_4x6x8.txt", dict(parse_L=True)), "txt2d": ("phout_2d_12x10.txt", {}),
         "bin": ("block_6x4x2.bin", dict(parse_L=True))}


@pytest.mark.parametrize("key", list(CASES))
def test_read_and_parse_match_reference(key):
    g = golden("reader_cases")
    fname, kw = CASES[key]
    d = read_density(FILES / fname, **kw)
    assert isinstance(d, Density)
    assert np.array_equal(d.data.values, g[f"{key}_data"])
    for f in ("NxNyNz", "lxlylz", "shape"):
        assert np.array_equal(getattr(d, f), g[f"{key}_{f}"])
    variants = {"plain": {}, "expand3": dict(expand=3), "t1": dict(target=1), "all": dict(target="all"),
                "expand_vec": dict(expand=[1, 2] if len(d.shape) == 2 else [1, 2, 1.5])}
    if len(d.shape) == 3:
        variants["slice"] = dict(slices=(1, 0))
        variants["permute"] = dict(permute=[2, 0, 1])
    for vname, vkw in variants.items():
        p = parse_density(d, **vkw)
        for f in ("mat", "NxNyNz", "lxlylz", "raw_NxNyNz", "raw_lxlylz", "expand"):
            assert np.array_equal(np.asarray(getattr(p, f)), g[f"{key}_{vname}_{f}"]), (vname, f)
        assert list(p.target) == list(g[f"{key}_{vname}_target"])
        assert p.mat.dtype == np.float64 and p.mat.flags.c_contiguous


def test_bytes_input_and_directory_path():
    raw = (FILES / "block_6x4x2.bin").read_bytes()
    a = read_density(raw, parse_L=True)
    b = read_density(FILES / "block_6x4x2.bin", parse_L=True)
    assert np.array_equal(a.data.values, b.data.values) and a.path is None
    with pytest.raises(ValueError):
        read_density(12345)


def test_header_guess_retry_and_errors(tmp_path):
    # file HAS an L line but the caller did not say so: ragged parse -> retry with parse_L (reader.py:173-186)
    d = read_density(FILES / "phin_with_mask_4x6x8.txt")
    assert np.allclose(d.lxlylz, [2.5, 3.0, 4.25]) and d.data.shape == (192, 5)
    with pytest.raises(AssertionError):
        parse_density(d, permute=[0, 1])
    with pytest.raises(AssertionError):
        parse_density(d, expand=[1, 2])


def test_printout_and_repair(tmp_path):
    g = golden("printout_C4")
    lines = ["x\n"] * 3 + ["Volume Fraction: 1.00000\n", "Chemical Potential: 45.16615\n",
                           "1.00000000 6.20325489 6.20325486 0.00000000 0.00000000 0.00000000\n",
                           "0.00000000 0.00000000 0.00000000 0.00000000 0.00000000 0.00000000\n",
                           "167 :\t1.167571355e+01  8.453360102e+00  4.516615474e+01 -4.194380129e+01  8.784508787e-09  1.683195709e-09\n"]
    (tmp_path / "printout.txt").write_text("".join(lines))
    po = read_printout(tmp_path)
    assert np.array_equal(po.box, g["box"]) and np.array_equal(po.lxlylz, g["lxlylz"])
    assert po.step == int(g["step"]) and po.freeEnergy == float(g["freeEnergy"]) and po.inCompMax == float(g["inCompMax"])
    d = read_density(FILES / "phout_2d_12x10.txt")
    assert np.array_equal(d.lxlylz, [1.0, 1.0])
    d.repair_from_printout(po)
    assert np.array_equal(d.lxlylz, g["lxlylz"][1:])  # only the axes with N != 1


@pytest.mark.reference
def test_reader_on_reference_demo_files():
    from oracle import ref_live

    sb = ref_live.load()
    for rel in ["demos/datasets/voronoi/phout.txt", "demos/datasets/density/bin_block/block.txt",
                "demos/datasets/scatter/phout_C6.txt"]:
        path = ref_live.REF_ROOT + "/" + rel
        a, b = sb.toolkit.read_density(path), read_density(path)
        assert np.array_equal(a.data.values, b.data.values)
        assert np.array_equal(a.NxNyNz, b.NxNyNz) and np.array_equal(a.lxlylz, b.lxlylz) and np.array_equal(a.shape, b.shape)
