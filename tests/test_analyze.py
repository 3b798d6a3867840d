"""CPU suite for the seed finder / analysis entry points that stay on the host as in the reference
(``VoronoiCell.OpenCV`` / ``Analyze`` / ``voronoi`` / ``triangle``, SimuBox/calculator/voronoi.py:31-241) and
``read_fet`` (toolkit/reader.py:289-313): golden vectors written by the live reference on the notebook's own example
(demos/scripts.ipynb cell 28; oracle/make_golden.py::gen_analyze), and the live reference itself when present."""
import numpy as np
import pandas as pd
import pytest

from conftest import golden


def demo_density():
    """demos/datasets/voronoi (2 x 64 x 128, target 0) rebuilt from the stored scatter golden input."""
    from simubox_b200 import Density

    g = golden("scatter_voronoi_demo_2x64x128")
    N = np.array(g["mat"].shape)
    return Density(path=None, data=pd.DataFrame(g["mat"].reshape(-1, 1)), NxNyNz=N, lxlylz=g["L"].copy(), shape=N)


def coord_pairs(coord_dict):
    return np.array([(int(k), int(v)) for k in sorted(int(k) for k in coord_dict) for v in sorted(coord_dict[k])], dtype=np.int32)


def test_opencv_centres_match_reference_golden():
    from simubox_b200 import VoronoiCell

    g = golden("analyze_voronoi_demo")
    res = VoronoiCell.OpenCV(demo_density(), expand=3, slices=(0, 0))
    assert np.array_equal(res.parsed_density.mat.squeeze(), g["tiled"])
    assert np.array_equal(res.parsed_density.lxlylz, g["tiled_lxlylz"])
    assert res.centers.shape == (67, 2)
    assert np.array_equal(res.centers, g["centers"])  # cv2 arithmetic, float32 centres widened to float64
    with pytest.raises(AssertionError):
        VoronoiCell.OpenCV(demo_density(), expand=1)  # 3D field without a slice: the reference asserts ndim == 2


def test_analyze_triangle_matches_reference_golden():
    from simubox_b200 import AnalyzeMode, VoronoiCell

    g = golden("analyze_voronoi_demo")
    res = VoronoiCell.Analyze(demo_density(), AnalyzeMode.TRIANGLE, slices=(0, 0), expand=3, interactive=False)
    assert np.array_equal(np.array(res.triangle, dtype=np.int32), g["triangles"])
    assert np.array_equal(coord_pairs(res.coord_dict), g["coord_pairs"])
    assert res.voronoi is None and res.cv_result.centers.shape == (67, 2)
    narrow = VoronoiCell.Analyze(demo_density(), "triangle", slices=(0, 0), expand=3, interactive=False, theta_max=np.pi / 3 * 1.2)
    assert np.array_equal(np.array(narrow.triangle, dtype=np.int32).reshape(-1, 3), g["triangles_theta72"].reshape(-1, 3))
    cut = VoronoiCell.set_point_limits(res.cv_result, g["centers_lxly"])
    assert np.array_equal(cut, g["point_cut"])


def test_analyze_voronoi_matches_reference_golden():
    from simubox_b200 import AnalyzeMode, VoronoiCell

    g = golden("analyze_voronoi_demo")
    res = VoronoiCell.Analyze(demo_density(), AnalyzeMode.VORONOI, slices=(0, 0), expand=3, interactive=False)
    assert np.array_equal(res.voronoi.points, g["centers_lxly"])
    assert np.array_equal(res.voronoi.vertices, g["vor_vertices"])  # Qhull, same scipy build as the generator
    assert np.array_equal(res.voronoi.ridge_points, g["vor_ridge_points"])
    assert res.triangle is None and res.coord_dict is None
    with pytest.raises(NotImplementedError):
        VoronoiCell.Analyze(demo_density(), "weighted", slices=(0, 0))


def test_read_fet_and_repair(tmp_path):
    from simubox_b200 import read_fet

    p = tmp_path / "fet.dat"
    p.write_text("0 lx 3.25\n1 ly 4.5\n\n2 lz 1\n3 freeEnergy -2.125\n4 stressX 1e-7\n")
    fet = read_fet(tmp_path)  # directory + default file name, like the reference's check_filepath
    assert (fet.lx, fet.ly, fet.lz, fet.freeEnergy, fet.stressX, fet.freeW) == (3.25, 4.5, 1.0, -2.125, 1e-7, None)
    assert fet.path == p
    d = demo_density()
    d.lxlylz = np.ones(3)
    d.repair_from_fet(fet)
    assert np.array_equal(d.lxlylz, [3.25, 4.5, 1.0])
    d.repair_from_fet(read_fet(p))  # a box that is already set stays
    assert np.array_equal(d.lxlylz, [3.25, 4.5, 1.0])
    (tmp_path / "bad.dat").write_text("0 lx 3.25 extra\n")
    with pytest.raises(ValueError):
        read_fet(tmp_path / "bad.dat")


@pytest.mark.reference
def test_live_reference_analyze_and_read_fet(tmp_path):
    """The same calls on the live reference (its read_fet raises: reader.py:306 unpacks readlines() into two names)."""
    import warnings

    from oracle import ref_live
    from simubox_b200 import AnalyzeMode, VoronoiCell

    sb = ref_live.load()
    ref = ref_live.REF_ROOT + "/demos/datasets/voronoi/"
    d = sb.toolkit.read_density(ref + "phout.txt")
    d.repair_from_printout(sb.toolkit.read_printout(ref + "printout.txt"))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = sb.calculator.VoronoiCell.Analyze(d, sb.schema.AnalyzeMode.TRIANGLE, slices=(0, 0), expand=3, interactive=False)
    mine = VoronoiCell.Analyze(demo_density(), AnalyzeMode.TRIANGLE, slices=(0, 0), expand=3, interactive=False)
    assert np.array_equal(mine.cv_result.centers, r.cv_result.centers)
    assert np.array_equal(np.array(mine.triangle), np.array(r.triangle))
    assert coord_pairs(mine.coord_dict).tolist() == coord_pairs(r.coord_dict).tolist()
    p = tmp_path / "fet.dat"
    p.write_text("0 lx 3.25\n1 ly 4.5\n2 lz 1\n")
    with pytest.raises(ValueError):
        sb.toolkit.read_fet(p)
