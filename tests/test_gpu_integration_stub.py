"""The reference-side binding shown in INTEGRATION.md is executed as written (the ```python blocks that define
``scatter_rows`` and ``label_image``; only the library name is replaced by the path of the in-tree .so) and checked
against goldens of the live reference: what a SimuBox maintainer would paste must work."""
import re

import numpy as np
import pytest

from conftest import ROOT, check_scatter_rows, golden, golden_names

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def stub(sbx_lib):
    from simubox_b200 import build

    text = (ROOT / "INTEGRATION.md").read_text()
    blocks = re.findall(r"```python\n(.*?)```", text, flags=re.S)
    code = "\n".join(b for b in blocks if "def scatter_rows" in b or "def label_image" in b)
    assert "def scatter_rows" in code and "def label_image" in code
    code = code.replace('"libsimubox_b200.so"', repr(str(build.LIB)))
    ns = {}
    exec(compile(code, "INTEGRATION.md", "exec"), ns)
    return ns


@pytest.mark.parametrize("name", ["scatter_synth_16_f0", "scatter_random_8x12x10_c10", "scatter_phoutC6_t0_unitbox"])
def test_documented_scatter_stub(stub, name):
    g = golden(name)
    qI = stub["scatter_rows"](g["mat"], g["L"], cutoff=int(g["cutoff"]), threshold=float(g["threshold"]))
    check_scatter_rows(qI, g["q_Intensity"], int(g["cutoff"]))


@pytest.mark.parametrize("name", golden_names("vor"))
def test_documented_voronoi_stub(stub, name):
    g = golden(name)
    w = g["weights"] if g["weights"].ndim else float(g["weights"])
    box = tuple(g["box"]) if g["box"].size else None
    lab = stub["label_image"](g["points"], w, tuple(int(n) for n in g["size"]), method=str(g["method"]), box=box)
    assert np.array_equal(lab, g["labels"])
