import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA sm_100 device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs the live reference at /root/reference (build container only)")


def pytest_collection_modifyitems(config, items):
    have_ref = os.path.isdir(os.environ.get("SIMUBOX_REF", "/root/reference") + "/SimuBox")
    for item in items:
        if "reference" in item.keywords and not have_ref:
            item.add_marker(pytest.mark.skip(reason="live reference not present on this machine"))


def golden(name):
    return np.load(GOLDEN / f"{name}.npz", allow_pickle=False)


def golden_names(prefix):
    return sorted(p.stem for p in GOLDEN.glob(f"{prefix}*.npz"))


_FIELDS = {}


def golden_field(g):
    """Input field of a scatter golden (stored, or regenerated for the large synthetic ones; the regenerated
    128^3 / 256^3 fields are kept for the session, a few hundred MB)."""
    if "mat" in g.files:
        return g["mat"]
    from simubox_b200 import synth

    shape = tuple(int(n) for n in g["shape"])
    key = (int(g["field_index"]), shape, int(g["cells"]))
    if key not in _FIELDS:
        mat = synth.field_numpy(key[0], shape, cells=key[2])
        assert float(mat.sum()) == float(g["checksum"]), "synthetic generator drifted from the golden input"
        if len(_FIELDS) >= 12:
            _FIELDS.pop(next(iter(_FIELDS)))
        _FIELDS[key] = mat
    return _FIELDS[key]


def check_scatter_rows(qI, ref, cutoff, member_I=None):
    """Parity rules 1-3 of SURVEY.md 8(c): q bit-exact; I within 1e-9*|ref| + 1e-12*scale for complete
    rows; the truncated last row must equal SOME member of its shell (``member_I`` = candidates).  ``scale`` is the
    strongest shell OTHER than q = 0 (the q = 0 row is ~N times the mean density squared and would make the absolute
    term meaningless for weak shells); a table with the q = 0 row only falls back to that row."""
    assert qI.shape == ref.shape
    assert np.array_equal(qI[:, 0], ref[:, 0]), "q column must be bit-exact"
    n = len(ref)
    full = n - 1 if n == cutoff + 1 else n
    scale = np.abs(ref[1:, 1]).max() if n > 1 else np.abs(ref[:, 1]).max()
    tol = 1e-9 * np.abs(ref[:full, 1]) + 1e-12 * scale
    err = np.abs(qI[:full, 1] - ref[:full, 1])
    assert np.all(err <= tol), f"shell intensity mismatch: worst {np.max(err / np.maximum(tol, 1e-300)):.3g} x tol"
    if full < n and member_I is not None:
        t = 1e-9 * np.abs(member_I) + 1e-12 * scale
        assert np.any(np.abs(member_I - qI[-1, 1]) <= t), "truncated row is not a member of its shell"
        assert np.any(np.abs(member_I - ref[-1, 1]) <= t)


@pytest.fixture(scope="session")
def sbx_lib():
    from simubox_b200 import _lib, build

    build.build()
    return _lib
