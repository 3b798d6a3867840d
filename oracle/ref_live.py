"""Import the LIVE reference (KangruiYuan/SimuBox at /root/reference) in this container.

Test infrastructure only (see oracle/__init__.py).  The reference's plotting / GUI
dependencies are not installed and cannot be (no network), so they are replaced by
MagicMock packages through a ``sys.meta_path`` finder; the numeric hot path
(numpy / scipy / pandas / pydantic / cv2) is the real code.

Used by ``oracle/make_golden.py`` and by the ``reference``-marked CPU tests, which
skip when ``/root/reference`` is absent (the GPU box).
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import os
import sys
import warnings
from unittest.mock import MagicMock

REF_ROOT = os.environ.get("SIMUBOX_REF", "/root/reference")

_STUBBED = (
    "matplotlib", "pyvista", "shapely", "skimage", "cycler", "streamlit",
    "plotly", "pyecharts", "simplejson", "mpl_toolkits", "networkx", "sympy",
    "sklearn", "tqdm", "PIL", "openpyxl",
)


class _MockLoader(importlib.abc.Loader):
    def create_module(self, spec):
        m = MagicMock(name=spec.name)
        m.__name__ = spec.name
        m.__path__ = []
        m.__spec__ = spec
        m.__loader__ = self
        return m

    def exec_module(self, module):
        pass


class _MockFinder(importlib.abc.MetaPathFinder):
    """Serve MagicMock modules for packages that are really missing."""

    def __init__(self):
        self._real = {}

    def _is_missing(self, top):
        if top not in self._real:
            self._real[top] = any(
                f is not self and getattr(f, "find_spec", None) and f.find_spec(top, None)
                for f in sys.meta_path
            )
        return not self._real[top]

    def find_spec(self, fullname, path=None, target=None):
        top = fullname.split(".")[0]
        if top in _STUBBED and self._is_missing(top):
            return importlib.machinery.ModuleSpec(fullname, _MockLoader(), is_package=True)
        return None


_loaded = None


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "SimuBox"))


def load():
    """Return the imported reference package ``SimuBox`` (cached)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"live reference not present at {REF_ROOT}")
    sys.meta_path.append(_MockFinder())
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import SimuBox  # noqa: F401
    _loaded = sys.modules["SimuBox"]
    return _loaded


def ref_scatter(mat, lxlylz, cutoff=30, threshold=1e-12):
    """Run the reference's ``Scatter.sacttering_peak`` on an in-memory field.

    Builds a ``Density`` the way ``read_density`` would (``toolkit/reader.py:218-230``)
    and calls ``calculator/scatter.py:68``.  Returns ``(q_Intensity, q_idx)``.
    """
    import numpy as np
    import pandas as pd

    sb = load()
    mat = np.ascontiguousarray(mat, dtype=np.float64)
    shape = np.array(mat.shape)
    dens = sb.schema.Density(
        path=None,
        data=pd.DataFrame(mat.reshape(-1, 1)),
        NxNyNz=shape,
        lxlylz=np.asarray(lxlylz, dtype=np.float64),
        shape=shape,
    )
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res = sb.calculator.Scatter.sacttering_peak(dens, cutoff=cutoff, threshold=threshold, target=0)
    return res.q_Intensity, res.q_idx, res


def ref_show_peak(res, **kw):
    """Reference ``Scatter.show_peak`` (``calculator/scatter.py:125``); plotting is mocked."""
    sb = load()
    sp = sb.calculator.Scatter.show_peak(res, **kw)
    return sp


def ref_label(points, weights, size, method="additive", images=None):
    """Label a grid with the reference's own ``additive``/``power`` + ``np.argmin``.

    Loops ``product(range(..))`` like ``calculator/voronoi.py:284-292``.  ``images``
    (optional, (M,d) offsets) replicates every seed seed-major for the periodic oracle
    (SURVEY.md section 8c); label = argmin // M.
    """
    import itertools
    import numpy as np

    sb = load()
    func = getattr(sb.calculator.VoronoiCell, method)
    points = np.asarray(points, dtype=np.float64)
    S = len(points)
    w = weights
    M = 1
    if images is not None:
        images = np.asarray(images, dtype=np.float64)
        M = len(images)
        points = (points[:, None, :] + images[None, :, :]).reshape(S * M, -1)
        if isinstance(w, np.ndarray):
            w = np.repeat(w, M)
    out = np.empty(size, dtype=np.int32)
    for arr in itertools.product(*[range(n) for n in size]):
        dis = func(points, np.array(arr, dtype=int), w)
        out[arr] = np.argmin(dis) // M
    return out
