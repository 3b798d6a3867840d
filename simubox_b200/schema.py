"""Data models of the density-analysis path, field-for-field compatible with the reference.

Mirrors ``SimuBox/schema/structs/DataFile.py:77-172`` (``Density``, ``DensityParseResult``,
``Printout``), ``ComputedRes.py:31-48`` (``ScatterResult``, ``ScatterPlot``, ``OpenCVResult``),
``MixinStructs.py`` and the enums of ``schema/enums/Modes.py:13-71`` that the entry points
accept.  ``fig`` / ``ax`` are typed ``Any`` because matplotlib is optional here (no graphics
stack on the GPU box); they are ``None`` unless matplotlib is importable and a plot is asked for.
"""
from __future__ import annotations

from enum import Enum
from pathlib import Path
from typing import Any, Optional, Sequence, Union

import numpy as np
import pandas as pd
from pydantic import BaseModel, ConfigDict

RealNum = Union[int, float]
PathLike = Union[Path, str]
FileLike = Union[Path, str, bytes]


class StrEnum(str, Enum):
    def __str__(self) -> str:  # reference enums print as their value
        return str(self.value)


class ColorMode(StrEnum):
    RGB = "rgb"
    HEX = "hex"
    L = "l"
    RANDOM = "random"


class AnalyzeMode(StrEnum):
    VORONOI = "voronoi"
    TRIANGLE = "triangle"
    WEIGHTED = "weighted"


class WeightedMode(StrEnum):
    ADDITIVE = "additive"
    POWER = "power"


class MixinBaseModel(BaseModel):
    model_config = ConfigDict(arbitrary_types_allowed=True)


class FigureAxesMixin(MixinBaseModel):
    fig: Optional[Any] = None
    ax: Optional[Any] = None


class Printout(MixinBaseModel):
    """Subset of the reference ``Printout`` that the density path reads (DataFile.py:17-33)."""
    path: Path
    box: np.ndarray
    lxlylz: np.ndarray
    step: int
    freeEnergy: float
    freeU: float
    freeW: float
    freeS: float
    freeWS: float
    inCompMax: float
    volumes: Optional[np.ndarray] = None
    potentials: Optional[np.ndarray] = None


class Fet(MixinBaseModel):
    """SCFT ``fet.dat`` key/value file (reference DataFile.py:38-74): every field optional, unknown keys ignored
    (pydantic's default, as in the reference)."""
    path: Optional[Path] = None
    xACN: Optional[float] = None
    xCAN: Optional[float] = None
    xABN: Optional[float] = None
    xBAN: Optional[float] = None
    xCBN: Optional[float] = None
    xBCN: Optional[float] = None
    stressX: Optional[float] = None
    stressY: Optional[float] = None
    stressZ: Optional[float] = None
    lx: Optional[float] = None
    ly: Optional[float] = None
    lz: Optional[float] = None
    lx_full: Optional[float] = None
    ly_full: Optional[float] = None
    lz_full: Optional[float] = None
    freeEnergy: Optional[float] = None
    freeAB_sum: Optional[float] = None
    freeAB: Optional[float] = None
    freeBA: Optional[float] = None
    freeAC: Optional[float] = None
    freeCA: Optional[float] = None
    freeBC: Optional[float] = None
    freeCB: Optional[float] = None
    freeW: Optional[float] = None
    freeS: Optional[float] = None
    freeDiff: Optional[float] = None
    inCompMax: Optional[float] = None
    Q0: Optional[float] = None
    VolumeFraction0: Optional[float] = None
    sumVolume: Optional[float] = None
    activity0: Optional[float] = None


class Density(MixinBaseModel):
    """Parsed phout/phin/block file (reference DataFile.py:77-86)."""
    path: Optional[Path] = None
    data: pd.DataFrame
    NxNyNz: Optional[np.ndarray] = None
    lxlylz: Optional[np.ndarray] = None
    shape: Optional[np.ndarray] = None

    def _axis_mask(self):
        return None if self.NxNyNz is None else (np.asarray(self.NxNyNz) != 1)

    def _repair_box(self, box: np.ndarray):
        # DataFile.py:88-115: only fill the box when the file gave none (all ones).
        keep = self._axis_mask()
        if self.lxlylz is None or (all(self.lxlylz == 1) and any(box != 1)):
            self.lxlylz = box if keep is None else box[keep]

    def repair_from_printout(self, printout: Printout):
        self._repair_box(np.asarray(printout.lxlylz))

    def repair_from_fet(self, fet):
        self._repair_box(np.array([fet.lx, fet.ly, fet.lz]))


class DensityParseResult(MixinBaseModel):
    """Reference DataFile.py:159-172."""
    path: Optional[Path] = None
    raw_NxNyNz: np.ndarray
    raw_lxlylz: np.ndarray
    raw_mat: np.ndarray
    lxlylz: np.ndarray
    NxNyNz: np.ndarray
    mat: np.ndarray
    expand: np.ndarray
    target: list[int]


class ScatterResult(MixinBaseModel):
    """Reference ComputedRes.py:31-34."""
    parsed_density: DensityParseResult
    q_Intensity: np.ndarray
    q_idx: dict


class ScatterPlot(FigureAxesMixin):
    """Reference ComputedRes.py:37-41."""
    peaks_location: np.ndarray
    peaks_height: dict
    plot_y: np.ndarray
    plot_x: np.ndarray


class OpenCVResult(MixinBaseModel):
    """Reference ComputedRes.py:44-48."""
    path: Optional[Path]
    parsed_density: DensityParseResult
    facets: Sequence
    centers: np.ndarray


class VoronoiAnalyzeResult(FigureAxesMixin):
    """Reference ComputedRes.py:51-56.  ``voronoi`` is the ``scipy.spatial.Voronoi`` object, ``triangle`` the list of
    kept Delaunay simplices, ``coord_dict`` centre -> set of neighbouring centres."""
    path: Optional[Path] = None
    cv_result: OpenCVResult
    voronoi: Optional[Any] = None
    triangle: Optional[list] = None
    coord_dict: Optional[dict] = None


class XML(MixinBaseModel):
    """Parsed particle snapshot (reference DataFile.py:175-182)."""
    path: Path
    NxNyNz: np.ndarray
    lxlylz: np.ndarray
    grid_spacing: np.ndarray
    data: pd.DataFrame
    atoms_type: list
    atoms_mapping: dict[str, int]


class XMLTransResult(MixinBaseModel):
    """Densities splatted from a particle snapshot (reference DataFile.py:185-221)."""
    xml: XML
    phi: np.ndarray

    def write(self, phi: Optional[np.ndarray] = None, path: Optional[Union[str, Path]] = None, scale: bool = False):
        """One row per grid point, one column per type, ``Nx Ny Nz`` header -- a phin-style file."""
        phi = self.phi if phi is None else phi
        if scale:
            phi = phi / phi.sum(axis=0)
        table = np.stack([np.asarray(p).reshape(-1) for p in phi], axis=1)
        if path is None:
            path = self.xml.path.with_suffix(".txt")
        np.savetxt(path, table, fmt="%.3f", delimiter=" ", header=" ".join(map(str, self.xml.NxNyNz)), comments="")
