"""simubox_b200 -- B200-native density-field analysis (scattering I(q)+peaks, grid Voronoi).

Drop-in for the corresponding SimuBox entry points (same names and signatures):
``read_density``, ``parse_density``, ``read_printout``, ``Scatter``, ``VoronoiCell`` and the
``Density`` / ``ScatterResult`` / ``ScatterPlot`` models.  All numerics run in
``libsimubox_b200.so`` (hand-written sm_100a CUDA + cuFFT); there is no CPU fallback.
"""
from .schema import (AnalyzeMode, ColorMode, Density, DensityParseResult, Fet, OpenCVResult, Printout,
                     ScatterPlot, ScatterResult, VoronoiAnalyzeResult, WeightedMode, XML, XMLTransResult)
from .reader import check_filepath, parse_density, read_density, read_fet, read_printout
from .peak import gaussian_expansion
from .scatter import Scatter
from .voronoi import VoronoiCell
from .xmltrans import XmlTransformer, splat_particles

__all__ = [
    "AnalyzeMode", "ColorMode", "Density", "DensityParseResult", "Fet", "OpenCVResult", "Printout", "ScatterPlot",
    "ScatterResult", "VoronoiAnalyzeResult", "WeightedMode", "check_filepath", "parse_density", "read_density", "read_fet",
    "read_printout",
    "gaussian_expansion", "Scatter", "VoronoiCell", "XML", "XMLTransResult", "XmlTransformer", "splat_particles",
]
