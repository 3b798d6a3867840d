"""Particle snapshot (HOOMD / GALAMOST style xml) -> grid densities, same call API as the reference's
``SimuBox/toolkit/xmltrans.py``: ``XmlTransformer(path).parse_xml(...)`` (:63-115, host code) and
``XmlTransformer.transform(xml, r_cut)`` (:117-153), whose particle x stencil Python loop runs here as
one CUDA kernel (``csrc/splat.cu``, bit-identical counts)."""
from __future__ import annotations

import re
import xml.dom.minidom
from pathlib import Path
from typing import Optional, Sequence, Union

import numpy as np
import pandas as pd

from . import _lib
from ._device import ptr, stream_ptr, to_device, torch_cuda
from .schema import XML, PathLike, XMLTransResult

__all__ = ["XmlTransformer", "splat_particles"]


def _words(node) -> list:
    return re.sub(r"[\n ]+", " ", node.firstChild.wholeText).strip(" ").split(" ")


def splat_particles(xyz, index, n_types: int, NxNyNz, lxlylz, r_cut: float = 2.0, device: int = 0, return_tensor: bool = False):
    """``phi[t, i, j, k]`` = number of type-``t`` particles within ``r_cut`` of grid point (i, j, k), with the
    reference's arithmetic (xmltrans.py:128-151).  ``xyz`` (P, 3) in [-L/2, L/2), ``index`` (P,) type ids."""
    torch = torch_cuda()
    pos = to_device(xyz if isinstance(xyz, torch.Tensor) else np.ascontiguousarray(xyz, dtype=np.float64), device, torch.float64).contiguous()
    typ = to_device(index if isinstance(index, torch.Tensor) else np.ascontiguousarray(index, dtype=np.int32), device, torch.int32).contiguous()
    if pos.ndim != 2 or pos.shape[1] != 3 or typ.numel() != pos.shape[0]:
        raise ValueError("xyz must be (P, 3) with one type index per particle")
    N = [int(n) for n in NxNyNz]
    phi = torch.empty((int(n_types), *N), dtype=torch.float64, device=pos.device)
    _, pn = _lib.i64(N)
    _, pl = _lib.f64([float(l) for l in lxlylz])
    _lib.check(_lib.lib().sbx_particle_splat(device, ptr(pos), ptr(typ), int(pos.shape[0]), int(n_types), pn, pl, float(r_cut),
                                             ptr(phi), stream_ptr(device)))
    return phi if return_tensor else phi.cpu().numpy()


class XmlTransformer:
    xml: Optional[XML] = None

    def __init__(self, path: PathLike):
        self.path = Path(path)

    @classmethod
    def collect_types(cls, root):
        return _words(root.getElementsByTagName("type")[0])

    @classmethod
    def collect_positions(cls, root):
        return np.array([float(w) for w in _words(root.getElementsByTagName("position")[0])]).reshape((-1, 3))

    @classmethod
    def collect_lxlylz(cls, root):
        box = root.getElementsByTagName("box")[0]
        return np.array([float(box.getAttribute(k)) for k in ("lx", "ly", "lz")])

    def parse_xml(self, NxNyNz: Optional[Union[np.ndarray, Sequence[int]]] = None, atoms_mapping: Optional[dict] = None,
                  merge: bool = True) -> XML:
        """Types, positions and box of the snapshot; grid defaults to ceil(L) rounded up to even (xmltrans.py:63-115)."""
        with self.path.open("r") as fh:
            root = xml.dom.minidom.parse(fh)
        lxlylz = self.collect_lxlylz(root)
        data = pd.DataFrame(data=self.collect_positions(root), columns=["x", "y", "z"])
        data["type"] = self.collect_types(root)
        atoms_type = sorted(data["type"].unique().tolist())
        if atoms_mapping is None:
            if merge:
                atoms_type = sorted(set(re.sub(r"\d", "", a) for a in atoms_type))
            atoms_mapping = {a: k for k, a in enumerate(atoms_type)}
        assert all(a in atoms_mapping for a in atoms_type), f"every atom type needs a mapping entry ({atoms_type})"
        values = list(atoms_mapping.values())
        assert all(np.diff(values) == 1) and min(values) == 0, "mapping values must be 0, 1, 2, ..."
        data["index"] = data["type"].map(atoms_mapping)
        if NxNyNz is None:
            NxNyNz = [n if n % 2 == 0 else n + 1 for n in np.ceil(lxlylz).astype(int)]
        NxNyNz = np.array(NxNyNz, dtype=int)
        self.xml = XML(path=self.path, data=data, NxNyNz=NxNyNz, lxlylz=lxlylz, grid_spacing=lxlylz / NxNyNz,
                       atoms_type=atoms_type, atoms_mapping=atoms_mapping)
        return self.xml

    @classmethod
    def transform(cls, xml: Optional[XML] = None, r_cut: Union[int, float] = 2.0, device: int = 0) -> XMLTransResult:
        xml = xml if xml is not None else cls.xml
        assert xml is not None
        phi = splat_particles(xml.data[["x", "y", "z"]].values, xml.data["index"].values, len(xml.atoms_type), xml.NxNyNz,
                              xml.lxlylz, r_cut, device=device)
        return XMLTransResult(xml=xml, phi=phi)
