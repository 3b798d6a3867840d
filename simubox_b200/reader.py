"""phin / phout / block density readers -- same call API as the reference's toolkit.

Mirrors ``SimuBox/toolkit/reader.py``: ``read_printout`` (:43-75), ``read_density`` (:78-230,
text and binary layouts) and ``parse_density`` (:345-415, target / permute / slices / periodic
``expand``).  Argument names, defaults, return models and error behaviour follow the
reference; the body is new (numpy's C text parser instead of the pandas python engine).

File formats:
  text   line 0 = ``Nx Ny Nz`` ints (left-padded with 1s to 3 entries), optional line 1 =
         ``lx ly lz`` (``parse_L``), then one row per grid point, C order, blank-separated.
  binary 3 x int32 (or 3 x int64 when any int32 value is > 1000 or <= 0), 3 x float64 box
         (``parse_L``), then float64 data stored column-contiguous (``order='F'``).
"""
from __future__ import annotations

import os
import re
from pathlib import Path
from typing import Iterable, Optional, Sequence, Union

import numpy as np
import pandas as pd

from .schema import Density, DensityParseResult, FileLike, PathLike, Printout, RealNum


def check_filepath(path: PathLike, filename: str) -> Path:
    path = Path(path)
    return path if path.is_file() else path / filename


def read_printout(path: PathLike, filename: str = "printout.txt", mode: str = "r",
                  encoding: str = "utf-8", **kwargs) -> Printout:
    """Parse the tail of a TOPS ``printout.txt`` (reference reader.py:43-75): third-from-last
    line is the box, last line the step / free-energy summary."""
    path = check_filepath(path, filename)
    with path.open(mode=mode, encoding=encoding) as fh:
        lines = fh.readlines()

    def floats(line: str, skip: int) -> np.ndarray:
        return np.array([float(tok) for tok in line.strip().split(" ")[skip:]])

    box = floats(lines[-3], 0)
    summary = [float(tok) for tok in re.findall("[.0-9e+-]+", lines[-1])]
    return Printout(
        path=path, box=box, lxlylz=box[:3].copy(),
        step=summary[0], freeEnergy=summary[1], freeU=summary[2], freeW=summary[3],
        freeS=summary[4], freeWS=summary[3] + summary[4], inCompMax=summary[-1],
        volumes=floats(lines[-5], 2), potentials=floats(lines[-4], 2),
    )


def _bad_grid(vec: np.ndarray) -> bool:
    return bool(np.any(vec > 1000) or np.any(vec <= 0))


def _load_text_table(path: Path, skip: int) -> np.ndarray:
    """Blank-separated numeric rows -> (rows, ncol) fp64.  numpy's C tokenizer parses with
    correctly rounded strtod, i.e. the same doubles Python's float() gives the reference."""
    return np.loadtxt(path, dtype=np.float64, skiprows=skip, ndmin=2, comments=None)


def read_density(path: FileLike, parse_N: bool = True, parse_L: bool = False,
                 filename: str = "phout.txt", binary: bool = False, mode: str = "r",
                 encoding: str = "utf-8", **kwargs) -> Density:
    """Read a density-like file into a :class:`Density` (reference reader.py:78-230)."""
    raw: Optional[bytes] = None
    header: list[str] = []
    if isinstance(path, bytes):
        raw, text = path, False
    elif isinstance(path, (str, Path)):
        path = check_filepath(path, filename)
        if binary or path.suffix == ".bin":
            text = False
        else:
            text = True
            try:
                with path.open(mode=mode, encoding=encoding) as fh:
                    header = [fh.readline() for _ in range(2)]
                    for _ in fh:  # the reference decodes the whole file; undecodable => binary
                        pass
            except UnicodeDecodeError:
                return read_density(path, parse_N=True, parse_L=True, filename=filename, mode=mode,
                                    encoding=encoding, binary=True, **kwargs)
    else:
        raise ValueError(type(path))

    def from_binary(dtype, count, offset):
        if raw is not None:
            return np.frombuffer(raw, dtype=dtype, count=count, offset=offset)
        return np.fromfile(path, dtype=dtype, count=count, offset=offset)

    skip = 0
    offset = 0
    if parse_N:
        if text:
            grid = [int(tok) for tok in header[skip].strip().split(" ")]
            grid = [1] * (3 - len(grid)) + grid
            NxNyNz = np.array(grid)
        else:
            NxNyNz = from_binary(np.int32, 3, 0)
            if _bad_grid(NxNyNz):
                NxNyNz = from_binary(np.int64, 3, 0)
                if _bad_grid(NxNyNz):
                    raise ValueError(f"cannot parse NxNyNz: {NxNyNz}")
                offset += 24
            else:
                offset += 12
        skip += 1
    else:
        NxNyNz = kwargs.get("NxNyNz")

    if parse_L:
        if text:
            lxlylz = np.array([float(tok) for tok in header[skip].strip().split(" ")])
        else:
            lxlylz = from_binary(np.float64, 3, offset)
            offset += 24
        skip += 1
    else:
        lxlylz = kwargs.get("lxlylz", np.ones(3))

    retry = dict(parse_N=True, parse_L=True, filename=filename, mode=mode, encoding=encoding, binary=binary)
    if text:
        try:
            table = _load_text_table(path, skip)
        except ValueError:
            # ragged first rows: the caller guessed the header length wrong (reader.py:173-186)
            if not parse_L or not parse_N:
                return read_density(path, **retry, **kwargs)
            raise
        data = pd.DataFrame(table)
    else:
        nbytes = len(raw) if raw is not None else os.path.getsize(path)
        count = (nbytes - offset) // 8
        flat = from_binary(np.float64, count, offset)
        assert NxNyNz is not None
        try:
            data = pd.DataFrame(flat.reshape((int(np.prod(NxNyNz)), -1), order=kwargs.get("order", "F")))
        except ValueError:
            if not parse_L or not parse_N:
                return read_density(path, **retry, **kwargs)
            raise

    if NxNyNz is not None:
        keep = NxNyNz != 1
        shape = NxNyNz[keep]
        lxlylz = np.asarray(lxlylz)[keep]
    else:
        shape = kwargs.get("shape", NxNyNz)

    return Density(path=path if isinstance(path, (str, Path)) else None, data=data,
                   NxNyNz=NxNyNz, lxlylz=lxlylz, shape=shape)


def parse_density(density: Density, target: Union[int, Iterable[int], str] = 0,
                  permute: Optional[Iterable[int]] = None, slices: Optional[tuple] = None,
                  expand: Union[RealNum, Sequence[RealNum]] = 1, **kwargs) -> DensityParseResult:
    """Column select, axis permutation, slicing and periodic extension (reference reader.py:345-415).

    NOTE (reference behaviour, kept): ``permute`` relabels ``shape`` / ``lxlylz`` before the flat
    column is reshaped -- it does not transpose data.  ``expand`` pads each side with
    ``around((expand-1)/2 * shape)`` wrapped samples and scales the box by ``expand``.
    """
    assert density.shape is not None, "Density.shape is required"
    shape = np.array(density.shape).copy()
    box = np.array(density.lxlylz).copy()
    if permute is None:
        order = np.arange(len(shape))
    else:
        order = np.array(permute)
        assert len(order) == len(shape), f"length of permute({len(order)}) and shape({len(shape)}) mismatch"
    shape = shape[order]
    box = box[order]

    if target == "all":
        target = list(density.data.columns)
    elif isinstance(target, (int, np.integer)):
        target = [int(target)]

    mats = []
    for t in target:
        m = density.data[t].values.reshape(shape)
        if slices is not None:
            m = np.take(m, indices=slices[0], axis=slices[1])
        mats.append(m)
    if slices is not None:
        shape = np.delete(shape, slices[1])
        box = np.delete(box, slices[1])

    if isinstance(expand, (int, float)):
        factor = np.array([max(expand, 1)] * len(shape))
    else:
        assert len(expand) == len(shape), f"expand has {len(expand)} entries for a {len(shape)}-d field"
        factor = np.array(expand)
        factor[factor < 1] = 1
    halo = np.around((factor - 1) / 2 * shape).astype(int)

    return DensityParseResult(
        path=density.path,
        raw_NxNyNz=shape.copy(), raw_lxlylz=box.copy(), raw_mat=np.stack(mats),
        lxlylz=box * factor, NxNyNz=shape + 2 * halo,
        mat=np.stack([np.pad(m, [(h, h) for h in halo], "wrap") for m in mats]),
        expand=factor.copy(), target=[int(t) for t in target],
    )
