"""phin / phout / block density readers -- same call API as the reference's toolkit.

Mirrors ``SimuBox/toolkit/reader.py``: ``read_printout`` (:43-75), ``read_density`` (:78-230,
text and binary layouts) and ``parse_density`` (:345-415, target / permute / slices / periodic
``expand``).  Argument names, defaults, return models and error behaviour follow the
reference; the body is new (the native multithreaded parser csrc/text_table.cpp instead of the pandas python engine).

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

from .schema import Density, DensityParseResult, Fet, FileLike, PathLike, Printout, RealNum


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


def read_fet(path: FileLike, filename: str = "fet.dat", mode: str = "r", encoding: str = "utf-8", **kwargs) -> Fet:
    """SCFT ``fet.dat``: one ``<index> <key> <value>`` triple per non-blank line -> :class:`Fet`
    (reference reader.py:289-313; feeds ``Density.repair_from_fet``, web/pages/2_Scatter.py:95).

    The reference's own body unpacks ``readlines()`` into two names (reader.py:306) and therefore raises
    ``ValueError`` on any file that does not have exactly two lines [probed]; this is the parse its next three
    lines spell out: drop the first token, ``key, value = rest``, ``float(value)``."""
    path = check_filepath(path, filename=filename)
    with path.open(mode=mode, encoding=encoding) as fh:
        rows = [line.strip().split()[1:] for line in fh if line.strip()]
    res = {}
    for key, val in rows:  # a line with another token count raises ValueError, as the reference's loop would
        res[key] = float(val)
    return Fet(path=path, **res)


def _bad_grid(vec: np.ndarray) -> bool:
    return bool(np.any(vec > 1000) or np.any(vec <= 0))


def _file_bytes(path: Path) -> np.ndarray:
    return np.fromfile(path, dtype=np.uint8)


def _load_text_table(raw: np.ndarray, skip: int, threads: int = 0) -> np.ndarray:
    """Blank-separated numeric rows -> (rows, ncol) fp64 through the native multithreaded parser
    (csrc/text_table.cpp restates pandas' ``precise_xstrtod``, the conversion behind the reference's
    ``pd.read_csv(..., engine="python")``, reader.py:170-172: the doubles are the reference's bit for bit, which beyond
    15 significant digits can be one ulp from the correctly rounded value ``float()`` would give).  ValueError on
    non-numeric or ragged rows.  One divergence: a row SHORTER than the first makes pandas NaN-pad without an error,
    here it raises and so takes the reference's header-guess retry (reader.py:173-186) -- benign for well-formed
    files, which is all the reference's writers produce."""
    import ctypes as C

    from . import _lib

    L_ = _lib.lib()
    ptr, n = C.c_void_p(raw.ctypes.data), C.c_int64(raw.size)
    rows, cols = C.c_int64(), C.c_int32()
    _lib.check(L_.sbx_text_table_shape(ptr, n, skip, threads, C.byref(rows), C.byref(cols)))
    if rows.value == 0 or cols.value == 0:
        return np.empty((0, 0), dtype=np.float64)
    out = np.empty((rows.value, cols.value), dtype=np.float64)
    rc = L_.sbx_text_table_parse(ptr, n, skip, rows, cols, out.ctypes.data_as(C.POINTER(C.c_double)), threads)
    if rc != 0:
        raise ValueError((L_.sbx_last_error() or b"").decode())
    return out


def _decodes(raw: np.ndarray, encoding: str) -> bool:
    """The reference decodes the whole file and treats a UnicodeDecodeError as "binary" (reader.py:107,
    115-125).  Pure-ASCII bytes decode under every ASCII-compatible codec: checked natively; otherwise decode."""
    import ctypes as C

    from . import _lib

    if _lib.lib().sbx_text_is_ascii(C.c_void_p(raw.ctypes.data), C.c_int64(raw.size)) == 1 and \
            encoding.lower().replace("_", "-") in ("utf-8", "utf8", "ascii", "latin-1", "latin1", "gbk", "gb2312", "gb18030", "cp936"):
        return True
    try:
        raw.tobytes().decode(encoding)
        return True
    except UnicodeDecodeError:
        return False


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
            text_raw = _file_bytes(path)
            if not _decodes(text_raw, encoding):
                return read_density(path, parse_N=True, parse_L=True, filename=filename, mode=mode,
                                    encoding=encoding, binary=True, **kwargs)
            pos = 0
            for _ in range(2):  # the two possible header lines
                nl = np.flatnonzero(text_raw[pos:pos + 4096] == 10)
                end = pos + int(nl[0]) + 1 if len(nl) else len(text_raw)
                header.append(text_raw[pos:end].tobytes().decode(encoding))
                pos = end
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
            table = _load_text_table(text_raw, skip, int(kwargs.get("threads", 0)))
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
