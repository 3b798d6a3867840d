"""CPU suite: the native multithreaded text parser behind read_density (csrc/text_table.cpp) against
Python's float() -- the conversion the reference's pandas python-engine parse performs (reader.py:170-172) --
and against the pandas call itself."""
import ctypes as C
import io

import numpy as np
import pandas as pd
import pytest

from simubox_b200 import read_density
from simubox_b200.reader import _load_text_table


def parse(text: str, skip=0, threads=0):
    return _load_text_table(np.frombuffer(text.encode(), dtype=np.uint8), skip, threads)


def pandas_column(tokens):
    """The reference's parse of one column (reader.py:170-172)."""
    return pd.read_csv(io.StringIO("\n".join(tokens)), header=None, sep="[ ]+", engine="python").values[:, 0].astype(np.float64)


def test_tokens_convert_like_the_reference_parse():
    toks = ["0", "-0.0", "+1.5", "1e-5", "1E+05", ".5", "5.", "0.1", "0.30000000000000004", "2.2250738585072014e-308",
            "4.9e-324", "1.7976931348623157e308", "123456789012345678901234567890", "0.045275193902445166",
            "0.000000000000000000000000000001234", "9007199254740993", "3.141592653589793238462643383279",
            "0.000123456789012345678", "12345678901234567.89", "1e-320", "1.5e300", "inf", "-inf", "nan"]
    got = parse("\n".join(toks) + "\n")[:, 0]
    want = pandas_column(toks)
    ok = ~np.isnan(want)
    assert np.array_equal(got[ok], want[ok]) and np.all(np.isnan(got[~ok]))
    # short mantissas (what the SCFT codes write) are also the correctly rounded doubles Python's float() gives
    short = ["%.6f" % v for v in np.linspace(-3, 7, 1001)] + ["%.10e" % v for v in np.geomspace(1e-12, 1e9, 500)]
    assert np.array_equal(parse("\n".join(short))[:, 0], [float(t) for t in short])


@pytest.mark.parametrize("fmt", ["%.6f", "%.10e", "%.15f", "%.15e", "%.16e", "%.17g", "%r"])
def test_long_mantissas_follow_the_reference_not_strtod(fmt):
    rng = np.random.default_rng(11)
    v = rng.random(4000) * rng.choice([1, 10, 1e-3, 100, 1e-9, 1e12], 4000) * rng.choice([1, -1], 4000)
    toks = [fmt % float(x) for x in v]
    assert np.array_equal(parse("\n".join(toks))[:, 0], pandas_column(toks))


def test_layout_forms():
    body = "1 2 3\n\n  4\t5   6  \r\n7 8 9"          # blank line, tabs, CRLF, no final newline, leading blanks
    assert np.array_equal(parse("8 8 8\n2.0 2.0 2.0\n" + body, skip=2), [[1, 2, 3], [4, 5, 6], [7, 8, 9]])
    assert parse("", 0).shape == (0, 0) and parse("4 4\n", 1).shape == (0, 0)
    with pytest.raises(ValueError, match="row 1"):
        parse("1 2 3\n4 5\n6 7 8\n")
    with pytest.raises(ValueError, match="row 2"):
        parse("1 2\n3 4\n5 x\n")
    with pytest.raises(ValueError):
        parse("1 2\n3 4 5\n")
    with pytest.raises(ValueError):
        parse("1 --2\n")


@pytest.mark.parametrize("threads", [1, 3, 16])
def test_large_table_matches_pandas_python_engine(tmp_path, threads):
    rng = np.random.default_rng(5)
    rows = 60000 if threads > 1 else 6000
    vals = np.concatenate([rng.random((rows, 4)), rng.integers(0, 3, (rows, 1))], axis=1)
    lines = ["40 50 %d" % (rows // 2000), "3.5 4.0 1.25"]
    lines += ["%.8f %.8f %.10e %r %d" % (r[0], r[1], r[2], float(r[3]), int(r[4])) for r in vals]
    text = "\n".join(lines) + "\n"
    got = parse(text, skip=2, threads=threads)
    want = pd.read_csv(io.StringIO("\n".join(lines[2:])), header=None, sep="[ ]+", engine="python").values
    assert np.array_equal(got, want)
    if threads == 16:  # through the public entry point
        (tmp_path / "phout.txt").write_text(text)
        d = read_density(tmp_path, parse_L=True, threads=threads)
        assert np.array_equal(d.data.values, want) and list(d.NxNyNz) == [40, 50, 30] and list(d.lxlylz) == [3.5, 4.0, 1.25]


def test_non_utf8_text_goes_to_the_binary_reader(tmp_path):
    """A file that does not decode is re-read as binary, like the reference (reader.py:107,115-125)."""
    grid = np.array([4, 3, 2], dtype=np.int32)
    box = np.array([1.0, 2.0, 3.0])
    data = np.arange(48, dtype=np.float64) + 0.5
    data[7] = -1.0e-300  # bytes >= 0x80 somewhere in the payload
    p = tmp_path / "phout.txt"
    p.write_bytes(grid.tobytes() + box.tobytes() + data.tobytes())
    d = read_density(p)
    assert list(d.NxNyNz) == [4, 3, 2] and np.array_equal(d.lxlylz, box)
    assert np.array_equal(d.data.values, data.reshape((24, 2), order="F"))
