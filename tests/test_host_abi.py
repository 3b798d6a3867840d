"""CPU suite: the C-ABI library loads, exports every symbol include/sbx.h declares, and its
host-side shell table (geometry-only plan, no CUDA call) reproduces the reference geometry."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

from conftest import ROOT, golden, golden_names
from oracle import scatter_np


def header_symbols():
    text = (ROOT / "include" / "sbx.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sbx_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(sbx_lib):
    lib = sbx_lib.lib()
    names = header_symbols()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), f"{name} declared in sbx.h but not exported"
    assert set(names) == set(sbx_lib._SIGNATURES), "ctypes table and header disagree"
    assert lib.sbx_version() == 100


def test_no_cpu_fallback(sbx_lib):
    """Without a device the compute entry points must fail loudly, not compute on the CPU."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(sbx_lib.SbxError) as ei:
        sbx_lib.ScatterPlan((8, 8, 8), (1, 1, 1), device=0)
    assert ei.value.code == -2
    plan = sbx_lib.ScatterPlan((8, 8, 8), (1, 1, 1), device=sbx_lib.SBX_DEVICE_NONE)
    buf = (ctypes.c_double * 512)()
    rc = sbx_lib.lib().sbx_scatter_execute(plan.handle, ctypes.addressof(buf), 1, ctypes.addressof(buf), None)
    assert rc == -2
    from simubox_b200 import Scatter

    with pytest.raises(RuntimeError):
        Scatter.sacttering_peak_batch(np.zeros((1, 8, 8, 8)), (1, 1, 1))


def test_argument_errors(sbx_lib):
    for N, L, cutoff, thr in [((7, 8, 8), (1, 1, 1), 30, 1e-12), ((8, 8, 8), (1, 0, 1), 30, 1e-12),
                              ((8, 8), (1, 1), -1, 1e-12), ((8, 8), (1, 1), 3, -1.0)]:
        with pytest.raises(sbx_lib.SbxError) as ei:
            sbx_lib.ScatterPlan(N, L, cutoff=cutoff, threshold=thr, device=sbx_lib.SBX_DEVICE_NONE)
        assert ei.value.code == -1


@pytest.mark.parametrize("name", golden_names("scatter_"))
def test_host_shell_table_matches_reference_geometry(sbx_lib, name):
    g = golden(name)
    N, L, cutoff, thr = tuple(int(n) for n in g["shape"]), g["L"], int(g["cutoff"]), float(g["threshold"])
    plan = sbx_lib.ScatterPlan(N, L, cutoff=cutoff, threshold=thr, device=sbx_lib.SBX_DEVICE_NONE)
    ref = g["q_Intensity"]
    assert plan.n_shells == len(ref)
    assert np.array_equal(plan.q, ref[:, 0]), "q must be bit-exact"
    counts = np.bincount(g["qidx_shell"], minlength=len(ref))
    assert np.array_equal(plan.count, counts)
    full = len(ref) - 1 if len(ref) == cutoff + 1 else len(ref)
    nd = len(N)
    for s in range(len(ref)):
        mine = plan.members(s)
        want = np.ascontiguousarray(g["qidx_vec"][g["qidx_shell"] == s])
        if s < full:
            a = np.sort(np.ascontiguousarray(mine).view([("", float)] * nd), axis=0)
            b = np.sort(want.view([("", float)] * nd), axis=0)
            assert np.array_equal(a, b), f"q_idx[{s}] multiset differs"
        else:  # truncated row: one member, must belong to the shell (same k^2 up to threshold)
            assert mine.shape == (1, nd)


@pytest.mark.parametrize("N,L,cutoff,thr", [
    ((8, 12, 10), (1.0, 2.0, 3.0), 10, 1e-12), ((64, 128), (3.3, 3.3), 10 ** 6, 1e-12), ((16, 16, 16), (4, 4, 4), 20, 0.3),
    ((2, 64, 128), (1.0, 3.7, 7.4), 30, 1e-12), ((16,), (3.0,), 30, 1e-12), ((32, 32, 32), (8, 8, 8), 200, 0.0)])
def test_host_shell_table_matches_numpy_table(sbx_lib, N, L, cutoff, thr):
    plan = sbx_lib.ScatterPlan(N, L, cutoff=cutoff, threshold=thr, device=sbx_lib.SBX_DEVICE_NONE)
    reps, _, _, members, kv = scatter_np.shell_table(N, L, cutoff, thr)
    assert np.array_equal(plan.q, np.sqrt(reps))
    assert np.array_equal(plan.count, [len(members[s]) for s in range(len(reps))])
    for s in range(len(reps)):
        sub = np.unravel_index(members[s], N)
        assert np.array_equal(plan.members(s), np.stack([kv[d][sub[d]] for d in range(len(N))], axis=1))


def test_large_grid_plan_is_cheap(sbx_lib):
    plan = sbx_lib.ScatterPlan((1024, 1024, 1024), (32.0, 32.0, 32.0), device=sbx_lib.SBX_DEVICE_NONE)
    assert plan.n_shells == 31 and plan.count[1] == 6 and plan.count[-1] == 1
