"""GPU parity suite for the scattering path: CUDA (through the C ABI) vs the golden vectors of
the live reference and vs the numpy oracle on seeded inputs; size-independent properties at
the BASELINE sizes.  Tolerances: q bit-exact; shell I within
1e-9*|ref| + 1e-12*max(ref); peak indices exact; plot_y within 1e-9 relative."""
import json

import numpy as np
import pytest

from conftest import check_scatter_rows, golden, golden_field, golden_names
from oracle import scatter_np
from test_oracle_golden import last_shell_members

pytestmark = pytest.mark.gpu

DENSE, GATHER, PRUNED, FUSED = 1, 2, 3, 4


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


def run_scatter(sbx_lib, torch, mats, L, cutoff, thr, path, peaks=None):
    from simubox_b200 import Scatter

    fields = torch.from_numpy(np.ascontiguousarray(np.stack(mats))).cuda()
    kw = dict(peaks=True, width=peaks["width"], step=peaks["step"], peak_cutoff=peaks["cutoff"],
              min_height=peaks["min_height"]) if peaks else {}
    try:
        out = Scatter.sacttering_peak_batch(fields, L, cutoff=cutoff, threshold=thr, path=path, max_batch=len(mats), **kw)
    except sbx_lib.SbxError as e:
        if path in (PRUNED, FUSED) and e.code == -1:
            pytest.skip("pruned-DFT path not applicable to this grid / cutoff")
        raise
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("path", [DENSE, GATHER, PRUNED, FUSED])
@pytest.mark.parametrize("name", golden_names("scatter_"))
def test_scatter_matches_reference_golden(sbx_lib, torch_cuda, name, path):
    g = golden(name)
    mat, L, cutoff, thr = golden_field(g), g["L"], int(g["cutoff"]), float(g["threshold"])
    out = run_scatter(sbx_lib, torch_cuda, [mat], L, cutoff, thr, path)
    qI = np.stack([out["q"], out["I"][0].cpu().numpy()], axis=1)
    check_scatter_rows(qI, g["q_Intensity"], cutoff, last_shell_members(mat, L, cutoff, thr))


@pytest.mark.parametrize("name", [n for n in golden_names("scatter_") if "peaks_index" in golden(n).files])
def test_fused_peaks_match_reference_golden(sbx_lib, torch_cuda, name):
    g = golden(name)
    kw = json.loads(str(g["peak_kw"]))
    mat, L, cutoff, thr = golden_field(g), g["L"], int(g["cutoff"]), float(g["threshold"])
    paths = [DENSE, GATHER] + ([PRUNED, FUSED] if "synth" in name else [])
    for path in paths:
        out = run_scatter(sbx_lib, torch_cuda, [mat, mat], L, cutoff, thr, path, peaks=kw)
        for b in range(2):
            n = int(out["n_peaks"][b])
            assert np.array_equal(out["peak_idx"][b, :n].cpu().numpy(), g["peaks_index"]), (path, b)


@pytest.mark.parametrize("name", [n for n in golden_names("scatter_") if "peaks_index" in golden(n).files])
def test_show_peak_api_matches_reference_golden(sbx_lib, torch_cuda, name):
    """Scatter.show_peak (stand-alone peaks kernel) fed the REFERENCE's q_Intensity."""
    from simubox_b200 import Scatter
    from simubox_b200.schema import DensityParseResult, ScatterResult

    g = golden(name)
    kw = json.loads(str(g["peak_kw"]))
    z = np.zeros(1)
    parsed = DensityParseResult(raw_NxNyNz=z, raw_lxlylz=z, raw_mat=z, lxlylz=z, NxNyNz=z, mat=z, expand=z, target=[0])
    sp = Scatter.show_peak(ScatterResult(parsed_density=parsed, q_Intensity=g["q_Intensity"], q_idx={}), **kw)
    assert np.array_equal(sp.plot_x, g["plot_x"])
    assert np.allclose(sp.plot_y, g["plot_y"], rtol=1e-9, atol=1e-9 * np.abs(g["plot_y"]).max())
    assert np.array_equal(sp.peaks_location, g["peaks_location"])
    assert np.allclose(sp.peaks_height["peak_heights"], g["peak_heights"], rtol=1e-9)


def test_drop_in_entry_point_on_density(sbx_lib, torch_cuda):
    """read_density-style Density -> Scatter.sacttering_peak -> ScatterResult, incl. expand and q_idx."""
    import pandas as pd
    from simubox_b200 import Density, Scatter

    g = golden("scatter_phoutC6_t0_unitbox")
    mat = g["mat"]
    dens = Density(data=pd.DataFrame(mat.reshape(-1, 1)), NxNyNz=np.array([1, *mat.shape]), lxlylz=g["L"].copy(),
                   shape=np.array(mat.shape))
    res = Scatter.sacttering_peak(dens, cutoff=int(g["cutoff"]))
    check_scatter_rows(res.q_Intensity, g["q_Intensity"], int(g["cutoff"]))
    counts = np.bincount(g["qidx_shell"], minlength=len(res.q_Intensity))
    assert [len(res.q_idx[s]) for s in range(len(counts))] == list(counts)
    res3 = Scatter.sacttering_peak(dens, cutoff=12, expand=3)
    tiled = np.pad(mat, [(n, n) for n in mat.shape], "wrap")
    ref, _ = scatter_np.scattering_peak(tiled, g["L"] * 3, 12, 1e-12)
    check_scatter_rows(res3.q_Intensity, ref, 12)


@pytest.mark.parametrize("shape,L,cutoff", [((20, 12, 14), (3.0, 2.0, 2.5), 40), ((24, 24, 24), (5.0, 5.0, 5.0), 10 ** 6),
                                            ((30, 18), (2.0, 3.0), 10 ** 6), ((64,), (7.0,), 10 ** 6), ((2, 2, 2), (1, 1, 1), 30)])
def test_random_fields_vs_oracle_all_paths(sbx_lib, torch_cuda, shape, L, cutoff):
    rng = np.random.default_rng(hash(shape) % 2 ** 32)
    mats = [rng.standard_normal(shape) for _ in range(3)]
    for path in (DENSE, GATHER):
        out = run_scatter(sbx_lib, torch_cuda, mats, L, cutoff, 1e-12, path)
        for b, mat in enumerate(mats):
            ref, _ = scatter_np.scattering_peak_table(mat, L, cutoff, 1e-12)
            qI = np.stack([out["q"], out["I"][b].cpu().numpy()], axis=1)
            check_scatter_rows(qI, ref, cutoff, last_shell_members(mat, L, cutoff, 1e-12))


def test_gather_path_is_bitwise_reproducible(sbx_lib, torch_cuda):
    from simubox_b200 import synth

    mats = [synth.field_numpy(i, (32, 32, 32)) for i in range(4)]
    a = run_scatter(sbx_lib, torch_cuda, mats, synth.box_for(2), 30, 1e-12, GATHER)["I"].cpu().numpy()
    b = run_scatter(sbx_lib, torch_cuda, mats[::-1], synth.box_for(2), 30, 1e-12, GATHER)["I"].cpu().numpy()
    assert np.array_equal(a, b[::-1])


def test_host_streaming_entry_point(sbx_lib, torch_cuda):
    """sbx_scatter_execute_host: numpy in, numpy out, ragged last chunk, pinned and pageable input."""
    from simubox_b200 import Scatter, synth

    mats = np.stack([synth.field_numpy(i, (32, 32, 32)) for i in range(7)])
    L = synth.box_for(2)
    dev = run_scatter(sbx_lib, torch_cuda, list(mats), L, 30, 1e-12, GATHER, peaks=dict(width=0.5, step=2000, cutoff=20, min_height=1))
    pinned = torch_cuda.from_numpy(mats).pin_memory()
    for src in (mats, pinned):
        out = Scatter.sacttering_peak_batch(src, L, peaks=True, max_batch=3, path=GATHER)
        assert np.array_equal(out["I"], dev["I"].cpu().numpy())
        assert np.array_equal(out["n_peaks"], dev["n_peaks"].cpu().numpy())
        for b in range(len(mats)):
            n = int(out["n_peaks"][b])
            assert np.array_equal(out["peak_idx"][b, :n], dev["peak_idx"][b, :n].cpu().numpy())


@pytest.mark.parametrize("n,cells", [(128, 2), (256, 4)])
def test_properties_at_baseline_sizes(sbx_lib, torch_cuda, n, cells):
    """Size-independent checks at the BASELINE grids (the numpy oracle needs minutes there):
    DC row = (sum rho)^2/N; a periodic shift leaves every shell sum unchanged; scaling the field
    by 2 scales I by exactly 4; dense and gather paths agree; first Bragg shells of the
    gyroid/BCC fields carry the intensity."""
    from simubox_b200 import Scatter, synth

    torch = torch_cuda
    L = synth.box_for(cells)
    f = synth.fields_torch(0, 2, (n, n, n), cells=cells)
    shifted = torch.roll(f, shifts=(5, n // 3, 1), dims=(1, 2, 3)).contiguous()
    outs = {p: Scatter.sacttering_peak_batch(f, L, path=p, max_batch=2)["I"].cpu().numpy() for p in (DENSE, GATHER, PRUNED, FUSED)}
    tol = 1e-9 * np.abs(outs[GATHER]) + 1e-12 * np.abs(outs[GATHER]).max()
    assert np.all(np.abs(outs[DENSE] - outs[GATHER]) <= tol)
    assert np.all(np.abs(outs[PRUNED] - outs[GATHER]) <= tol)
    assert np.all(np.abs(outs[FUSED] - outs[GATHER]) <= tol)
    again = Scatter.sacttering_peak_batch(f, L, path=FUSED, max_batch=1)["I"].cpu().numpy()
    assert np.array_equal(again, outs[FUSED]), "fused path must be bitwise reproducible whatever the unit schedule"
    Ip = Scatter.sacttering_peak_batch(shifted, L, path=FUSED, max_batch=2)["I"].cpu().numpy()
    assert np.all(np.abs(Ip[:, :-1] - outs[GATHER][:, :-1]) <= tol[:, :-1])
    I = outs[GATHER]
    dc = (f.sum(dim=(1, 2, 3)) ** 2 / n ** 3).cpu().numpy()
    assert np.allclose(I[:, 0], dc, rtol=1e-11)
    Is = Scatter.sacttering_peak_batch(shifted, L, path=GATHER, max_batch=2)["I"].cpu().numpy()
    assert np.all(np.abs(Is[:, :-1] - I[:, :-1]) <= 1e-9 * np.abs(I[:, :-1]) + 1e-12 * np.abs(I).max())
    I2 = Scatter.sacttering_peak_batch(2.0 * f, L, path=GATHER, max_batch=2)["I"].cpu().numpy()
    assert np.array_equal(I2, 4.0 * I)
    # both level sets (sin X cos Y + ... / cos X cos Y + ...) are built from the {110} star: |k|^2 = 2 (2 pi cells / L)^2
    q = Scatter.sacttering_peak_batch(f, L, path=GATHER, max_batch=2)["q"]
    unit = 2 * np.pi * cells / L[0]
    assert abs(q[1 + np.argmax(I[0, 1:])] / unit - np.sqrt(2)) < 1e-9
    assert abs(q[1 + np.argmax(I[1, 1:])] / unit - np.sqrt(2)) < 1e-9


@pytest.mark.parametrize("shape,L,cutoff", [((16, 16, 16), (4.0, 4.0, 4.0), 30), ((32, 24, 20), (8.0, 6.0, 5.0), 30),
                                            ((20, 32, 48), (3.0, 5.0, 7.0), 40), ((64, 64, 64), (8.0, 8.0, 8.0), 30),
                                            ((18, 40, 22), (2.0, 9.0, 4.0), 12), ((6, 16, 32), (2.0, 4.0, 8.0), 20),
                                            ((10, 24, 16), (5.0, 6.0, 4.0), 30)])
@pytest.mark.parametrize("path", [PRUNED, FUSED])
def test_pruned_dft_vs_oracle(sbx_lib, torch_cuda, shape, L, cutoff, path):
    """Hand-written pruned DFT (DMMA stages + finish kernel) against the numpy oracle, incl. non-cubic grids
    where the meshgrid 'xy' pairing (quirk Q1) decides which modes are needed."""
    rng = np.random.default_rng(sum(shape))
    mats = [rng.standard_normal(shape) for _ in range(3)]
    out = run_scatter(sbx_lib, torch_cuda, mats, L, cutoff, 1e-12, path, peaks=dict(width=0.5, step=500, cutoff=1e9, min_height=0.0))
    assert out["path"] == ("pruned_dft" if path == PRUNED else "pruned_dft_fused")
    for b, mat in enumerate(mats):
        ref, _ = scatter_np.scattering_peak_table(mat, L, cutoff, 1e-12)
        qI = np.stack([out["q"], out["I"][b].cpu().numpy()], axis=1)
        check_scatter_rows(qI, ref, cutoff, last_shell_members(mat, L, cutoff, 1e-12))


@pytest.mark.parametrize("shape", [(256, 256, 32), (384, 384, 16), (48, 48, 64), (128, 128, 128), (32, 32, 32),
                                   (64, 64, 128), (128, 128, 256), (256, 256, 128), (192, 192, 384)])
def test_blocked_pruned_kernel_vs_gather(sbx_lib, torch_cuda, monkeypatch, shape):
    """The row-blocked kernels (the forms used for 512^3 / 1024^3 planes and for slabs), forced on grids the oracle and
    the cuFFT gather path can check: 1-3 row blocks, in-kernel finish with peaks, two fields per launch.  Grids with
    N2 % 128 == 0 and blocks of 64 k rows run the bulk-staged kernel (TMA pieces of 1, 2 or 3 segments per half row);
    it keeps the summation order of the LDG blocked kernel, so the two must agree bit for bit."""
    from simubox_b200 import Scatter
    from simubox_b200.scatter import clear_plans

    monkeypatch.setenv("SBX_FUSED_BLOCKED", "1")  # the switches are read when a plan is created: drop the cached ones
    clear_plans()
    rng = np.random.default_rng(sum(shape))
    mats = [rng.random(shape) for _ in range(2)]
    L = (6.0, 6.0, 6.0)
    pk = dict(width=0.5, step=500, cutoff=1e9, min_height=0.0)
    out = run_scatter(sbx_lib, torch_cuda, mats, L, 30, 1e-12, FUSED, peaks=pk)
    from simubox_b200.scatter import get_plan
    yb_rows = 128 if shape[1] % 128 == 0 else shape[1]
    bulk = shape[2] % 128 == 0 and yb_rows % 64 == 0
    assert get_plan(shape, L, 30, 1e-12, max_batch=2, path=FUSED).kernel == ("pruned_bulk_kernel" if bulk else "pruned_blocked_kernel")
    if bulk:
        monkeypatch.setenv("SBX_BULK", "0")
        clear_plans()
        ldg = run_scatter(sbx_lib, torch_cuda, mats, L, 30, 1e-12, FUSED, peaks=pk)
        assert get_plan(shape, L, 30, 1e-12, max_batch=2, path=FUSED).kernel == "pruned_blocked_kernel"
        monkeypatch.delenv("SBX_BULK")
        assert np.array_equal(out["I"].cpu().numpy(), ldg["I"].cpu().numpy())
    monkeypatch.delenv("SBX_FUSED_BLOCKED")
    clear_plans()
    ref = run_scatter(sbx_lib, torch_cuda, mats, L, 30, 1e-12, GATHER, peaks=pk)
    a, b = out["I"].cpu().numpy(), ref["I"].cpu().numpy()
    assert np.all(np.abs(a[:, :-1] - b[:, :-1]) <= 1e-9 * np.abs(b[:, :-1]) + 1e-12 * np.abs(b).max())
    assert np.array_equal(out["n_peaks"].cpu().numpy(), ref["n_peaks"].cpu().numpy())
    for b in range(2):  # entries beyond a field's own count are unspecified
        n = int(out["n_peaks"][b])
        assert np.array_equal(out["peak_idx"][b, :n].cpu().numpy(), ref["peak_idx"][b, :n].cpu().numpy())
    if shape == (32, 32, 32):
        want, _ = scatter_np.scattering_peak_table(mats[0], L, 30, 1e-12)
        check_scatter_rows(np.stack([out["q"], a[0]], axis=1), want, 30, last_shell_members(mats[0], L, 30, 1e-12))


def test_giant_grid_uses_the_blocked_kernel(sbx_lib, torch_cuda):
    """512^3 (1 GB per field): AUTO picks the pruned path with the row-blocked (bulk-staged) kernel; checked against the cuFFT gather
    path and through size-independent properties (DC row, x4 under x2 scaling)."""
    from simubox_b200 import Scatter, synth
    from simubox_b200.scatter import get_plan

    torch = torch_cuda
    n, cells = 512, 8
    L = synth.box_for(cells)
    f = synth.fields_torch(0, 1, (n, n, n), cells=cells)
    out = Scatter.sacttering_peak_batch(f, L, max_batch=1)
    assert out["path"] == "pruned_dft_fused" and get_plan((n, n, n), L, 30, 1e-12, max_batch=1).kernel == "pruned_bulk_kernel"
    I = out["I"][0].cpu().numpy()
    ref = Scatter.sacttering_peak_batch(f, L, max_batch=1, path=GATHER)["I"][0].cpu().numpy()
    assert np.all(np.abs(I[:-1] - ref[:-1]) <= 1e-9 * np.abs(ref[:-1]) + 1e-12 * np.abs(ref).max())
    total = float(f.sum().item())
    assert abs(I[0] - total * total / n ** 3) <= 1e-9 * total * total / n ** 3
    I2 = Scatter.sacttering_peak_batch(2.0 * f, L, max_batch=1)["I"][0].cpu().numpy()
    assert np.array_equal(I2, 4.0 * I)


@pytest.mark.parametrize("width", [0.5, 0.25, 0.3, 1.7, 2.0])
def test_peak_width_forms_vs_oracle(sbx_lib, torch_cuda, width):
    """Gaussian broadening with power-of-two widths (exact reciprocal multiply on the device) and general widths
    (true division): plot_y within 1e-12 relative of the numpy oracle, peak indices exact; all four paths."""
    from simubox_b200 import synth

    shape, L = (32, 32, 32), synth.box_for(2)
    mat = synth.field_numpy(7, shape, cells=2)
    ref, _ = scatter_np.scattering_peak_table(mat, L, 30, 1e-12)
    kw = dict(width=width, step=700, cutoff=20, min_height=1e-6)
    want = scatter_np.show_peak(ref, **kw)
    for path in (DENSE, GATHER, PRUNED, FUSED):
        out = run_scatter(sbx_lib, torch_cuda, [mat], L, 30, 1e-12, path, peaks=kw)
        n = int(out["n_peaks"][0])
        assert np.array_equal(out["peak_idx"][0, :n].cpu().numpy(), want["peaks_index"]), path


def test_auto_path_selection(sbx_lib, torch_cuda):
    from simubox_b200.scatter import get_plan

    assert get_plan((64, 64, 64), (8, 8, 8), 30, 1e-12).path == FUSED           # few low modes: fused pruned DFT
    assert get_plan((40, 40, 40), (8, 8, 8), 30, 1e-12).path == PRUNED          # N2 % 16 != 0: two-kernel pruned DFT
    assert get_plan((64, 64, 64), (8, 8, 8), 60, 1e-12).path == GATHER          # modes beyond the 8x8 tile: cuFFT + gather
    assert get_plan((16, 16, 16), (4, 4, 4), 10 ** 6, 1e-12).path == DENSE      # every mode binned: cuFFT + dense scan
    assert get_plan((192, 192), (6.2, 6.2), 300, 1e-12).path in (GATHER, DENSE)  # 2D: cuFFT paths
