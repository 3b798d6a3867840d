"""GPU parity for the giant-grid path (csrc/slab.cu): the P-rank slab decomposition (2D D2Z ->
pack -> all-to-all -> 1D Z2Z -> fused shell histogram -> all-reduce) against the numpy oracle.
One GPU: all P ranks emulated on the device (``sbx_scatter_slab_emulate``, device memcpys as the
exchange).  >= 2 GPUs: the real NCCL path, one process per GPU under torchrun."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, check_scatter_rows
from oracle import scatter_np
from test_oracle_golden import last_shell_members

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,L,cutoff,world", [
    ((16, 16, 16), (4.0, 4.0, 4.0), 30, 1), ((16, 16, 16), (4.0, 4.0, 4.0), 30, 2), ((16, 16, 16), (4.0, 4.0, 4.0), 30, 4),
    ((8, 12, 10), (1.0, 2.0, 3.0), 10, 2),     # non-cubic, unequal box: quirk Q1 under the row remap
    ((12, 8, 6), (2.0, 1.0, 1.5), 30, 4), ((32, 32, 32), (8.0, 8.0, 8.0), 400, 8),
    ((64, 64, 64), (8.0, 8.0, 8.0), 30, 8)])
@pytest.mark.parametrize("slab_path", ["fft", "auto"])
def test_emulated_slab_matches_oracle(sbx_lib, monkeypatch, N, L, cutoff, world, slab_path):
    """"fft": the slab FFT + exchange for every case; "auto": the pruned-DFT slab mode wherever it applies
    (low modes only, N2 % 16 == 0, planes per rank divisible by the unit size), the FFT mode elsewhere."""
    from simubox_b200 import distributed as D

    if slab_path == "fft":
        monkeypatch.setenv("SBX_SLAB_PATH", "fft")
    else:
        monkeypatch.delenv("SBX_SLAB_PATH", raising=False)
    rng = np.random.default_rng(hash((N, world)) % 2 ** 32)
    mat = rng.random(N)
    I = D.slab_emulate(mat, L, world, cutoff=cutoff)
    ref, _ = scatter_np.scattering_peak(mat, L, cutoff=cutoff)
    plan = sbx_lib.ScatterPlan(N, L, cutoff, 1e-12, 1, sbx_lib.SBX_DEVICE_NONE)
    qI = np.stack([plan.q, I], axis=1)
    check_scatter_rows(qI, ref, cutoff, last_shell_members(mat, L, cutoff, 1e-12))


@pytest.mark.parametrize("chunks,allcols", [(1, 0), (3, 0), (8, 1), (16, 0), (2, 1)])
@pytest.mark.parametrize("N,L,cutoff,world", [((32, 32, 32), (8.0, 8.0, 8.0), 400, 2), ((12, 8, 6), (2.0, 1.0, 1.5), 30, 4),
                                              ((48, 16, 64), (3.0, 2.0, 5.0), 60, 2), ((16, 16, 16), (4.0, 4.0, 4.0), 5, 1)])
def test_emulated_slab_fft_chunks_and_column_pruning(sbx_lib, monkeypatch, N, L, cutoff, world, chunks, allcols):
    """Slab FFT mode with the local planes cut into 1 ... 16 chunks (the count is lowered to a divisor of the planes
    per rank) and with / without the pruning of the exchanged columns (only c with kv2[c] <= the largest listed k^2
    travel): the rows must not depend on either."""
    from simubox_b200 import distributed as D

    monkeypatch.setenv("SBX_SLAB_PATH", "fft")
    monkeypatch.setenv("SBX_SLAB_CHUNKS", str(chunks))
    monkeypatch.setenv("SBX_SLAB_ALLCOLS", str(allcols))
    rng = np.random.default_rng(hash((N, world)) % 2 ** 32)
    mat = rng.random(N)
    I = D.slab_emulate(mat, L, world, cutoff=cutoff)
    ref, _ = scatter_np.scattering_peak(mat, L, cutoff=cutoff)
    plan = sbx_lib.ScatterPlan(N, L, cutoff, 1e-12, 1, sbx_lib.SBX_DEVICE_NONE)
    check_scatter_rows(np.stack([plan.q, I], axis=1), ref, cutoff, last_shell_members(mat, L, cutoff, 1e-12))


def test_emulated_slab_equals_single_gpu_plan(sbx_lib):
    import torch

    from simubox_b200 import Scatter, distributed as D, synth

    mat = synth.field_numpy(3, (64, 64, 64), cells=2)
    L = synth.box_for(2)
    one = Scatter.sacttering_peak_batch(torch.from_numpy(mat[None]).cuda(), L, path=1)["I"][0].cpu().numpy()
    for world in (2, 4):
        I = D.slab_emulate(mat, L, world)
        tol = 1e-9 * np.abs(one[:-1]) + 1e-12 * np.abs(one).max()
        assert np.all(np.abs(I[:-1] - one[:-1]) <= tol)


@pytest.mark.parametrize("N,world", [((256, 256, 32), 4), ((384, 384, 16), 2), ((48, 48, 64), 4), ((128, 128, 128), 8)])
def test_pruned_slab_mode_blocked_rows(sbx_lib, monkeypatch, N, world):
    """Pruned slab mode forced (a plan error if it does not apply): row-blocked planes (N1 = 2 or 3 blocks of 128 rows,
    or one odd-sized block; N0 = N1, else quirk Q1 scrambles the mode set), x_base offsets of up to 8 ranks, against the cuFFT gather path of the single-GPU plan."""
    import torch

    from simubox_b200 import Scatter, distributed as D

    monkeypatch.setenv("SBX_SLAB_PATH", "pruned")
    rng = np.random.default_rng(sum(N))
    mat = rng.random(N)
    L = (6.0, 6.0, 6.0)
    I = D.slab_emulate(mat, L, world)
    one = Scatter.sacttering_peak_batch(torch.from_numpy(mat[None]).cuda(), L, path=2)["I"][0].cpu().numpy()
    tol = 1e-9 * np.abs(one[:-1]) + 1e-12 * np.abs(one).max()
    assert np.all(np.abs(I[:-1] - one[:-1]) <= tol)


def test_slab_rejects_bad_partition(sbx_lib):
    from simubox_b200 import distributed as D

    with pytest.raises(sbx_lib.SbxError) as ei:
        D.slab_emulate(np.zeros((12, 8, 8)), (1, 1, 1), 8)
    assert ei.value.code == -1


@pytest.mark.parametrize("port,slab_path,extra", [
    (29517, "fft", {"SBX_EXPECT_EXCHANGE": "p2p_fused"}), (29518, "pruned", {"SBX_EXPECT_EXCHANGE": "none"}),
    (29519, "fft", {"SBX_SLAB_CHUNKS": "8", "SBX_SLAB_ALLCOLS": "1", "SBX_EXPECT_EXCHANGE": "p2p_fused"}),
    (29520, "fft", {"SBX_SLAB_CHUNKS": "1", "SBX_EXPECT_EXCHANGE": "p2p_fused"}),
    (29521, "fft", {"SBX_SLAB_EXCHANGE": "nccl", "SBX_EXPECT_EXCHANGE": "nccl_send_recv"}),
    (29522, "fft", {"SBX_SLAB_EXCHANGE": "nccl", "SBX_SLAB_CHUNKS": "8", "SBX_SLAB_ALLCOLS": "1", "SBX_EXPECT_EXCHANGE": "nccl_send_recv"})])
def test_nccl_slab_two_ranks(port, slab_path, extra):
    """The real multi-process path, one process per GPU.  Slab FFT mode with the fused pack + exchange over peer memory
    (default: CUDA IPC-mapped receive buffers, chunk pushes on a second stream overlapping the 2D transforms; also with 8
    chunks and every column exchanged, and unchunked), the same with the NCCL send/recv exchange, and pruned mode
    (all-reduce only).  The worker asserts which exchange the plan ended up with."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (the single-GPU emulation test covers the decomposition)")
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(ROOT / "tests" / "slab_worker.py"), "--grid", "64"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=str(ROOT), env=dict(os.environ, SBX_SLAB_PATH=slab_path, **extra))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "slab parity ok" in res.stdout
