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
def test_emulated_slab_matches_oracle(sbx_lib, N, L, cutoff, world):
    from simubox_b200 import distributed as D

    rng = np.random.default_rng(hash((N, world)) % 2 ** 32)
    mat = rng.random(N)
    I = D.slab_emulate(mat, L, world, cutoff=cutoff)
    ref, _ = scatter_np.scattering_peak(mat, L, cutoff=cutoff)
    plan = sbx_lib.ScatterPlan(N, L, cutoff, 1e-12, 1, sbx_lib.SBX_DEVICE_NONE)
    qI = np.stack([plan.q, I], axis=1)
    check_scatter_rows(qI, ref, cutoff, last_shell_members(mat, L, cutoff, 1e-12))


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


def test_slab_rejects_bad_partition(sbx_lib):
    from simubox_b200 import distributed as D

    with pytest.raises(sbx_lib.SbxError) as ei:
        D.slab_emulate(np.zeros((12, 8, 8)), (1, 1, 1), 8)
    assert ei.value.code == -1


def test_nccl_slab_two_ranks():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (the single-GPU emulation test covers the decomposition)")
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", str(ROOT / "tests" / "slab_worker.py"), "--grid", "64"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "slab parity ok" in res.stdout
