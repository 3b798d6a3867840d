"""torchrun worker: the NCCL slab path on WORLD_SIZE GPUs against the numpy oracle (rank 0)."""
import argparse
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    import torch
    import torch.distributed as dist

    from simubox_b200 import distributed as D, synth

    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=64)
    ap.add_argument("--cutoff", type=int, default=30)
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = D.rank_world()
    n = args.grid
    mat = synth.field_numpy(5, (n, n, n), cells=2)
    L = synth.box_for(2)
    slab = D.SlabScatter((n, n, n), L, cutoff=args.cutoff, device=local)
    want = os.environ.get("SBX_EXPECT_EXCHANGE")
    assert not want or slab.exchange == want, f"exchange is {slab.exchange}, expected {want}"
    lo, hi = slab.planes
    local_slab = torch.from_numpy(np.ascontiguousarray(mat[lo:hi])).cuda(local)
    qI = slab.q_intensity(local_slab)
    qI2 = slab.q_intensity(local_slab)  # plan re-use
    # the dense histogram accumulates with atomics: order, hence the last bits, may differ between runs
    assert np.array_equal(qI[:, 0], qI2[:, 0]) and np.allclose(qI[:, 1], qI2[:, 1], rtol=1e-12, atol=1e-12 * np.abs(qI[:, 1]).max())
    rows = [None] * world
    dist.all_gather_object(rows, qI[:, 1].tolist())
    assert all(r == rows[0] for r in rows), "ranks disagree on the all-reduced shell sums"
    if rank == 0:
        from oracle import scatter_np

        ref, _ = scatter_np.scattering_peak(mat, L, cutoff=args.cutoff)
        assert np.array_equal(qI[:, 0], ref[:, 0])
        tol = 1e-9 * np.abs(ref[:-1, 1]) + 1e-12 * np.abs(ref[:, 1]).max()
        assert np.all(np.abs(qI[:-1, 1] - ref[:-1, 1]) <= tol), "slab shell sums differ from the oracle"
        print(f"slab parity ok: {world} ranks, {n}^3, exchange {slab.exchange}")
    slab.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
