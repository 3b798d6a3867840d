"""CPU suite for the N>1 host logic: world_size-2 ``gloo`` process group (no GPU).  Covers the
sharding of a field sweep over ranks, the all-gather of the per-field result rows, and the slab
partition arithmetic the giant-grid path uses (the NCCL transport itself is GPU-only)."""
import os
import socket

import numpy as np
import pytest

from simubox_b200 import distributed as D


def test_shard_partitions_are_exact_covers():
    for total in (0, 1, 7, 256, 2048):
        for world in (1, 2, 3, 8):
            rr = np.concatenate([D.shard_indices(total, r, world) for r in range(world)])
            assert sorted(rr.tolist()) == list(range(total))
            spans = [D.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        D.shard_indices(4, 2, 2)
    assert D.slab_planes(1024, 3, 8) == (384, 512)
    with pytest.raises(ValueError):
        D.slab_planes(10, 0, 4)


def test_gather_rows_without_process_group():
    rows = np.arange(6.0).reshape(3, 2)
    out = D.gather_rows(rows, [0, 1, 2], 3)
    assert np.array_equal(out, rows)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert D.rank_world() == (rank, world)
        mine = D.shard_indices(total, rank, world)
        # stand-in for the per-field GPU result: row i = [i, i^2, rank that computed it]
        rows = np.stack([mine.astype(float), mine.astype(float) ** 2, np.full(len(mine), float(rank))], axis=1)
        full = D.gather_rows(rows, mine, total)
        q.put((rank, full))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [5, 8])
def test_gloo_world2_sweep_gather(total):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    i = np.arange(total, dtype=float)
    want = np.stack([i, i ** 2, i % 2], axis=1)
    for rank in (0, 1):
        assert np.array_equal(got[rank], want), f"rank {rank} sees a different gathered table"


def _planes_worker(rank, world, port, shape, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = D.shard_range(shape[0], rank, world)
        # stand-in for the labels of this rank's planes: value = global linear voxel index
        full = np.arange(int(np.prod(shape)), dtype=np.int32).reshape(shape)
        q.put((rank, D.gather_planes(full[lo:hi].copy(), (lo, hi), shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("shape", [(7, 3, 5), (8, 6)])
def test_gloo_world2_voxel_slab_gather(shape):
    """Partition logic of the sharded Voronoi labelling (uneven plane counts included): every rank ends up with
    the whole label grid."""
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_planes_worker, args=(r, 2, port, shape, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = np.arange(int(np.prod(shape)), dtype=np.int32).reshape(shape)
    for rank in (0, 1):
        assert np.array_equal(got[rank], want)


def test_gather_planes_single_process_and_errors():
    a = np.arange(24, dtype=np.int32).reshape(4, 6)
    assert np.array_equal(D.gather_planes(a, (0, 4), (4, 6)), a)
    with pytest.raises(ValueError):
        D.gather_planes(a[:2], (0, 3), (4, 6))
    with pytest.raises(ValueError):
        D.gather_planes(a[:2].copy(), (0, 2), (4, 6))
