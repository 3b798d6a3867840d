"""Multi-GPU host logic: one process per GPU (``torch.distributed`` is the plumbing only).

Two cases:

* batches / sweeps of independent fields -- ``shard_range`` / ``shard_indices`` give each rank its
  fields, ``gather_rows`` brings the small per-field results to every rank.  No data-path
  collective.
* one Voronoi grid -- ``label_grid_sharded``: seeds replicated, rank r labels the planes ``shard_range(N0, r, P)``
  of axis 0 (``sbx_voronoi_plan_execute_range``); labels stay sharded, ``gather_planes`` all-gathers them if a full
  copy is wanted.  No collective in the labelling itself.
* one giant field -- ``SlabScatter`` owns an NCCL communicator created by libsimubox_b200.so
  (``sbx_nccl_unique_id`` on rank 0, broadcast through the existing process group,
  ``sbx_nccl_comm_create`` everywhere) and runs the slab-decomposed FFT with an all-to-all
  (``sbx_scatter_slab_execute``, csrc/slab.cu).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence, Tuple

import numpy as np

from . import _lib


def rank_world() -> Tuple[int, int]:
    """(rank, world) from an initialised process group, else from the torchrun environment."""
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def shard_indices(total: int, rank: int, world: int) -> np.ndarray:
    """Round-robin sharding of a sweep: field i goes to rank ``i % world``."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    return np.arange(rank, int(total), world, dtype=np.int64)


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous sharding: rank r owns [lo, hi); the first ``total % world`` ranks get one extra."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    base, extra = divmod(int(total), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def slab_planes(N0: int, rank: int, world: int) -> Tuple[int, int]:
    """Planes [lo, hi) of axis 0 owned by ``rank`` in the slab decomposition (N0 % world == 0)."""
    if N0 % world:
        raise ValueError(f"N0={N0} must be divisible by the number of ranks ({world})")
    n0 = N0 // world
    return rank * n0, (rank + 1) * n0


def gather_rows(local_rows: np.ndarray, indices: Sequence[int], total: int) -> np.ndarray:
    """All-gather per-field result rows (small: ~0.5 kB per field) into a (total, width) array on
    every rank.  ``indices`` are the global field numbers of ``local_rows``.  Works on any
    backend (gloo on CPU, nccl on GPU); with no process group it is the identity."""
    import torch
    import torch.distributed as dist

    local_rows = np.ascontiguousarray(local_rows, dtype=np.float64)
    if local_rows.ndim == 1:
        local_rows = local_rows[:, None]
    width = local_rows.shape[1]
    out = np.zeros((int(total), width), dtype=np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        out[np.asarray(indices, dtype=np.int64)] = local_rows
        return out
    world = dist.get_world_size()
    per = (int(total) + world - 1) // world
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    rows = torch.zeros((per, width), dtype=torch.float64, device=dev)
    idx = torch.full((per,), -1, dtype=torch.int64, device=dev)
    n = len(indices)
    rows[:n] = torch.from_numpy(local_rows).to(dev)
    idx[:n] = torch.as_tensor(np.asarray(indices, dtype=np.int64)).to(dev)
    all_rows = [torch.empty_like(rows) for _ in range(world)]
    all_idx = [torch.empty_like(idx) for _ in range(world)]
    dist.all_gather(all_rows, rows)
    dist.all_gather(all_idx, idx)
    for r, i in zip(all_rows, all_idx):
        i = i.cpu().numpy()
        keep = i >= 0
        out[i[keep]] = r.cpu().numpy()[keep]
    return out


def label_grid_sharded(points, weights=0, shape: Sequence[int] = (256, 256, 256), method="additive", box=None, mask=None,
                       device: int | None = None, gather: bool = False, rank: int | None = None, world: int | None = None):
    """Voronoi labelling of ONE grid spread over the ranks by voxel slab (the loop it replaces is
    SimuBox/calculator/voronoi.py:284-292, split over the first pixel coordinate).  Every rank passes the complete
    ``points`` (and ``weights``); ``mask`` may be the whole-grid mask or this rank's planes.  Returns
    ``(labels_local, (x_lo, x_hi))`` with ``labels_local`` a CUDA int32 tensor of shape ``(x_hi - x_lo, *shape[1:])``;
    with ``gather=True`` the full ``shape`` array (numpy) on every rank instead of the local tensor."""
    from .voronoi import VoronoiCell

    r, w = rank_world()
    rank = r if rank is None else int(rank)
    world = w if world is None else int(world)
    dev = int(device if device is not None else os.environ.get("LOCAL_RANK", "0"))
    lo, hi = shard_range(int(shape[0]), rank, world)
    if hi <= lo:
        raise ValueError(f"rank {rank} of {world} owns no plane of a grid with {shape[0]} planes")
    local = VoronoiCell.label_grid(points, weights, shape=shape, method=method, box=box, mask=mask, device=dev,
                                   return_tensor=True, planes=(lo, hi))
    if not gather:
        return local, (lo, hi)
    return gather_planes(local.cpu().numpy(), (lo, hi), tuple(int(n) for n in shape)), (lo, hi)


def gather_planes(local: np.ndarray, planes: Tuple[int, int], shape: Sequence[int]) -> np.ndarray:
    """All-gather the plane slabs of a sharded grid into the full array on every rank (any backend; identity
    without a process group).  Slabs may differ by one plane (``shard_range``): they are padded to the largest."""
    import torch
    import torch.distributed as dist

    shape = tuple(int(n) for n in shape)
    local = np.ascontiguousarray(local)
    lo, hi = int(planes[0]), int(planes[1])
    if tuple(local.shape) != (hi - lo,) + shape[1:]:
        raise ValueError(f"slab shape {local.shape} does not match planes {planes} of {shape}")
    out = np.empty(shape, dtype=local.dtype)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        if (lo, hi) != (0, shape[0]):
            raise ValueError("a single process must hold every plane")
        out[:] = local
        return out
    world = dist.get_world_size()
    per = (shape[0] + world - 1) // world
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    buf = torch.zeros((per,) + shape[1:], dtype=torch.from_numpy(local).dtype, device=dev)
    buf[:hi - lo] = torch.from_numpy(local).to(dev)
    span = torch.tensor([lo, hi], dtype=torch.int64, device=dev)
    bufs = [torch.empty_like(buf) for _ in range(world)]
    spans = [torch.empty_like(span) for _ in range(world)]
    dist.all_gather(bufs, buf)
    dist.all_gather(spans, span)
    for b, sp in zip(bufs, spans):
        a, e = (int(v) for v in sp.cpu().tolist())
        out[a:e] = b[:e - a].cpu().numpy()
    return out


def _nccl_library_hint():
    """Point the dlopen in csrc/slab.cu at torch's bundled NCCL when nothing else is set."""
    if os.environ.get("SBX_NCCL_LIB"):
        return
    try:
        import nvidia.nccl as n

        for base in list(getattr(n, "__path__", [])):
            cand = os.path.join(base, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["SBX_NCCL_LIB"] = cand
                return
    except Exception:
        pass


class SlabScatter:
    """Scattering rows of ONE field spread over the ranks of the current process group."""

    def __init__(self, N, L, cutoff: int = 30, threshold: float = 1e-12, device: int | None = None):
        import torch
        import torch.distributed as dist

        self.rank, self.world = rank_world()
        self.device = int(device if device is not None else os.environ.get("LOCAL_RANK", "0"))
        self.N = tuple(int(n) for n in N)
        if len(self.N) != 3:
            raise ValueError("the slab decomposition is for 3D fields")
        L_ = _lib.lib()
        self._comm = C.c_void_p()
        self._h = C.c_void_p()
        if self.world > 1:
            if not (dist.is_available() and dist.is_initialized()):
                raise RuntimeError("SlabScatter with WORLD_SIZE > 1 needs torch.distributed.init_process_group first")
            _nccl_library_hint()
            ident = [None]
            if self.rank == 0:
                buf = (C.c_char * 128)()
                _lib.check(L_.sbx_nccl_unique_id(buf))
                ident[0] = bytes(buf.raw)
            dist.broadcast_object_list(ident, src=0)
            buf = (C.c_char * 128).from_buffer_copy(ident[0])
            torch.cuda.set_device(self.device)
            _lib.check(L_.sbx_nccl_comm_create(C.byref(self._comm), buf, self.rank, self.world, self.device))
        _, pn = _lib.i64(self.N)
        _, pl = _lib.f64([float(l) for l in L])
        _lib.check(L_.sbx_scatter_slab_plan_create(C.byref(self._h), self._comm, self.rank, self.world, self.device,
                                                   pn, pl, int(cutoff), float(threshold)))
        n = C.c_int()
        _lib.check(L_.sbx_scatter_slab_shells(self._h, C.byref(n), None, None))
        self.n_shells = n.value
        self.q = np.empty(self.n_shells)
        self.count = np.empty(self.n_shells, dtype=np.int64)
        _lib.check(L_.sbx_scatter_slab_shells(self._h, C.byref(n), self.q.ctypes.data_as(C.POINTER(C.c_double)),
                                              self.count.ctypes.data_as(C.POINTER(C.c_int64))))
        self.planes = slab_planes(self.N[0], self.rank, self.world)
        # 4: pruned DFT (mode sums + one small all-reduce); 1: slab FFT with all-to-all.  SBX_SLAB_PATH=fft|pruned forces one.
        self.path = int(L_.sbx_scatter_slab_path(self._h))
        # slab FFT mode: map every rank's receive buffer (CUDA IPC) so that the pack kernel stores straight into the
        # peers over NVLink (fused pack + exchange).  All ranks must agree: if the mapping fails anywhere, every rank
        # detaches and the NCCL send/recv exchange is used.  SBX_SLAB_EXCHANGE=nccl skips the attempt.
        if self.world > 1 and self.path != _lib.PATH_PRUNED_FUSED and os.environ.get("SBX_SLAB_EXCHANGE", "p2p") != "nccl":
            hb = (C.c_char * 64)()
            ok = L_.sbx_scatter_slab_ipc_export(self._h, hb) == 0
            blobs = [None] * self.world
            dist.all_gather_object(blobs, bytes(hb.raw) if ok else None)
            if all(b is not None for b in blobs):
                allh = (C.c_char * (64 * self.world)).from_buffer_copy(b"".join(blobs))
                ok = L_.sbx_scatter_slab_ipc_attach(self._h, allh) == 0
            else:
                ok = False
            flag = torch.tensor([1 if ok else 0], device=torch.device("cuda", self.device), dtype=torch.int32)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 0:
                L_.sbx_scatter_slab_ipc_detach(self._h)
        self.exchange = {0: "none", 1: "nccl_send_recv", 2: "p2p_fused"}.get(int(L_.sbx_scatter_slab_exchange(self._h)), "?")

    @property
    def kernel(self) -> str:
        """Dominant hand-written kernel of the last execute in pruned mode ('' in slab FFT mode)."""
        return (_lib.lib().sbx_scatter_slab_kernel(self._h) or b"").decode()

    def execute(self, local_slab, out=None):
        """``local_slab``: CUDA fp64 tensor (N0/world, N1, N2) = this rank's planes.  Returns the
        (n_shells,) CUDA tensor of shell intensities (identical on every rank); asynchronous on
        the current stream."""
        import torch

        lo, hi = self.planes
        want = (hi - lo, self.N[1], self.N[2])
        if tuple(local_slab.shape) != want or local_slab.dtype != torch.float64 or not local_slab.is_cuda:
            raise ValueError(f"local_slab must be a CUDA float64 tensor of shape {want}")
        slab = local_slab.contiguous()
        if out is None:
            out = torch.empty(self.n_shells, dtype=torch.float64, device=slab.device)
        st = C.c_void_p(torch.cuda.current_stream(slab.device).cuda_stream)
        _lib.check(_lib.lib().sbx_scatter_slab_execute(self._h, C.c_void_p(slab.data_ptr()), C.c_void_p(out.data_ptr()), st))
        return out

    def q_intensity(self, local_slab) -> np.ndarray:
        """(n_shells, 2) array [q, I] like ``ScatterResult.q_Intensity``."""
        I = self.execute(local_slab).cpu().numpy()
        return np.stack([self.q, I], axis=1)

    def set_profiling(self, enable: bool):
        _lib.check(_lib.lib().sbx_scatter_slab_set_profiling(self._h, int(enable)))

    def profile(self) -> dict:
        ms = (C.c_double * 4)()
        _lib.check(_lib.lib().sbx_scatter_slab_profile(self._h, ms))
        if self.path == _lib.PATH_PRUNED_FUSED:
            return dict(pruned_planes_ms=ms[0], all_reduce_ms=ms[1], finish_ms=ms[3])
        return dict(fft2d_pack_ms=ms[0], all_to_all_ms=ms[1], fft1d_ms=ms[2], histogram_allreduce_ms=ms[3])

    def close(self):
        L_ = _lib.lib()
        if self._h:
            if getattr(self, "exchange", "") == "p2p_fused":
                # unmap the peers' buffers on every rank before any rank frees its own (the exporter must outlive the mappings)
                L_.sbx_scatter_slab_ipc_detach(self._h)
                try:
                    import torch.distributed as dist

                    if dist.is_available() and dist.is_initialized():
                        dist.barrier()
                except Exception:
                    pass
                self.exchange = "nccl_send_recv"
            L_.sbx_scatter_slab_plan_destroy(self._h)
            self._h = C.c_void_p()
        if self._comm:
            L_.sbx_nccl_comm_destroy(self._comm)
            self._comm = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def slab_emulate(field, L, world: int, cutoff: int = 30, threshold: float = 1e-12, device: int = 0) -> np.ndarray:
    """Run the ``world``-rank slab decomposition on ONE GPU (``sbx_scatter_slab_emulate``): the
    single-device check of csrc/slab.cu.  ``field``: full (N0, N1, N2) fp64 array or CUDA tensor.
    Returns the (n_shells,) intensities as numpy."""
    from ._device import ptr, stream_ptr, to_device, torch_cuda

    torch = torch_cuda()
    f = to_device(field, device, torch.float64)
    N = tuple(int(n) for n in f.shape)
    plan = _lib.ScatterPlan(N, L, cutoff, threshold, 1, _lib.SBX_DEVICE_NONE)
    out = torch.empty(plan.n_shells, dtype=torch.float64, device=f.device)
    _, pn = _lib.i64(N)
    _, pl = _lib.f64([float(l) for l in L])
    _lib.check(_lib.lib().sbx_scatter_slab_emulate(device, int(world), pn, pl, int(cutoff), float(threshold), ptr(f), ptr(out),
                                                   stream_ptr(device)))
    plan.close()
    return out.cpu().numpy()


def gpu_numa_cpus(device: int):
    """CPUs local to ``cuda:device`` according to sysfs (``/sys/bus/pci/devices/<bus id>/local_cpulist``), with the NUMA
    node (-1 = the platform reports none).  Returns (sorted cpu list, node) or ([], -1) when sysfs has no answer."""
    import torch

    try:
        prop = torch.cuda.get_device_properties(device)
        bus = f"{getattr(prop, 'pci_domain_id', 0):04x}:{prop.pci_bus_id:02x}:{prop.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bus}"
        with open(f"{base}/local_cpulist") as fh:
            text = fh.read().strip()
        try:
            with open(f"{base}/numa_node") as fh:
                node = int(fh.read().strip())
        except Exception:
            node = -1
        cpus = []
        for part in filter(None, text.split(",")):
            a, _, b = part.partition("-")
            cpus.extend(range(int(a), int(b or a) + 1))
        return sorted(cpus), node
    except Exception:
        return [], -1


def bind_to_gpu_numa(device: int) -> dict:
    """Pin this process to the CPUs local to its GPU BEFORE it allocates pinned host memory (first touch then places
    the staging buffers on the GPU's NUMA node).  No-op when sysfs gives no list or the list does not intersect the
    CPUs the process may use.  Returns what was done, for the bench record."""
    cpus, node = gpu_numa_cpus(device)
    allowed = os.sched_getaffinity(0)
    use = sorted(set(cpus) & allowed)
    info = {"numa_node": node, "local_cpus": len(cpus), "allowed_cpus": len(allowed), "bound": False}
    if use and len(use) < len(allowed):
        os.sched_setaffinity(0, use)
        info["bound"] = True
        info["bound_cpus"] = len(use)
    return info
