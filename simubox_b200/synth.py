"""Synthetic density fields and domain-centre seeds for tests and ``bench.py``.

Closed-form gyroid / BCC three-component fields as specified in SURVEY.md 8(d) config 2,
plus counter-based noise.  ``field_numpy`` is the definition; ``fields_torch`` produces the
same shapes on a torch device for bulk benchmark data (torch is only the buffer carrier /
data generator here -- it is not on the analysis path).  The two generators agree to
~1e-15 (libm vs CUDA sin/cos), so parity tests always feed BOTH arms the same bytes.
"""
from __future__ import annotations

import numpy as np

_MASK64 = (1 << 64) - 1


def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def noise_numpy(index: int, n: int) -> np.ndarray:
    """hash64(index, voxel) / 2^64 - 0.5 for voxel = 0..n-1 (fp64)."""
    with np.errstate(over="ignore"):
        v = np.arange(n, dtype=np.uint64) + np.uint64((index * 0x632BE59BD9B4E019) & _MASK64)
        h = _splitmix64(v)
    return (h >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53)) - 0.5


def field_params(index: int):
    t = 0.6 + 0.4 * (index % 16) / 16
    w = 0.05 + 0.1 * ((index // 16) % 16) / 16
    return t, w


def level_set_numpy(index: int, shape, cells: int = 2) -> np.ndarray:
    axes = [2 * np.pi * cells * np.arange(n) / n for n in shape]
    X, Y, Z = np.meshgrid(*axes, indexing="ij")
    if index % 2 == 0:  # gyroid
        return np.sin(X) * np.cos(Y) + np.sin(Y) * np.cos(Z) + np.sin(Z) * np.cos(X)
    return np.cos(X) * np.cos(Y) + np.cos(Y) * np.cos(Z) + np.cos(Z) * np.cos(X)  # BCC


def field_numpy(index: int, shape=(128, 128, 128), cells: int = 2, component: int = 0, noise: float = 1e-3):
    """phi_A (component 0), phi_B (1) or phi_C (2) of synthetic field ``index``; fp64 C order."""
    s = level_set_numpy(index, shape, cells)
    t, w = field_params(index)
    sig = lambda v: 1.0 / (1.0 + np.exp(-v))
    phiA = sig((s - t) / w) + noise * noise_numpy(index, s.size).reshape(shape)
    if component == 0:
        return np.ascontiguousarray(phiA)
    phiC = sig((-s - t) / w)
    if component == 2:
        return np.ascontiguousarray(phiC)
    return np.ascontiguousarray(1.0 - phiA - phiC)


def box_for(cells: int = 2, ndim: int = 3):
    """L = 4.0 * cells per axis."""
    return np.full(ndim, 4.0 * cells)


def fields_torch(first: int, count: int, shape=(128, 128, 128), cells: int = 2, device="cuda", noise: float = 1e-3):
    """(count, *shape) fp64 tensor of phi_A of fields first..first+count-1, generated on ``device``."""
    import torch

    dev = torch.device(device)
    axes = [2 * torch.pi * cells * torch.arange(n, device=dev, dtype=torch.float64) / n for n in shape]
    X, Y, Z = torch.meshgrid(*axes, indexing="ij")
    gyr = torch.sin(X) * torch.cos(Y) + torch.sin(Y) * torch.cos(Z) + torch.sin(Z) * torch.cos(X)
    bcc = torch.cos(X) * torch.cos(Y) + torch.cos(Y) * torch.cos(Z) + torch.cos(Z) * torch.cos(X)
    n = gyr.numel()
    out = torch.empty((count, *shape), device=dev, dtype=torch.float64)
    vox = torch.arange(n, device=dev, dtype=torch.int64)

    def wrap(c):  # two's-complement view of a uint64 constant
        return c - (1 << 64) if c >= (1 << 63) else c

    def lsr(z, k):  # logical shift right on int64
        return (z >> k) & ((1 << (64 - k)) - 1)

    for j in range(count):
        idx = first + j
        t, w = field_params(idx)
        s = gyr if idx % 2 == 0 else bcc
        z = vox + wrap((idx * 0x632BE59BD9B4E019) & _MASK64) + wrap(0x9E3779B97F4A7C15)
        z = (z ^ lsr(z, 30)) * wrap(0xBF58476D1CE4E5B9)
        z = (z ^ lsr(z, 27)) * wrap(0x94D049BB133111EB)
        z = z ^ lsr(z, 31)
        u = lsr(z, 11).to(torch.float64) * (1.0 / (1 << 53)) - 0.5
        out[j] = torch.sigmoid((s - t) / w) + noise * u.reshape(shape)
    return out


def field_planes_torch(index: int, shape, x0: int, x1: int, cells: int = 2, device="cuda", noise: float = 1e-3):
    """Planes x0..x1-1 of field ``index`` as a (x1-x0, N1, N2) fp64 tensor: the same values ``fields_torch``
    produces, generated slab by slab (a 1024^3 field never exists in one piece on one rank)."""
    import torch

    dev = torch.device(device)
    N0, N1, N2 = (int(n) for n in shape)
    ax = [2 * torch.pi * cells * torch.arange(n, device=dev, dtype=torch.float64) / n for n in (N0, N1, N2)]
    X, Y, Z = torch.meshgrid(ax[0][x0:x1], ax[1], ax[2], indexing="ij")
    if index % 2 == 0:
        s = torch.sin(X) * torch.cos(Y) + torch.sin(Y) * torch.cos(Z) + torch.sin(Z) * torch.cos(X)
    else:
        s = torch.cos(X) * torch.cos(Y) + torch.cos(Y) * torch.cos(Z) + torch.cos(Z) * torch.cos(X)
    t, w = field_params(index)
    vox = torch.arange(x0 * N1 * N2, x1 * N1 * N2, device=dev, dtype=torch.int64)

    def wrap(c):
        return c - (1 << 64) if c >= (1 << 63) else c

    def lsr(z, k):
        return (z >> k) & ((1 << (64 - k)) - 1)

    z = vox + wrap((index * 0x632BE59BD9B4E019) & _MASK64) + wrap(0x9E3779B97F4A7C15)
    z = (z ^ lsr(z, 30)) * wrap(0xBF58476D1CE4E5B9)
    z = (z ^ lsr(z, 27)) * wrap(0x94D049BB133111EB)
    z = z ^ lsr(z, 31)
    u = lsr(z, 11).to(torch.float64) * (1.0 / (1 << 53)) - 0.5
    return torch.sigmoid((s - t) / w) + noise * u.reshape(x1 - x0, N1, N2)


def seeds_lattice(shape=(256, 256, 256), per_axis: int = 16, jitter: float = 4.0, seed: int = 1234):
    """Jittered lattice of domain centres: centres at
    spacing*(a+1/2) voxel units plus uniform jitter in [-jitter, jitter), a-major order."""
    d = len(shape)
    sp = [n / per_axis for n in shape]
    grids = np.meshgrid(*[sp[a] * (np.arange(per_axis) + 0.5) for a in range(d)], indexing="ij")
    pts = np.stack([g.reshape(-1) for g in grids], axis=1).astype(np.float64)
    if jitter:
        rng = np.random.default_rng(seed)
        pts = pts + rng.uniform(-jitter, jitter, size=pts.shape)
    return np.ascontiguousarray(pts)
