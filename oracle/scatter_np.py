"""CPU restatement of the reference scattering path -- TEST INFRASTRUCTURE ONLY.

Follows ``/root/reference/SimuBox/calculator/scatter.py`` operation by operation
(numpy/scipy, fp64).  Pinned against the live reference by ``oracle/make_golden.py``
and ``tests/test_oracle_golden.py`` (every golden, plus the live reference when present); fixtures in ``tests/golden/``.

Two restatements are kept side by side:

* ``scattering_peak``       -- literal: full-spectrum FFT, argsort, Python merge loop
                               (``scatter.py:78-120``).  This is what ``bench.py`` times
                               as the CPU baseline ("port").
* ``scattering_peak_table`` -- the sort-free shell-table formulation the CUDA path uses
                               (field-independent shell geometry + binning), written
                               independently in numpy so that the GPU kernels can be
                               compared against something that shares no code with them.
"""
from __future__ import annotations

from itertools import product

import numpy as np


# --------------------------------------------------------------------------- k tables
def k_vector(N: int) -> np.ndarray:
    """``Scatter.k_vector`` (scatter.py:40-47): [0..N/2, -N/2+1..-1] as float64."""
    if -N / 2 + 1 > -1:
        return np.arange(0, N / 2 + 0.5, 1)
    if -N / 2 + 1 == -1:
        return np.hstack([np.arange(0, N / 2 + 0.5, 1), np.array([-1])])
    return np.hstack([np.arange(0, N / 2 + 0.5, 1), np.arange(-N / 2 + 1, 0, 1)])


def k_space(NxNyNz, lxlylz):
    """``Scatter.k_space`` (scatter.py:50-65).  Returns (kxyz, k_vecs).

    ``k_vecs[d] = (2*pi*kvec/l)**2`` (SQUARED components); ``kxyz = sum(meshgrid(*k_vecs))``
    with numpy's default ``indexing='xy'`` -- shape (N1, N0, N2): quirk Q1 of SURVEY.md 8(a).
    """
    k_vecs = []
    for N, l in zip(NxNyNz, lxlylz):
        N = int(N)
        assert N % 2 == 0 or N == 1, f"grid sizes must be even or 1, got {NxNyNz}"
        if N != 1:
            k_vecs.append((2 * np.pi * k_vector(N) / l) ** 2)
    K = np.meshgrid(*k_vecs)
    kxyz = sum(K)
    return kxyz, k_vecs


# --------------------------------------------------------------------------- literal port
def scattering_peak(mat, lxlylz, cutoff: int = 30, threshold: float = 1e-12, literal_txyz: bool = False):
    """Literal port of ``Scatter.sacttering_peak`` (scatter.py:68-122) on an array.

    ``mat`` is ``parsed.mat.squeeze()`` (fp64, C order), ``lxlylz`` the box.  Returns
    ``(q_Intensity (rows,2), q_idx {shell: [d-vectors]})``.

    ``literal_txyz=True`` builds ``txyz`` with ``list(itertools.product(...))`` exactly as
    scatter.py:82 does (52 % of the reference's wall time at 128^3); the default builds the
    same array with vectorised numpy -- identical values, a favour to the CPU baseline.
    """
    mat = np.asarray(mat, dtype=np.float64)
    NxNyNz = np.array(mat.shape)
    kxyz, k_vecs = k_space(NxNyNz, lxlylz)
    Nxyz = int(np.prod(NxNyNz))
    if literal_txyz:
        txyz = np.asarray(list(product(*k_vecs)))
    else:
        grids = np.meshgrid(*k_vecs, indexing="ij")
        txyz = np.stack([g.reshape(-1) for g in grids], axis=1)

    factor = 1 / Nxyz
    ipha = np.fft.fftn(mat)                                   # scatter.py:85
    phaNS = ipha.reshape((Nxyz, 1))
    kxyzNS = kxyz.reshape((Nxyz, 1))                          # Q1: flattened in (N1,N0,N2) order
    kpp = np.hstack([np.real(kxyzNS), factor * np.abs(phaNS ** 2), np.real(txyz)])
    kpp = kpp[np.argsort(kpp[:, 0]), :]                       # scatter.py:97 (unstable sort, Q3)

    q_Intensity = [kpp[0, 0:2]]
    q_idx = {0: [kpp[0, 2:]]}
    for i in range(1, kpp.shape[0]):                          # scatter.py:104-115
        n = len(q_Intensity)
        if n > cutoff:
            break                                             # Q2: last shell keeps one member
        if np.abs(q_Intensity[n - 1][0] - np.abs(kpp[i, 0])) > threshold:
            q_Intensity.append(kpp[i, 0:2])
            q_idx[n] = [kpp[i, 2:]]
        else:
            q_Intensity[n - 1][1] += kpp[i, 1]
            q_idx.setdefault(n - 1, []).append(kpp[i, 2:])
    q_Intensity = np.asarray(q_Intensity)
    q_Intensity[:, 0] = np.sqrt(q_Intensity[:, 0])
    return q_Intensity, q_idx


# --------------------------------------------------------------------------- shell-table port
def mode_k2(NxNyNz, lxlylz):
    """k^2 assigned by the reference to every FULL-spectrum element F[a,b,c], shape = mat.shape.

    Restates quirk Q1: with plane index m = a*N1 + b, F[a,b,c] is paired with
    ``(kv0[m % N0] + kv1[m // N0]) + kv2[c]``.
    """
    dims = [int(n) for n in NxNyNz if int(n) != 1]
    ls = [l for n, l in zip(NxNyNz, lxlylz) if int(n) != 1]
    kv = [(2 * np.pi * k_vector(n) / l) ** 2 for n, l in zip(dims, ls)]
    if len(dims) == 1:
        return kv[0].copy(), kv
    N0, N1 = dims[0], dims[1]
    m = np.arange(N0 * N1)
    plane = kv[0][m % N0] + kv[1][m // N0]
    if len(dims) == 2:
        return plane.reshape(N0, N1), kv
    return (plane[:, None] + kv[2][None, :]).reshape(N0, N1, dims[2]), kv


def shell_table(NxNyNz, lxlylz, cutoff: int = 30, threshold: float = 1e-12):
    """Field-independent shell geometry.

    Returns ``reps`` (k^2 of the first -- smallest -- member of each returned shell),
    ``shell_of`` (int32 array over the full spectrum: shell id, or -1 for modes beyond
    the table), ``truncated`` (True when the last row is the single-member row of Q2)
    and the dict of member index arrays.  The single member kept for a truncated last
    shell is the one with the LOWEST flat index (what a stable sort would keep).
    """
    k2, kv = mode_k2(NxNyNz, lxlylz)
    flat = k2.reshape(-1)
    order = np.argsort(flat, kind="stable")
    sv = flat[order]
    reps = [sv[0]]
    starts = [0]
    i = 1
    n = flat.size
    truncated = False
    while i < n:
        if len(reps) > cutoff:
            truncated = True
            break
        if np.abs(reps[-1] - np.abs(sv[i])) > threshold:
            reps.append(sv[i])
            starts.append(i)
        i += 1
    end = i
    shell_of = np.full(n, -1, dtype=np.int32)
    members = {}
    bounds = starts + [end]
    for s in range(len(reps)):
        idx = order[bounds[s]:bounds[s + 1]]
        if truncated and s == len(reps) - 1:
            idx = idx[:1]
        shell_of[idx] = s
        members[s] = np.sort(idx)
    return np.asarray(reps), shell_of.reshape(k2.shape), truncated, members, kv


def scattering_peak_table(mat, lxlylz, cutoff: int = 30, threshold: float = 1e-12):
    """Sort-free formulation: bin |F|^2/N into the shell table.  Same outputs as ``scattering_peak``
    (q bit-exact; I equal up to fp64 summation order; q_idx as a multiset)."""
    mat = np.asarray(mat, dtype=np.float64)
    reps, shell_of, truncated, members, kv = shell_table(mat.shape, lxlylz, cutoff, threshold)
    F = np.fft.fftn(mat)
    inten = (1 / mat.size) * np.abs(F ** 2)
    sel = shell_of.reshape(-1) >= 0
    I = np.bincount(shell_of.reshape(-1)[sel], weights=inten.reshape(-1)[sel], minlength=len(reps))
    qI = np.stack([np.sqrt(reps), I], axis=1)
    q_idx = {}
    for s, idx in members.items():
        sub = np.unravel_index(idx, mat.shape)
        q_idx[s] = [np.array([kv[d][sub[d][j]] for d in range(mat.ndim)]) for j in range(len(idx))]
    return qI, q_idx


# --------------------------------------------------------------------------- show_peak
def gaussian_expansion(array, amp, ctr, wid):
    """``peak.gaussian_expansion`` (calculator/peak.py:11-17)."""
    return amp * np.exp(-(((array - ctr) / wid) ** 2))


def show_peak(q_Intensity, width: float = 0.5, step: int = 2000, cutoff: float = 20, min_height: float = 1):
    """Numeric part of ``Scatter.show_peak`` (scatter.py:137-151).

    Returns dict(peaks_index, peaks_location, peak_heights, plot_x, plot_y, q).
    ``scipy.signal.find_peaks`` is the same third-party call the reference makes.
    """
    import scipy.signal as sg

    qI = np.array(q_Intensity, dtype=np.float64, copy=True)
    qI[0, 1] = 0
    qI = qI[qI[:, 0] <= cutoff]
    x = np.linspace(qI[0, 0], qI[-1, 0], step)
    plot_y = 0
    for q_i in qI:
        plot_y = plot_y + gaussian_expansion(array=x, ctr=q_i[0], amp=q_i[1], wid=width)
    peaks_location, props = sg.find_peaks(plot_y, height=min_height)
    q = x.copy()
    x = x / x[peaks_location[props["peak_heights"].argmax()]]
    return dict(
        peaks_index=peaks_location,
        peaks_location=x[peaks_location],
        peak_heights=props["peak_heights"],
        plot_x=x,
        plot_y=plot_y,
        q=q,
    )


def find_peaks_height(y, height):
    """Restatement of ``scipy.signal.find_peaks(y, height=h)`` (scipy 1.11-1.18,
    ``_peak_finding_utils._local_maxima_1d`` + ``_select_by_property``): strict rise into a
    plateau, strict fall out of it, plateau midpoint (floor), edges excluded; keep peaks with
    ``y[p] >= height``."""
    y = np.asarray(y, dtype=np.float64)
    n = y.size
    out = []
    i = 1
    i_max = n - 1
    while i < i_max:
        if y[i - 1] < y[i]:
            i_ahead = i + 1
            while i_ahead < i_max and y[i_ahead] == y[i]:
                i_ahead += 1
            if y[i_ahead] < y[i]:
                left, right = i, i_ahead - 1
                out.append((left + right) // 2)
                i = i_ahead
        i += 1
    out = np.asarray(out, dtype=np.int64)
    return out[y[out] >= height] if out.size else out
