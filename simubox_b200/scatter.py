"""``Scatter`` -- scattering I(q) and Bragg peaks, same entry points as the reference
(SimuBox/calculator/scatter.py:38-178), computed by libsimubox_b200.so on a B200.

``sacttering_peak`` [sic, reference spelling] = FFT + |F|^2/N + |q|-shell sums;
``show_peak`` = Gaussian broadening + peak search.  ``sacttering_peak_batch`` is an addition
for sweeps (many fields of one grid) and is what ``bench.py`` drives.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Union

import numpy as np

from . import _lib
from ._device import ptr, stream_ptr, to_device, torch_cuda
from .reader import parse_density
from .schema import Density, PathLike, ScatterPlot, ScatterResult

_PLANS: dict = {}
_MAX_PEAKS = 256


def get_plan(N, L, cutoff=30, threshold=1e-12, max_batch=1, device=0, path=_lib.PATH_AUTO) -> _lib.ScatterPlan:
    """Cached scatter plan (cuFFT plan + shell table) for one grid / box / cutoff."""
    key = (tuple(int(n) for n in N), tuple(float(l) for l in L), int(cutoff), float(threshold), int(max_batch),
           int(device), int(path))
    plan = _PLANS.get(key)
    if plan is None:
        if len(_PLANS) >= 16:
            _PLANS.pop(next(iter(_PLANS))).close()
        plan = _PLANS[key] = _lib.ScatterPlan(key[0], key[1], cutoff, threshold, max_batch, device, path)
    return plan


def clear_plans():
    while _PLANS:
        _PLANS.popitem()[1].close()


class Scatter:
    @staticmethod
    def k_vector(N: int):
        """Signed integer frequencies [0..N/2, -N/2+1..-1] as float (reference scatter.py:40-47)."""
        n = int(N)
        pos = np.arange(0, n // 2 + 1, dtype=np.float64)
        neg = np.arange(-(n // 2) + 1, 0, dtype=np.float64)
        return np.concatenate([pos, neg]) if n > 1 else pos[:1] if n == 1 else pos

    @classmethod
    def k_space(cls, *args):
        """(kxyz, k_vecs) as the reference returns them (scatter.py:50-65), including the default
        ``meshgrid`` 'xy' layout of ``kxyz`` (shape (N1, N0, N2))."""
        half = len(args) // 2
        k_vecs = []
        for N, l in zip(args[:half], args[half:]):
            assert N % 2 == 0 or N == 1, f"grid sizes must be even or 1, got {args[:half]}"
            if N != 1:
                k_vecs.append((2 * np.pi * cls.k_vector(N) / l) ** 2)
        kxyz = sum(np.meshgrid(*k_vecs))
        return kxyz, k_vecs

    # ------------------------------------------------------------------ I(q)
    @classmethod
    def sacttering_peak(cls, density: Density, cutoff: int = 30, threshold: float = 1e-12, **kwargs) -> ScatterResult:
        """Drop-in for ``Scatter.sacttering_peak`` (scatter.py:68-122).  kwargs go to
        ``parse_density`` (target / permute / slices / expand); ``device=`` selects the GPU."""
        device = int(kwargs.pop("device", 0))
        parsed = parse_density(density, **kwargs)
        mat = parsed.mat.squeeze()
        dims = [int(n) for n in parsed.NxNyNz]
        assert all(n % 2 == 0 or n == 1 for n in dims), f"grid sizes must be even or 1, got {dims}"
        keep = [i for i, n in enumerate(dims) if n != 1]
        N = [dims[i] for i in keep]
        L = [float(parsed.lxlylz[i]) for i in keep]
        if mat.shape != tuple(N):
            raise ValueError(f"sacttering_peak needs a single target; got an array of shape {mat.shape} for grid {N}")
        plan = get_plan(N, L, cutoff, threshold, max_batch=1, device=device)
        torch = torch_cuda()
        field = to_device(np.ascontiguousarray(mat, dtype=np.float64), device)
        out = torch.empty((1, plan.n_shells), dtype=torch.float64, device=field.device)
        _lib.check(_lib.lib().sbx_scatter_execute(plan.handle, ptr(field), 1, ptr(out), stream_ptr(device)))
        q_Intensity = np.stack([plan.q, out[0].cpu().numpy()], axis=1)
        return ScatterResult(parsed_density=parsed, q_Intensity=q_Intensity, q_idx=plan.q_idx())

    @classmethod
    def sacttering_peak_batch(cls, fields, lxlylz, cutoff: int = 30, threshold: float = 1e-12, peaks: bool = False,
                              width: float = 0.5, step: int = 2000, peak_cutoff: float = 20, min_height: float = 1,
                              max_batch: int = 16, device: int = 0, path: int = _lib.PATH_AUTO):
        """Many fields of one grid.  ``fields``: (B, *N) fp64 -- a numpy array (host; pinned or
        not) is streamed to the GPU in chunks with copy/compute overlap, a CUDA tensor is used
        in place (the plan is then taken on the tensor's device, whatever ``device`` says).  Plans are cached per
        (grid, box, cutoff, batch, device, path) and own one workspace: use one stream / thread per plan.
        Returns dict(q, I (B, n_shells), [peak_idx (B, max_peaks), n_peaks (B)])."""
        torch = torch_cuda()
        is_tensor = isinstance(fields, torch.Tensor)
        shape = tuple(fields.shape)
        B, N = int(shape[0]), shape[1:]
        if is_tensor and fields.is_cuda:
            device = fields.device.index  # the plan (tables, workspace, stream) must live where the data is
        plan = get_plan(N, lxlylz, cutoff, threshold, max_batch=min(max_batch, B), device=device, path=path)
        L_ = _lib.lib()
        res = {"q": plan.q.copy(), "path": _lib.PATH_NAMES.get(plan.path, "?")}
        if is_tensor and fields.is_cuda:
            f = fields.contiguous()
            assert f.dtype == torch.float64
            I = torch.empty((B, plan.n_shells), dtype=torch.float64, device=f.device)
            pidx = torch.empty((B, _MAX_PEAKS), dtype=torch.int32, device=f.device) if peaks else None
            npk = torch.empty((B,), dtype=torch.int32, device=f.device) if peaks else None
            st = stream_ptr(f.device.index)
            if peaks:
                _lib.check(L_.sbx_scatter_peaks_execute(plan.handle, ptr(f), B, ptr(I), width, step, peak_cutoff,
                                                        min_height, None, ptr(pidx), ptr(npk), _MAX_PEAKS, st))
            else:
                _lib.check(L_.sbx_scatter_execute(plan.handle, ptr(f), B, ptr(I), st))
            res["I"] = I
            if peaks:
                res["peak_idx"], res["n_peaks"] = pidx, npk
            return res
        if is_tensor:
            host = fields.contiguous()
            assert host.dtype == torch.float64
            hp = C.c_void_p(host.data_ptr())
        else:
            host = np.ascontiguousarray(fields, dtype=np.float64)
            hp = C.c_void_p(host.ctypes.data)
        I = np.empty((B, plan.n_shells), dtype=np.float64)
        pidx = np.full((B, _MAX_PEAKS), -1, dtype=np.int32) if peaks else None
        npk = np.zeros((B,), dtype=np.int32) if peaks else None
        _lib.check(L_.sbx_scatter_execute_host(
            plan.handle, hp, B, C.c_void_p(I.ctypes.data), width, step, peak_cutoff, min_height,
            C.c_void_p(pidx.ctypes.data) if peaks else None, C.c_void_p(npk.ctypes.data) if peaks else None,
            _MAX_PEAKS if peaks else 0))
        res["I"] = I
        if peaks:
            res["peak_idx"], res["n_peaks"] = pidx, npk
        return res

    # ------------------------------------------------------------------ peaks
    @classmethod
    def show_peak(cls, res: ScatterResult, width: float = 0.5, step: int = 2000, cutoff: int = 20, min_height: int = 1,
                  save: Union[PathLike, bool] = False, interactive: bool = False, **kwargs) -> ScatterPlot:
        """Drop-in for ``Scatter.show_peak`` (scatter.py:125-178): broadened curve + peak search on
        the GPU.  The matplotlib figure is produced only when matplotlib is importable and
        ``plot=True`` is passed (no graphics stack on the GPU box); ``fig``/``ax`` are None otherwise."""
        device = int(kwargs.pop("device", 0))
        torch = torch_cuda()
        qI = np.ascontiguousarray(res.q_Intensity, dtype=np.float64)
        n = qI.shape[0]
        q = to_device(qI[:, 0].copy(), device)
        I = to_device(qI[:, 1].copy(), device)
        plot_y = torch.empty((step,), dtype=torch.float64, device=q.device)
        pidx = torch.empty((_MAX_PEAKS,), dtype=torch.int32, device=q.device)
        npk = torch.zeros((1,), dtype=torch.int32, device=q.device)
        _lib.check(_lib.lib().sbx_peaks_execute(ptr(q), ptr(I), n, 1, float(width), int(step), float(cutoff),
                                                float(min_height), ptr(plot_y), ptr(pidx), ptr(npk), _MAX_PEAKS,
                                                stream_ptr(device)))
        count = int(npk.item())
        if count < 0:
            raise IndexError("no shell with q <= cutoff")  # the reference fails on q_Intensity[0, 0] here
        if count > _MAX_PEAKS:
            raise RuntimeError(f"{count} peaks exceed the {_MAX_PEAKS}-entry peak buffer")
        y = plot_y.cpu().numpy()
        peaks_location = pidx[:count].cpu().numpy().astype(np.int64)
        kept = qI[qI[:, 0] <= cutoff, 0]
        x = np.linspace(kept[0], kept[-1], step)
        heights = {"peak_heights": y[peaks_location]}
        qx = x.copy()
        x = x / x[peaks_location[heights["peak_heights"].argmax()]]  # ValueError when no peak, as the reference
        plot_x = x if kwargs.get("scale", True) else qx
        fig = ax = None
        if kwargs.get("plot", False) or save:
            fig, ax = _plot_curve(plot_x, y, kwargs)
        return ScatterPlot(peaks_location=x[peaks_location], peaks_height=heights, plot_x=plot_x, plot_y=y, fig=fig, ax=ax)


def _plot_curve(x, y, kwargs):
    try:
        import matplotlib.pyplot as plt
    except Exception:  # no graphics stack here: figures are optional
        return None, None
    fig = plt.figure(figsize=kwargs.get("figsize", (8, 6)))
    ax = plt.gca()
    plt.plot(x, y, lw=2, c="k")
    plt.ylabel("Intensity", fontsize=20)
    plt.xlabel(r"$q/R_g^{-1}$", fontsize=20)
    plt.tight_layout()
    return fig, ax
