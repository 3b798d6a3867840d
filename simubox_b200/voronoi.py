"""``VoronoiCell`` -- grid Voronoi labelling of domain centres on a B200, with the reference's
entry points (SimuBox/calculator/voronoi.py:29-327).

``additive`` / ``power`` keep the reference signatures (host helpers for a single pixel).
``label_grid`` is the addition that returns what the reference only paints: the int32 label
array, computed by ``sbx_voronoi_label`` (2D or 3D, optional periodic box and uint8 mask).
``weighted_voronoi_diagrams`` keeps its signature and routes the per-pixel loop
(voronoi.py:284-292) through ``label_grid``.  ``OpenCV`` (the 2D seed finder) and ``Analyze`` with its ``voronoi`` /
``triangle`` modes (voronoi.py:70-241: centres -> box coordinates, Qhull Voronoi, angle-filtered Delaunay,
``coord_dict``) stay on the CPU with cv2 / scipy as in the reference -- 67 centres, milliseconds; their figures are
drawn only when matplotlib is importable (no graphics stack on this box), the numeric results always.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple, Union

import numpy as np

from . import _lib
from ._device import ptr, stream_ptr, to_device, torch_cuda
from .reader import parse_density
from .schema import AnalyzeMode, ColorMode, Density, OpenCVResult, VoronoiAnalyzeResult, WeightedMode

_MODES = {"additive": 0, "power": 1}
_PLANS: dict = {}


class _VoronoiPlan:
    """Owner of an ``sbx_voronoi_plan`` (candidate-list workspace of one grid size on one device)."""

    def __init__(self, device: int, shape):
        self._h = C.c_void_p()
        self._keep, pn = _lib.i64(shape)
        _lib.check(_lib.lib().sbx_voronoi_plan_create(C.byref(self._h), int(device), len(shape), pn))

    def close(self):
        if self._h:
            _lib.lib().sbx_voronoi_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _get_plan(device: int, shape, slot: int = 0) -> _VoronoiPlan:
    """Cached plan (workspace) per device, grid and ``slot``.  A plan serves one stream at a time; labellings issued
    on several streams at once must use different slots."""
    key = (int(device), tuple(shape), int(slot))
    plan = _PLANS.get(key)
    if plan is None:
        if len(_PLANS) >= 16:
            _PLANS.pop(next(iter(_PLANS))).close()
        plan = _PLANS[key] = _VoronoiPlan(device, shape)
    return plan


def _mode_id(method) -> int:
    name = method.value if isinstance(method, WeightedMode) else str(method)
    if name not in _MODES:
        raise AttributeError(f"VoronoiCell has no weighting method {name!r}")  # reference: getattr(cls, name)
    return _MODES[name]


class VoronoiCell:
    @staticmethod
    def additive(points, arr, weights):
        """||p - x||_2 - w per seed (reference voronoi.py:244-246)."""
        return np.linalg.norm(points - arr, axis=1) - weights

    @staticmethod
    def power(points, arr, weights):
        """(||p - x + s||_2)^2 - w per seed, seed index s added to every coordinate
        (reference voronoi.py:249-255; kept as defined there)."""
        return np.linalg.norm(points - arr + np.arange(len(points))[:, np.newaxis], axis=1) ** 2 - weights

    # ------------------------------------------------------------------ GPU labelling
    @classmethod
    def label_grid(cls, points, weights=0, shape: Sequence[int] = (500, 500),
                   method: Union[WeightedMode, str] = WeightedMode.ADDITIVE, box: Optional[Sequence[float]] = None,
                   mask=None, device: int = 0, return_tensor: bool = False, stats: bool = False,
                   planes: Optional[Tuple[int, int]] = None, plan_slot: int = 0):
        """Label every voxel of ``product(range(shape[0]), range(shape[1])[, range(shape[2])])``
        with the index of the seed minimising ``method`` (first minimum, like ``np.argmin``).

        ``box``: periodic box lengths in voxel units (minimum image), None = non-periodic as the
        reference.  ``mask``: uint8/bool array of ``shape``; voxels where it is 0 get label -1.
        ``planes=(x_lo, x_hi)``: label only those planes of the first axis (one rank's slab of a grid
        sharded across GPUs, see ``distributed.label_grid_sharded``); the result then has shape
        ``(x_hi - x_lo, *shape[1:])`` and ``mask`` may be the whole-grid mask or the slab's.
        ``plan_slot``: the call is asynchronous on the current CUDA stream and uses the workspace of plan number
        ``plan_slot`` of this grid; labellings in flight on different streams need different slots.
        Returns an int32 array of ``shape`` (numpy, or a CUDA tensor with ``return_tensor``);
        with ``stats=True`` also a dict(evaluations, tiles, overflow_tiles, path).
        """
        torch = torch_cuda()
        shape = tuple(int(n) for n in shape)
        nd = len(shape)
        if not isinstance(points, torch.Tensor) and not np.isfinite(np.asarray(points, dtype=np.float64)).all():
            raise ValueError("seed coordinates must be finite")
        x_lo, x_hi = (0, shape[0]) if planes is None else (int(planes[0]), int(planes[1]))
        if not (0 <= x_lo < x_hi <= shape[0]):
            raise ValueError(f"planes {planes} outside [0, {shape[0]})")
        local = (x_hi - x_lo,) + shape[1:]
        pts = to_device(np.ascontiguousarray(points, dtype=np.float64) if not isinstance(points, torch.Tensor) else points,
                        device, torch.float64)
        if pts.ndim != 2 or pts.shape[1] != nd:
            raise ValueError(f"points must be (S, {nd}), got {tuple(pts.shape)}")
        S = int(pts.shape[0])
        if isinstance(weights, (int, float)):
            w = None if weights == 0 else to_device(np.full(S, float(weights)), device)
        elif isinstance(weights, (list, np.ndarray)) or isinstance(weights, torch.Tensor):
            w = to_device(weights if isinstance(weights, torch.Tensor) else np.asarray(weights, dtype=np.float64), device, torch.float64)
            if w.numel() != S:
                raise ValueError(f"weights must have one entry per seed ({S}), got {w.numel()}")
        else:
            raise ValueError("Weights must be an array")  # reference voronoi.py:280
        m = None
        if mask is not None:
            m = to_device(mask if isinstance(mask, torch.Tensor) else np.ascontiguousarray(mask), device)
            m = (m != 0).to(torch.uint8) if m.dtype != torch.uint8 else m
            if tuple(m.shape) == shape:
                m = m[x_lo:x_hi]
            elif tuple(m.shape) != local:
                raise ValueError(f"mask shape {tuple(m.shape)} != grid {shape}")
            m = m.contiguous()
        labels = torch.empty(local, dtype=torch.int32, device=pts.device)
        pb = None
        if box is not None:
            _keep, pb = _lib.f64(box)
            if len(_keep) != nd:
                raise ValueError(f"box must have {nd} entries")
        st_arr = np.zeros(3, dtype=np.int64) if stats else None
        plan = _get_plan(device, shape, plan_slot)
        _lib.check(_lib.lib().sbx_voronoi_plan_execute_range(
            plan._h, ptr(pts), S, ptr(w), _mode_id(method), pb, ptr(m), ptr(labels), x_lo, x_hi, stream_ptr(device),
            st_arr.ctypes.data_as(C.POINTER(C.c_int64)) if stats else None))
        out = labels if return_tensor else labels.cpu().numpy()
        if stats:
            path = {0: "tiles", 1: "cells"}.get(int(_lib.lib().sbx_voronoi_plan_last_path(plan._h)), "?")
            return out, dict(evaluations=int(st_arr[0]), tiles=int(st_arr[1]), overflow_tiles=int(st_arr[2]), path=path)
        return out

    @classmethod
    def domain_centres(cls, field, threshold: Optional[float] = None, periodic: bool = True, min_voxels: int = 1,
                       device: int = 0, return_labels: bool = False) -> dict:
        """Domain centres of a 2D or 3D field: connected components of ``field > threshold`` (faces,
        wrap-around when ``periodic``) and their centroids in voxel units -- the 3D-capable counterpart
        of ``VoronoiCell.OpenCV`` (reference voronoi.py:31-68, 2D only, 3 x 3 tiling + cv2 moments).

        ``threshold`` defaults to the mid-range ``(min + max) / 2`` of the field.  Components with fewer
        than ``min_voxels`` voxels are dropped from ``centres`` / ``volumes`` (their ids stay in
        ``labels``).  Returns dict(centres (S, d), volumes (S,), ids (S,), n_components[, labels]).
        Feed ``centres`` to ``label_grid(points=centres, box=shape)``.
        """
        torch = torch_cuda()
        f = to_device(field if isinstance(field, torch.Tensor) else np.ascontiguousarray(field, dtype=np.float64), device, torch.float64)
        f = f.contiguous()
        shape = tuple(int(n) for n in f.shape)
        nd = len(shape)
        if nd not in (2, 3):
            raise ValueError(f"field must be 2D or 3D, got shape {shape}")
        if threshold is None:
            threshold = 0.5 * (float(f.min().item()) + float(f.max().item()))
        comp = torch.empty(shape, dtype=torch.int32, device=f.device) if return_labels else None
        _, pn = _lib.i64(shape)
        cap = 4096
        while True:
            centres = torch.empty((cap, nd), dtype=torch.float64, device=f.device)
            volumes = torch.empty(cap, dtype=torch.int64, device=f.device)
            n = C.c_int64()
            _lib.check(_lib.lib().sbx_domain_centres(device, nd, pn, ptr(f), float(threshold), int(bool(periodic)), ptr(comp),
                                                     cap, ptr(centres), ptr(volumes), C.byref(n), stream_ptr(device)))
            if n.value <= cap:
                break
            cap = int(n.value)
        S = int(n.value)
        c, v = centres[:S].cpu().numpy(), volumes[:S].cpu().numpy()
        keep = v >= int(min_voxels)
        out = dict(centres=c[keep], volumes=v[keep], ids=np.flatnonzero(keep).astype(np.int32), n_components=S, threshold=float(threshold))
        if return_labels:
            out["labels"] = comp.cpu().numpy()
        return out

    @classmethod
    def cell_statistics(cls, labels, field=None, n_cells: Optional[int] = None, periodic: bool = True,
                        neighbours: bool = True, device: int = 0) -> dict:
        """Per-cell statistics of a label grid (``label_grid`` output; label < 0 = masked voxel).

        Returns ``volume`` (int64 voxels per cell), ``mass`` (sum of ``field`` per cell, when a field is
        given) and, with ``neighbours``, the face-adjacency graph of the cells: ``degree`` (coordination
        number per cell), ``indptr`` / ``indices`` (CSR neighbour lists, ascending) and ``coord_dict``
        (cell -> set of neighbours, the structure ``VoronoiCell.triangle`` returns from Delaunay
        triangles, reference voronoi.py:178-182).  ``periodic``: neighbours wrap around the grid.
        """
        torch = torch_cuda()
        lab = to_device(labels if isinstance(labels, torch.Tensor) else np.ascontiguousarray(labels, dtype=np.int32), device, torch.int32)
        lab = lab.contiguous()
        shape = tuple(int(n) for n in lab.shape)
        if len(shape) not in (2, 3):
            raise ValueError(f"labels must be a 2D or 3D grid, got shape {shape}")
        f = None
        if field is not None:
            f = to_device(field if isinstance(field, torch.Tensor) else np.ascontiguousarray(field, dtype=np.float64), device, torch.float64)
            f = f.contiguous()
            if tuple(f.shape) != shape:
                raise ValueError(f"field shape {tuple(f.shape)} != labels {shape}")
        if n_cells is None:
            n_cells = int(lab.max().item()) + 1
        n_cells = int(n_cells)
        if n_cells < 1:
            raise ValueError("no labelled voxel")
        words = (n_cells + 31) // 32
        vol = torch.empty(n_cells, dtype=torch.int64, device=lab.device)
        mass = torch.empty(n_cells, dtype=torch.float64, device=lab.device) if f is not None else None
        adj = torch.empty((n_cells, words), dtype=torch.int32, device=lab.device) if neighbours else None
        deg = torch.empty(n_cells, dtype=torch.int32, device=lab.device) if neighbours else None
        _, pn = _lib.i64(shape)
        _lib.check(_lib.lib().sbx_cell_stats(device, len(shape), pn, ptr(lab), ptr(f), n_cells, int(bool(periodic)),
                                             ptr(vol), ptr(mass), ptr(adj), ptr(deg), stream_ptr(device)))
        out = dict(volume=vol.cpu().numpy(), n_cells=n_cells)
        if mass is not None:
            out["mass"] = mass.cpu().numpy()
        if neighbours:
            bits = np.unpackbits(adj.cpu().numpy().view(np.uint8), axis=1, bitorder="little")[:, :n_cells]
            rows, cols = np.nonzero(bits)
            degree = deg.cpu().numpy()
            indptr = np.zeros(n_cells + 1, dtype=np.int64)
            np.cumsum(degree, out=indptr[1:])
            out.update(degree=degree, indptr=indptr, indices=cols.astype(np.int32),
                       coord_dict={int(s): set(cols[indptr[s]:indptr[s + 1]].tolist()) for s in range(n_cells) if degree[s]})
        return out

    @classmethod
    def weighted_voronoi_diagrams(cls, points, weights=0, method: WeightedMode = WeightedMode.ADDITIVE,
                                  size: Tuple = (500, 500), color_mode: ColorMode = ColorMode.L,
                                  interactive: bool = True, num_threads: int = 8, **kwargs):
        """Reference signature (voronoi.py:258-327).  The per-pixel loop runs on the GPU; the
        figure is drawn only if matplotlib is importable (``fig, ax`` are None otherwise).  The
        label image ``labels[x, y]`` is attached as ``ax.sbx_labels`` when a figure exists and is
        always available from ``label_grid``."""
        if isinstance(points, list):
            points = np.array(points)
        if isinstance(weights, (list, np.ndarray)):
            weights = np.array(weights)
        elif not isinstance(weights, (int, float)):
            raise ValueError("Weights must be an array")
        labels = cls.label_grid(points, weights, shape=size, method=method, device=int(kwargs.pop("device", 0)))
        try:
            import matplotlib.pyplot as plt
        except Exception:
            return None, None
        fig = plt.figure(figsize=kwargs.get("figsize", (5, 5)))
        ax = plt.gca()
        n = max(len(points) - 1, 1)
        plt.imshow((labels.T * (255 // n)).astype(np.uint8), "gray", zorder=5, origin="lower")
        plt.scatter(points[:, 0], points[:, 1], s=max(size) * 0.1, c="white", zorder=10, edgecolors="black")
        ax.sbx_labels = labels
        if interactive:
            plt.show()
        return fig, ax

    # ------------------------------------------------------------------ Analyze (reference voronoi.py:70-241)
    @staticmethod
    def in_lim(point, xlim, ylim):
        """Points strictly inside the window (reference voronoi.py:70-77)."""
        return (xlim[0] < point[:, 0]) & (point[:, 0] < xlim[1]) & (ylim[0] < point[:, 1]) & (point[:, 1] < ylim[1])

    @staticmethod
    def get_lim_attributes(key: str, default: Sequence, expand, **kwargs):
        """Window in units of the raw box, shifted by half the expansion (reference voronoi.py:103-106)."""
        return np.array(kwargs.get(key, default)) + expand / 2

    @classmethod
    def set_point_limits(cls, density: OpenCVResult, centers: np.ndarray, **kwargs):
        """Centres inside the plotted window (reference voronoi.py:215-230)."""
        pd_ = density.parsed_density
        xlim = cls.get_lim_attributes("point_xlim", (-0.1, 1.1), pd_.expand[0], **kwargs)
        ylim = cls.get_lim_attributes("point_ylim", (-0.1, 1.1), pd_.expand[1], **kwargs)
        return centers[cls.in_lim(point=centers, xlim=pd_.raw_lxlylz[1] * xlim, ylim=pd_.raw_lxlylz[0] * ylim)]

    @classmethod
    def set_axis_limits(cls, density: OpenCVResult, **kwargs):
        """Axis window of the figure (reference voronoi.py:232-241); returns (xlim, ylim) and applies them when a
        matplotlib figure is current."""
        pd_ = density.parsed_density
        xlim = pd_.raw_lxlylz[1] * cls.get_lim_attributes("axis_xlim", (-0.5, 1.5), pd_.expand[0], **kwargs)
        ylim = pd_.raw_lxlylz[0] * cls.get_lim_attributes("axis_ylim", (-0.5, 1.5), pd_.expand[1], **kwargs)
        try:
            import matplotlib.pyplot as plt

            plt.xlim(xlim)
            plt.ylim(ylim)
        except Exception:
            pass
        return xlim, ylim

    @classmethod
    def Analyze(cls, density: Density, mode: AnalyzeMode, **kwargs) -> VoronoiAnalyzeResult:
        """Domain centres of the tiled field -> box coordinates -> Voronoi diagram or filtered Delaunay
        triangulation (reference voronoi.py:79-101).  ``kwargs`` go to ``OpenCV`` / ``parse_density`` and to the
        mode function exactly as in the reference (``expand`` is consumed by ``OpenCV``)."""
        cv_res = cls.OpenCV(density, **kwargs)
        kwargs.pop("expand", None)
        pd_ = cv_res.parsed_density
        vecs = [np.linspace(0, l, N) for l, N in zip(pd_.lxlylz, pd_.NxNyNz)][::-1]
        X, Y = np.meshgrid(*vecs)
        centers = np.around(cv_res.centers).astype(int)
        centers_lxly = np.stack([X[centers[:, 1], centers[:, 0]], Y[centers[:, 1], centers[:, 0]]], axis=1)
        if mode == AnalyzeMode.VORONOI:
            return cls.voronoi(cv_res, centers_lxly, **kwargs)
        if mode == AnalyzeMode.TRIANGLE:
            return cls.triangle(cv_res, centers_lxly, **kwargs)
        raise NotImplementedError(f"analysis mode {mode} is not implemented")

    @classmethod
    def voronoi(cls, cv_res: OpenCVResult, centers: np.ndarray, interactive: bool = True, save=False, **kwargs) -> VoronoiAnalyzeResult:
        """``scipy.spatial.Voronoi(centers, incremental=True)`` (Qhull; reference voronoi.py:108-148).  The figure
        (``voronoi_plot_2d`` + the centres inside the point window) is drawn when matplotlib is importable."""
        from scipy.spatial import Voronoi

        centers_cut = cls.set_point_limits(cv_res, centers, **kwargs)
        vor = Voronoi(centers, incremental=True)
        fig = ax = None
        try:
            import matplotlib.pyplot as plt
            from scipy.spatial import voronoi_plot_2d

            asp = cv_res.parsed_density.lxlylz[1] / cv_res.parsed_density.lxlylz[0]
            plt.figure()
            ax = plt.gca()
            fig = voronoi_plot_2d(vor, ax, show_points=False, show_vertices=False, line_colors=kwargs.get("lc", "k"))
            if kwargs.get("point", True):
                plt.scatter(centers_cut[:, 0], centers_cut[:, 1], s=20, c=kwargs.get("pc", "b"))
            cls.set_axis_limits(cv_res, **kwargs)
            ax.set_aspect(1 / asp)
            if interactive:
                plt.show()
        except Exception:
            fig = ax = None
        return VoronoiAnalyzeResult(path=cv_res.path, cv_result=cv_res, fig=fig, ax=ax, voronoi=vor)

    @classmethod
    def triangle(cls, cv_res: OpenCVResult, centers: np.ndarray, interactive: bool = True, save=False, **kwargs) -> VoronoiAnalyzeResult:
        """Delaunay triangles of the centres without those whose largest angle exceeds ``theta_max`` (default
        pi / 2), and the neighbour sets they induce (reference voronoi.py:150-213).  Arithmetic as there: edge
        lengths from ``cdist``, ``a <= b <= c`` sorted, ``acos((a^2 + b^2 - c^2) / (2ab))``."""
        from collections import defaultdict
        from math import acos

        from scipy.spatial import Delaunay, distance

        centers_cut = cls.set_point_limits(cv_res, centers, **kwargs)
        theta_max = kwargs.get("theta_max", np.pi / 2)
        dist_matrix = distance.cdist(centers, centers, "euclidean")
        triangles = []
        for t in Delaunay(centers).simplices:
            a, b, c = sorted([dist_matrix[t[0], t[1]], dist_matrix[t[0], t[2]], dist_matrix[t[1], t[2]]])
            if acos((a ** 2 + b ** 2 - c ** 2) / (2 * a * b)) > theta_max:
                continue
            triangles.append(t)
        coord_dict = defaultdict(set)
        for t in triangles:
            coord_dict[t[0]].update([t[1], t[2]])
            coord_dict[t[1]].update([t[0], t[2]])
            coord_dict[t[2]].update([t[0], t[1]])
        fig = ax = None
        try:
            import matplotlib.pyplot as plt

            asp = cv_res.parsed_density.lxlylz[1] / cv_res.parsed_density.lxlylz[0]
            fig = plt.figure(figsize=kwargs.get("figsize", (6, 6 / asp)))
            ax = plt.gca()
            plt.triplot(centers[:, 0], centers[:, 1], triangles, linewidth=1, color=kwargs.get("lc", "k"))
            if kwargs.get("point", True):
                plt.scatter(centers_cut[:, 0], centers_cut[:, 1], s=20, c=kwargs.get("pc", "b"))
            cls.set_axis_limits(cv_res, **kwargs)
            ax.set_aspect(asp)
            if interactive:
                plt.show()
        except Exception:
            fig = ax = None
        return VoronoiAnalyzeResult(path=cv_res.path, cv_result=cv_res, fig=fig, ax=ax, triangle=triangles, coord_dict=coord_dict)

    # ------------------------------------------------------------------ CPU seed finder (as the reference)
    @staticmethod
    def OpenCV(density: Density, expand: Union[int, Sequence[int]] = 3, **kwargs) -> OpenCVResult:
        """Domain centres of a periodically tiled 2D field via cv2 contours + moments
        (reference voronoi.py:31-68; third-party arithmetic, unchanged, CPU)."""
        import cv2

        parsed = parse_density(density, expand=expand, **kwargs)
        mat = parsed.mat.squeeze()
        assert mat.ndim == 2, f"only 2D Voronoi analysis is supported, got a {mat.ndim}-d field"
        gray = np.array(mat * 255, dtype=np.uint8)
        h, w = gray.shape
        subdiv = cv2.Subdiv2D((0, 0, w, h))
        _, binary = cv2.threshold(gray, 127, 255, 0)
        contours, _ = cv2.findContours(binary, cv2.RETR_TREE, cv2.CHAIN_APPROX_SIMPLE)
        for contour in contours:
            mom = cv2.moments(contour)
            subdiv.insert((mom["m10"] / (mom["m00"] + 1e-6), mom["m01"] / (mom["m00"] + 1e-6)))
        facets, centers = subdiv.getVoronoiFacetList([])
        return OpenCVResult(path=density.path, parsed_density=parsed, facets=facets, centers=np.array(centers, dtype=float))
