"""ctypes binding of ``libsimubox_b200.so`` (C ABI declared in ``include/sbx.h``).

The library is built in-tree by ``simubox_b200/build.py`` (nvcc, sm_100a).  There is no CPU
fallback: if the library is missing or no sm_100 device is present the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("SIMUBOX_B200_LIB", PKG / "libsimubox_b200.so"))

SBX_DEVICE_NONE = -1
PATH_AUTO, PATH_DENSE, PATH_GATHER, PATH_PRUNED, PATH_PRUNED_FUSED = 0, 1, 2, 3, 4
PATH_NAMES = {0: "auto", 1: "fft_dense", 2: "fft_gather", 3: "pruned_dft", 4: "pruned_dft_fused"}

_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)
_i32p = C.POINTER(C.c_int32)
_vp = C.c_void_p

_SIGNATURES = {
    "sbx_version": (C.c_int, []),
    "sbx_scatter_slab_path": (C.c_int, [_vp]),
    "sbx_scatter_slab_kernel": (C.c_char_p, [_vp]),
    "sbx_particle_splat": (C.c_int, [C.c_int, _vp, _vp, C.c_int64, C.c_int, _i64p, _f64p, C.c_double, _vp, _vp]),
    "sbx_domain_centres": (C.c_int, [C.c_int, C.c_int, _i64p, _vp, C.c_double, C.c_int, _vp, C.c_int64, _vp, _vp, _i64p, _vp]),
    "sbx_cell_stats": (C.c_int, [C.c_int, C.c_int, _i64p, _vp, _vp, C.c_int64, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "sbx_text_is_ascii": (C.c_int, [_vp, C.c_int64]),
    "sbx_text_table_shape": (C.c_int, [_vp, C.c_int64, C.c_int, C.c_int, _i64p, C.POINTER(C.c_int32)]),
    "sbx_text_table_parse": (C.c_int, [_vp, C.c_int64, C.c_int, C.c_int64, C.c_int32, _f64p, C.c_int]),
    "sbx_last_error": (C.c_char_p, []),
    "sbx_device_count": (C.c_int, []),
    "sbx_scatter_plan_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, _i64p, _f64p, C.c_int, C.c_double, C.c_int]),
    "sbx_scatter_plan_create_ex": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, _i64p, _f64p, C.c_int, C.c_double, C.c_int, C.c_int]),
    "sbx_scatter_plan_destroy": (C.c_int, [_vp]),
    "sbx_scatter_plan_shells": (C.c_int, [_vp, C.POINTER(C.c_int), _f64p, _i64p]),
    "sbx_scatter_plan_members": (C.c_int, [_vp, C.c_int, _i64p, _f64p, C.c_int64]),
    "sbx_scatter_plan_path": (C.c_int, [_vp]),
    "sbx_scatter_plan_kernel": (C.c_char_p, [_vp]),
    "sbx_scatter_plan_set_profiling": (C.c_int, [_vp, C.c_int]),
    "sbx_scatter_plan_get_profile": (C.c_int, [_vp, _f64p, _f64p, _i64p, _i64p, C.c_int]),
    "sbx_scatter_execute": (C.c_int, [_vp, _vp, C.c_int, _vp, _vp]),
    "sbx_peaks_execute": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_double, C.c_double, _vp, _vp, _vp, C.c_int, _vp]),
    "sbx_scatter_peaks_execute": (C.c_int, [_vp, _vp, C.c_int, _vp, C.c_double, C.c_int, C.c_double, C.c_double, _vp, _vp, _vp, C.c_int, _vp]),
    "sbx_scatter_execute_host": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_double, C.c_int, C.c_double, C.c_double, _vp, _vp, C.c_int]),
    "sbx_voronoi_label": (C.c_int, [C.c_int, C.c_int, _i64p, _vp, C.c_int64, _vp, C.c_int, _f64p, _vp, _vp, _vp, _i64p]),
    "sbx_voronoi_plan_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, _i64p]),
    "sbx_voronoi_plan_execute": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int, _f64p, _vp, _vp, _vp, _i64p]),
    "sbx_voronoi_plan_execute_range": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int, _f64p, _vp, _vp, C.c_int64, C.c_int64, _vp, _i64p]),
    "sbx_voronoi_plan_last_path": (C.c_int, [_vp]),
    "sbx_voronoi_plan_destroy": (C.c_int, [_vp]),
    "sbx_voronoi_label_noalloc": (C.c_int, [C.c_int, C.c_int, _i64p, _vp, C.c_int64, _f64p, _vp, _vp, C.c_int64, C.c_int64, _vp]),
    "sbx_scatter_slab_plan_create": (C.c_int, [C.POINTER(_vp), _vp, C.c_int, C.c_int, C.c_int, _i64p, _f64p, C.c_int, C.c_double]),
    "sbx_scatter_slab_shells": (C.c_int, [_vp, C.POINTER(C.c_int), _f64p, _i64p]),
    "sbx_scatter_slab_execute": (C.c_int, [_vp, _vp, _vp, _vp]),
    "sbx_scatter_slab_plan_destroy": (C.c_int, [_vp]),
    "sbx_scatter_slab_ipc_export": (C.c_int, [_vp, _vp]),
    "sbx_scatter_slab_ipc_attach": (C.c_int, [_vp, _vp]),
    "sbx_scatter_slab_ipc_detach": (C.c_int, [_vp]),
    "sbx_scatter_slab_exchange": (C.c_int, [_vp]),
    "sbx_scatter_slab_set_profiling": (C.c_int, [_vp, C.c_int]),
    "sbx_scatter_slab_profile": (C.c_int, [_vp, _f64p]),
    "sbx_scatter_slab_emulate": (C.c_int, [C.c_int, C.c_int, _i64p, _f64p, C.c_int, C.c_double, _vp, _vp, _vp]),
    "sbx_nccl_unique_id": (C.c_int, [_vp]),
    "sbx_nccl_comm_create": (C.c_int, [C.POINTER(_vp), _vp, C.c_int, C.c_int, C.c_int]),
    "sbx_nccl_comm_destroy": (C.c_int, [_vp]),
}

_lib = None


class SbxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libsimubox_b200 error {code}: {msg}")
        self.code = code


def lib() -> C.CDLL:
    """Load the shared library (once).  Raises if it has not been built -- never falls back."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -m simubox_b200.build` "
                "(nvcc, sm_100a).  simubox_b200 has no CPU fallback.")
        handle = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so does not export it
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(code: int) -> None:
    if code != 0:
        raise SbxError(code, lib().sbx_last_error().decode("utf-8", "replace"))


def i64(values):
    arr = np.ascontiguousarray(values, dtype=np.int64)
    return arr, arr.ctypes.data_as(_i64p)


def f64(values):
    arr = np.ascontiguousarray(values, dtype=np.float64)
    return arr, arr.ctypes.data_as(_f64p)


class ScatterPlan:
    """Owner of an ``sbx_scatter_plan`` handle plus its field-independent geometry."""

    def __init__(self, N, L, cutoff=30, threshold=1e-12, max_batch=1, device=0, path=PATH_AUTO):
        L_ = lib()
        self.N = tuple(int(n) for n in N)
        self.L = tuple(float(l) for l in L)
        self.ndim = len(self.N)
        self.cutoff, self.threshold, self.max_batch, self.device = int(cutoff), float(threshold), int(max_batch), int(device)
        self._h = _vp()
        _, pn = i64(self.N)
        _, pl = f64(self.L)
        check(L_.sbx_scatter_plan_create_ex(C.byref(self._h), self.device, self.ndim, pn, pl, self.cutoff,
                                            self.threshold, self.max_batch, int(path)))
        n = C.c_int()
        check(L_.sbx_scatter_plan_shells(self._h, C.byref(n), None, None))
        self.n_shells = n.value
        self.q = np.empty(self.n_shells, dtype=np.float64)
        self.count = np.empty(self.n_shells, dtype=np.int64)
        check(L_.sbx_scatter_plan_shells(self._h, C.byref(n), self.q.ctypes.data_as(_f64p), self.count.ctypes.data_as(_i64p)))
        self.path = L_.sbx_scatter_plan_path(self._h)

    @property
    def handle(self):
        return self._h

    @property
    def kernel(self) -> str:
        """Dominant hand-written kernel of the last execute (empty before the first one)."""
        return (lib().sbx_scatter_plan_kernel(self._h) or b"").decode()

    def members(self, shell: int) -> np.ndarray:
        """(n, ndim) squared-k component vectors of the modes in ``shell`` (= q_idx[shell])."""
        n = C.c_int64()
        check(lib().sbx_scatter_plan_members(self._h, shell, C.byref(n), None, 0))
        out = np.empty((n.value, self.ndim), dtype=np.float64)
        check(lib().sbx_scatter_plan_members(self._h, shell, C.byref(n), out.ctypes.data_as(_f64p), n.value))
        return out

    def q_idx(self) -> dict:
        return {s: list(self.members(s)) for s in range(self.n_shells)}

    def set_profiling(self, enable: bool):
        check(lib().sbx_scatter_plan_set_profiling(self._h, int(enable)))

    def profile(self, reset: bool = True) -> dict:
        a, b = C.c_double(), C.c_double()
        n, m = C.c_int64(), C.c_int64()
        check(lib().sbx_scatter_plan_get_profile(self._h, C.byref(a), C.byref(b), C.byref(n), C.byref(m), int(reset)))
        return dict(transform_ms=a.value, fused_ms=b.value, own_launches=n.value, transform_calls=m.value)

    def close(self):
        if self._h:
            lib().sbx_scatter_plan_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
