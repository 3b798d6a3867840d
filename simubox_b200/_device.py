"""torch as the buffer carrier: device tensors, pinned host memory and the current stream.
No torch op takes part in the analysis itself."""
from __future__ import annotations

import ctypes as C

import numpy as np


def torch_cuda():
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError("simubox_b200 needs a CUDA device (sm_100); there is no CPU fallback")
    return torch


def ptr(t) -> C.c_void_p:
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(None)


def stream_ptr(device) -> C.c_void_p:
    torch = torch_cuda()
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def to_device(arr, device, dtype=None):
    """numpy array or torch tensor -> contiguous torch tensor on ``cuda:device``."""
    torch = torch_cuda()
    dev = torch.device("cuda", device)
    if isinstance(arr, torch.Tensor):
        t = arr.to(dev)
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        return t.contiguous()
    a = np.ascontiguousarray(arr)
    t = torch.from_numpy(a).to(dev)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t
