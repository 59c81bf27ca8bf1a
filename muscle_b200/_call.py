"""Shared plumbing of the op wrappers: argument checks and the C-ABI call convention."""
from __future__ import annotations

import torch

from . import _lib


def stream_ptr(device):
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda_f32(t: torch.Tensor, name: str):
    if not torch.is_tensor(t) or not t.is_cuda:
        raise RuntimeError("muscle_b200: `%s` must be a CUDA tensor (there is no CPU path)" % name)
    if t.dtype != torch.float32:
        raise TypeError("muscle_b200: `%s` must be float32 (the reference path is fp32), got %s" % (name, t.dtype))


def call(name, device, *args):
    """Invokes C-ABI function `name` with the current stream of `device` appended; raises on error."""
    lib = _lib.load()
    with torch.cuda.device(device):
        rc = getattr(lib, name)(*args, stream_ptr(device))
    _lib.check(rc, name)


def ptr(t):
    return None if t is None else t.data_ptr()
