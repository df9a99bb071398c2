"""Shared helpers for the op wrappers: dtype codes, raw pointers, current stream, native/eager dispatch rule."""
from __future__ import annotations

import torch

from .. import lib
from ..framework.flags import flag

_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


def dtype_code(t) -> int:
    d = t if isinstance(t, torch.dtype) else t.dtype
    try:
        return _DT[d]
    except KeyError:
        raise TypeError(f"native kernels support float32/bfloat16/float16, got {d}") from None


def ptr(t):
    return None if t is None else t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def use_native(*tensors) -> bool:
    """True when the hand-written kernel must run: all tensors on a CUDA device and the switch is on.
    A GPU tensor with the library missing raises instead of silently taking the torch path."""
    t0 = next((t for t in tensors if isinstance(t, torch.Tensor)), None)
    if t0 is None or t0.device.type != "cuda":
        return False
    if not flag("use_native_kernels"):
        return False
    if t0.dtype not in _DT:
        return False
    lib.kernels.require()
    return True


def K():
    """The kernel library through its launch-counting proxy."""
    lib.kernels.require()
    return lib.kernels


check = lib.check
