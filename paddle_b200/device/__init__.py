"""paddle.device (reference: python/paddle/device/__init__.py, device/cuda/__init__.py)."""
from __future__ import annotations

import torch

from ..framework.place import (CPUPlace, CUDAPinnedPlace, CUDAPlace, get_device, is_compiled_with_cuda,  # noqa: F401
                               set_device)
from . import cuda  # noqa: F401


def synchronize(device=None):
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def get_all_device_type():
    return ["cpu"] + (["gpu"] if torch.cuda.is_available() else [])


def get_available_device():
    return ["cpu"] + [f"gpu:{i}" for i in range(torch.cuda.device_count())]


def is_compiled_with_xpu():
    return False


def is_compiled_with_rocm():
    return False


def is_compiled_with_custom_device(device_type=None):
    return False


def is_compiled_with_distribute():
    return True


def device_count():
    return torch.cuda.device_count()


Stream = torch.cuda.Stream
Event = torch.cuda.Event
current_stream = torch.cuda.current_stream
stream_guard = torch.cuda.stream


def get_cudnn_version():
    import torch
    v = torch.backends.cudnn.version()
    return int(v) if v else None


def is_compiled_with_ipu():
    return False


def is_compiled_with_cinn():
    return False


def get_all_custom_device_type():
    return []


def get_available_custom_device():
    return []


def set_stream(stream=None):
    """Make ``stream`` the current CUDA stream; returns the previous one."""
    import torch
    prev = torch.cuda.current_stream()
    if stream is not None:
        torch.cuda.set_stream(stream)
    return prev


class XPUPlace:
    def __init__(self, *a, **k):
        raise NotImplementedError("XPU devices are not supported: this framework targets NVIDIA B200 only")


class IPUPlace(XPUPlace):
    pass
