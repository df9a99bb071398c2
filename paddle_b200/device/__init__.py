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
