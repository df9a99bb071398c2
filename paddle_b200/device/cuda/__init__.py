"""paddle.device.cuda (reference: python/paddle/device/cuda/__init__.py; memory stats from paddle/fluid/memory/stats.cc)."""
import torch

Stream = torch.cuda.Stream
Event = torch.cuda.Event
current_stream = torch.cuda.current_stream
stream_guard = torch.cuda.stream
synchronize = torch.cuda.synchronize
device_count = torch.cuda.device_count
empty_cache = torch.cuda.empty_cache
max_memory_allocated = torch.cuda.max_memory_allocated
max_memory_reserved = torch.cuda.max_memory_reserved
memory_allocated = torch.cuda.memory_allocated
memory_reserved = torch.cuda.memory_reserved
reset_max_memory_allocated = torch.cuda.reset_peak_memory_stats
reset_max_memory_reserved = torch.cuda.reset_peak_memory_stats

from . import graphs  # noqa: E402,F401
from .graphs import CUDAGraph, wrap_cuda_graph  # noqa: E402,F401


def get_device_properties(device=None):
    return torch.cuda.get_device_properties(device if device is not None else torch.cuda.current_device())


def get_device_name(device=None):
    return torch.cuda.get_device_name(device)


def get_device_capability(device=None):
    return torch.cuda.get_device_capability(device)
