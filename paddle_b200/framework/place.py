"""Places and device selection (reference: paddle/phi/common/place.h, python/paddle/device/__init__.py)."""
from __future__ import annotations

import os

import torch


class Place:
    def __init__(self, kind: str, idx: int = 0):
        self._kind = kind
        self._idx = idx

    def is_gpu_place(self):
        return self._kind == "gpu"

    def is_cpu_place(self):
        return self._kind == "cpu"

    def is_cuda_pinned_place(self):
        return self._kind == "gpu_pinned"

    def gpu_device_id(self):
        return self._idx

    def get_device_id(self):
        return self._idx

    def to_torch(self) -> torch.device:
        if self._kind == "gpu":
            return torch.device("cuda", self._idx)
        return torch.device("cpu")

    def __eq__(self, other):
        return isinstance(other, Place) and (self._kind, self._idx) == (other._kind, other._idx)

    def __hash__(self):
        return hash((self._kind, self._idx))

    def __repr__(self):
        if self._kind == "gpu":
            return f"Place(gpu:{self._idx})"
        if self._kind == "gpu_pinned":
            return "Place(gpu_pinned)"
        return "Place(cpu)"

    __str__ = __repr__


class CPUPlace(Place):
    def __init__(self):
        super().__init__("cpu", 0)


class CUDAPlace(Place):
    def __init__(self, idx: int = 0):
        super().__init__("gpu", int(idx))


class CUDAPinnedPlace(Place):
    def __init__(self):
        super().__init__("gpu_pinned", 0)


_current: Place | None = None


def is_compiled_with_cuda():
    return torch.cuda.is_available()


def _default_place() -> Place:
    if torch.cuda.is_available():
        sel = os.environ.get("FLAGS_selected_gpus", "")
        if sel == "":
            sel = os.environ.get("LOCAL_RANK", "0")
        idx = int(sel.split(",")[0]) if sel else 0
        return CUDAPlace(idx % max(torch.cuda.device_count(), 1))
    return CPUPlace()


def set_device(device):
    """paddle.set_device('gpu:0' | 'cpu' | 'gpu')."""
    global _current
    if isinstance(device, Place):
        _current = device
    else:
        d = str(device).lower()
        if d.startswith("gpu") or d.startswith("cuda"):
            idx = int(d.split(":")[1]) if ":" in d else _default_place().get_device_id() if torch.cuda.is_available() else 0
            if not torch.cuda.is_available():
                raise ValueError("The device should not be 'gpu', since this process sees no CUDA device")
            _current = CUDAPlace(idx)
        elif d.startswith("cpu"):
            _current = CPUPlace()
        else:
            raise ValueError(f"unsupported device {device!r} (this build is sm_100a + cpu only)")
    if _current.is_gpu_place():
        torch.cuda.set_device(_current.get_device_id())
    return _current


def get_device() -> str:
    p = current_place()
    return f"gpu:{p.get_device_id()}" if p.is_gpu_place() else "cpu"


def current_place() -> Place:
    global _current
    if _current is None:
        _current = _default_place()
        if _current.is_gpu_place():
            torch.cuda.set_device(_current.get_device_id())
    return _current


def current_torch_device() -> torch.device:
    return current_place().to_torch()


def place_of(t: torch.Tensor) -> Place:
    if t.device.type == "cuda":
        return CUDAPlace(t.device.index or 0)
    if t.device.type == "cpu" and t.is_pinned():
        return CUDAPinnedPlace()
    return CPUPlace()


def to_torch_device(place) -> torch.device:
    if place is None:
        return current_torch_device()
    if isinstance(place, Place):
        return place.to_torch()
    if isinstance(place, torch.device):
        return place
    s = str(place).lower()
    if s.startswith("gpu"):
        return torch.device("cuda", int(s.split(":")[1]) if ":" in s else current_place().get_device_id())
    return torch.device(s)
