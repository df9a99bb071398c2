"""paddle.base.core: the subset of the pybind module that user-level distributed code touches
(reference: paddle/fluid/pybind/*.cc).  Native pieces are bound from libpdb200_runtime.so."""
from __future__ import annotations

import torch

from ..framework.place import CPUPlace, CUDAPinnedPlace, CUDAPlace, Place  # noqa: F401
from ..framework.place import is_compiled_with_cuda  # noqa: F401

eager = type("eager", (), {"Tensor": torch.Tensor})


def is_compiled_with_nccl():
    import torch.distributed as dist
    return dist.is_nccl_available()


def is_compiled_with_xpu():
    return False


def is_compiled_with_rocm():
    return False


def is_compiled_with_custom_device(name=None):
    return False


def is_compiled_with_dist():
    return True


def get_cuda_device_count():
    return torch.cuda.device_count()


def nccl_version():
    try:
        v = torch.cuda.nccl.version()
        return v[0] * 1000 + v[1] * 100 + v[2]
    except Exception:  # noqa: BLE001
        return 0


def globals():  # noqa: A001
    from ..framework import flags
    return {("FLAGS_" + k): v["value"] for k, v in flags._REGISTRY.items()}
