"""paddle_b200: a Blackwell-native (sm_100a) training framework exposing PaddlePaddle's public API.

``import paddle_b200 as paddle`` or simply ``import paddle`` (the alias package at the repo root) gives the
namespace a user of the reference expects: ``paddle.nn``, ``paddle.optimizer``, ``paddle.amp``, ``paddle.io``,
``paddle.distributed`` (+ ``fleet``), ``paddle.incubate``, ``paddle.vision``, ``paddle.profiler`` ...
PyTorch is the tensor/autograd/allocator layer; hot ops are hand-written CUDA kernels for sm_100a in
``csrc/kernels`` (C ABI in ``include/pdb200_kernels.h``); the host runtime (store, watchdog, reducer,
symmetric-memory arena) is C++ in ``csrc/runtime``.  See DESIGN.md.
"""
from __future__ import annotations

__version__ = "0.1.0"

from . import _alias

_alias.install()

import torch as _torch  # noqa: E402

from .framework import tensor_patch as _tensor_patch  # noqa: E402,F401  (patches torch.Tensor)
from .framework.dtype import (bfloat16, bool, complex64, complex128, dtype, finfo, float8_e4m3fn, float8_e5m2,  # noqa: E402,F401,A004
                              float16, float32, float64, get_default_dtype, iinfo, int8, int16, int32, int64,
                              set_default_dtype, uint8)
from .framework.flags import get_flags, set_flags  # noqa: E402,F401
from .framework.io import async_save, load, save  # noqa: E402,F401
from .framework.param_attr import ParamAttr  # noqa: E402,F401
from .framework.place import (CPUPlace, CUDAPinnedPlace, CUDAPlace, get_device, is_compiled_with_cuda,  # noqa: E402,F401
                              set_device)
from .framework.random import get_cuda_rng_state, get_rng_state, seed, set_cuda_rng_state, set_rng_state  # noqa: E402,F401
from .tensor import *  # noqa: E402,F401,F403
from .tensor import creation as _creation  # noqa: E402
from .autograd import enable_grad, grad, is_grad_enabled, no_grad, set_grad_enabled  # noqa: E402,F401
from . import (amp, autograd, base, device, framework, incubate, io, metric, nn, optimizer, profiler, regularizer,  # noqa: E402,F401
               utils, vision)
from . import distributed  # noqa: E402,F401
from .distributed.parallel import DataParallel  # noqa: E402,F401
from .hapi import Model, summary  # noqa: E402,F401
from . import models, ops, parallel  # noqa: E402,F401
from . import (audio, callbacks, cost_model, distribution, fft, geometric, hub, inference, jit, linalg, onnx, quantization, signal,  # noqa: E402,F401
               sparse, static, sysconfig, text)
from .batch import batch  # noqa: E402,F401

from .tensor.compat import install_inplace_variants as _iiv  # noqa: E402
from .tensor._inplace_names import INPLACE as _INPLACE  # noqa: E402

_iiv(globals(), _INPLACE)
from .framework import tensor_methods as _tensor_methods  # noqa: E402
_tensor_methods.install(globals())       # x.max(axis=...), x.split(n, axis=...), ...: Paddle keywords -> Paddle semantics
Tensor = _torch.Tensor
version = type("version", (), {"full_version": __version__, "cuda": staticmethod(lambda: _torch.version.cuda),
                                "show": staticmethod(lambda: print(__version__))})


def in_dynamic_mode():
    return True


def enable_static():
    raise NotImplementedError("static-graph mode is out of scope for this framework (SURVEY §2.4 R10)")


def disable_static(place=None):
    if place is not None:
        set_device(place)


def is_compiled_with_xpu():
    return False


def is_compiled_with_rocm():
    return False


def is_compiled_with_cinn():
    return False


def is_compiled_with_distribute():
    return True


def is_compiled_with_custom_device(name=None):
    return False


def disable_signal_handler():
    pass


def set_printoptions(precision=None, threshold=None, edgeitems=None, sci_mode=None, linewidth=None):
    _torch.set_printoptions(precision=precision, threshold=threshold, edgeitems=edgeitems, sci_mode=sci_mode,
                            linewidth=linewidth)


def set_grad_enabled_(mode):
    return _torch.set_grad_enabled(mode)
