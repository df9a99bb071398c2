"""Paddle's Tensor surface on ``torch.Tensor``.

PyTorch is this framework's tensor/autograd/allocator layer; ``paddle.Tensor`` *is*
``torch.Tensor`` with the Paddle-only attributes and methods attached here: ``stop_gradient``, ``name``,
``place``, ``astype``, ``clear_grad``, ``transpose(perm)``, the storage tricks the sharding code relies on
(``_clear_data``, ``_share_buffer_to``) and so on.  Reference behaviour: python/paddle/base/dygraph/
tensor_patch_methods.py, paddle/fluid/pybind/eager_method.cc.
"""
from __future__ import annotations

import numpy as np
import torch

from ..utils import unique_name
from . import dtype as _dt
from .place import place_of, to_torch_device

Tensor = torch.Tensor


# ---------------------------------------------------------------- properties
def _get_stop_gradient(self):
    return not self.requires_grad


def _set_stop_gradient(self, v):
    v = bool(v)
    if v:
        if self.requires_grad:
            if self.is_leaf:
                self.requires_grad_(False)
            else:
                self.detach_()
    else:
        if not self.requires_grad and (self.is_floating_point() or self.is_complex()):
            self.requires_grad_(True)


def _get_name(self):
    n = self.__dict__.get("_pd_name")
    if n is None:
        n = unique_name.generate("generated_tensor")
        self.__dict__["_pd_name"] = n
    return n


def _set_name(self, v):
    self.__dict__["_pd_name"] = v


def _get_persistable(self):
    return self.__dict__.get("_pd_persistable", False)


def _set_persistable(self, v):
    self.__dict__["_pd_persistable"] = bool(v)


Tensor.stop_gradient = property(_get_stop_gradient, _set_stop_gradient)
Tensor.name = property(_get_name, _set_name)
Tensor.persistable = property(_get_persistable, _set_persistable)
Tensor.place = property(lambda self: place_of(self))


# ---------------------------------------------------------------- methods
def _astype(self, dtype):
    return self.to(_dt.to_torch_dtype(dtype))


def _clear_grad(self, set_to_zero=False):
    if self.grad is not None:
        if set_to_zero:
            self.grad.zero_()
        else:
            self.grad = None


def _clear_gradient(self, set_to_zero=True):
    _clear_grad(self, set_to_zero)


def _gradient(self):
    return None if self.grad is None else _numpy(self.grad)


_orig_numpy = Tensor.numpy


def _numpy(self, *a, **k):
    t = self.detach()
    if t.device.type != "cpu":
        t = t.cpu()
    if t.dtype == torch.bfloat16:
        # the reference hands bf16 out as uint16 ndarrays (framework/io.py, test_collective_api_base.py:382)
        return _orig_numpy(t.view(torch.uint16))
    return _orig_numpy(t.resolve_conj().resolve_neg(), *a, **k)


_orig_transpose = Tensor.transpose


def _transpose(self, *args, **kwargs):
    if len(args) == 1 and isinstance(args[0], (list, tuple)):
        return self.permute(*args[0])
    if "perm" in kwargs:
        return self.permute(*kwargs["perm"])
    return _orig_transpose(self, *args, **kwargs)


def _clear_data(self):
    """Free the storage but keep the Python object/autograd identity (used by ZeRO-3 to drop full params)."""
    self.data = torch.empty(0, dtype=self.dtype, device=self.device)


_clear_data_alias = _clear_data


def _share_buffer_to(self, other):
    other.data = self.data


def _is_initialized(self):
    return self.numel() > 0 or self.dim() == 0


def _set_value(self, value):
    if isinstance(value, np.ndarray):
        if value.dtype == np.uint16 and self.dtype == torch.bfloat16:
            value = torch.from_numpy(value.copy()).view(torch.bfloat16)
        else:
            value = torch.from_numpy(np.ascontiguousarray(value))
    elif not isinstance(value, torch.Tensor):
        value = torch.as_tensor(value)
    with torch.no_grad():
        if tuple(value.shape) != tuple(self.shape):
            raise ValueError(f"set_value shape mismatch: {tuple(value.shape)} vs {tuple(self.shape)} for {self.name}")
        self.copy_(value.to(device=self.device, dtype=self.dtype))
    return self


def _copy_to(self, place, blocking=True):
    return self.to(to_torch_device(place), non_blocking=not blocking)


def _scale_(self, scale=1.0, bias=0.0, bias_after_scale=True):
    with torch.no_grad():
        if bias_after_scale:
            self.mul_(scale)
            if bias:
                self.add_(bias)
        else:
            if bias:
                self.add_(bias)
            self.mul_(scale)
    return self


def _scale(self, scale=1.0, bias=0.0, bias_after_scale=True):
    return self * scale + bias if bias_after_scale else (self + bias) * scale


def _is_dense(self):
    return self.layout == torch.strided


def _value(self):
    return self


def _get_tensor(self):
    return self


def _ndimension(self):
    return self.dim()


def _cast_(self, dtype):
    self.data = self.data.to(_dt.to_torch_dtype(dtype))
    return self


def _pd_cuda(self, device_id=None, blocking=True):
    dev = torch.device("cuda", device_id if device_id is not None else torch.cuda.current_device())
    return self.to(dev, non_blocking=not blocking)


def _unsqueeze_(self, axis):
    self.data = self.data.unsqueeze(axis)
    return self


def _slice_param(self, begin, end):
    """View of rows [begin, end) sharing storage (reference: Tensor._slice)."""
    return self.detach()[begin:end]


def _numel_prop(self):
    return self.numel()


for _n, _f in {
    "astype": _astype,
    "cast": _astype,
    "clear_grad": _clear_grad,
    "clear_gradient": _clear_gradient,
    "gradient": _gradient,
    "numpy": _numpy,
    "transpose": _transpose,
    "_clear_data": _clear_data,
    "_clear": _clear_data,
    "_share_buffer_to": _share_buffer_to,
    "_is_initialized": _is_initialized,
    "set_value": _set_value,
    "_copy_to": _copy_to,
    "scale_": _scale_,
    "scale": _scale,
    "is_dense": _is_dense,
    "value": _value,
    "get_tensor": _get_tensor,
    "cast_": _cast_,
    "_slice": _slice_param,
}.items():
    setattr(Tensor, _n, _f)

if not hasattr(Tensor, "ndimension"):
    Tensor.ndimension = _ndimension


def _apply(self, func):
    """``func(tensor)`` -> a new tensor; refused on tensors that take part in autograd (tensor_patch_methods.py apply)."""
    if self.requires_grad:
        raise RuntimeError("Cannot apply function on a tensor that required gradient.")
    return func(self)


def _apply_(self, func):
    """In-place ``apply``: the tensor takes the function's result."""
    if self.requires_grad:
        raise RuntimeError("Cannot apply function on a tensor that required gradient.")
    with torch.no_grad():
        self.copy_(func(self))
    return self


if not hasattr(Tensor, "apply"):
    Tensor.apply = _apply
if getattr(Tensor.apply_, "__name__", "") != "_apply_":
    _torch_apply_ = Tensor.apply_

    def _apply_dispatch(self, func):            # torch's apply_ maps a Python scalar function element by element (CPU only)
        try:
            return _apply_(self, func)
        except (TypeError, RuntimeError):
            if self.requires_grad:
                raise
            return _torch_apply_(self, func)
    _apply_dispatch.__name__ = "_apply_"
    Tensor.apply_ = _apply_dispatch
if not hasattr(Tensor, "inplace_version"):
    Tensor.inplace_version = property(lambda self: self._version)


# ---------------------------------------------------------------- parameters
class Parameter(torch.nn.Parameter):
    """EagerParamBase equivalent (reference: python/paddle/base/framework.py EagerParamBase).

    Carries ``name``, ``trainable``, ``optimize_attr``, ``regularizer``, ``need_clip``, ``is_distributed`` and the
    attributes parallel code attaches (``split_axis``, ``no_sync``, ``sequence_parallel``, ``main_grad``...).
    """

    def __new__(cls, data=None, requires_grad=True, name=None, trainable=True, **attrs):
        if data is None:
            data = torch.empty(0)
        p = torch.Tensor._make_subclass(cls, data.detach() if isinstance(data, torch.Tensor) else torch.as_tensor(data),
                                        requires_grad and trainable)
        return p

    def __init__(self, data=None, requires_grad=True, name=None, trainable=True, **attrs):
        self.__dict__["_pd_name"] = name or unique_name.generate("param")
        self.__dict__["_pd_persistable"] = True
        self.optimize_attr = attrs.pop("optimize_attr", {"learning_rate": 1.0})
        self.regularizer = attrs.pop("regularizer", None)
        self.need_clip = attrs.pop("need_clip", True)
        self.is_distributed = attrs.pop("is_distributed", False)
        self.do_model_average = attrs.pop("do_model_average", None)
        for k, v in attrs.items():
            setattr(self, k, v)

    @property
    def trainable(self):
        return self.requires_grad

    @trainable.setter
    def trainable(self, v):
        self.requires_grad_(bool(v))

    def __deepcopy__(self, memo):
        new = Parameter(self.data.clone(memory_format=torch.preserve_format), self.requires_grad,
                        name=unique_name.generate(self.name.rsplit("_", 1)[0] if "_" in self.name else self.name))
        for k, v in self.__dict__.items():
            if k not in ("_pd_name",):
                new.__dict__.setdefault(k, v)
        memo[id(self)] = new
        return new

    def __reduce_ex__(self, proto):
        return (_rebuild_param, (self.data, self.requires_grad, self.name, dict(
            optimize_attr=self.optimize_attr, need_clip=self.need_clip, is_distributed=self.is_distributed)))

    def __repr__(self):
        return f"Parameter containing:\n{super(torch.nn.Parameter, self).__repr__()}"


def _rebuild_param(data, requires_grad, name, attrs):
    return Parameter(data, requires_grad, name=name, **attrs)


EagerParamBase = Parameter
ParamBase = Parameter
