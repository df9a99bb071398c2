"""Shape manipulation and random sampling with Paddle signatures
(reference: python/paddle/tensor/{manipulation,random}.py)."""
from __future__ import annotations

import numpy as np
import torch

from ..framework import dtype as _dt
from ..framework.place import current_torch_device
from .creation import _shape


def _ints(v):
    if isinstance(v, torch.Tensor):
        return [int(a) for a in v.tolist()] if v.dim() else int(v.item())
    if isinstance(v, (list, tuple)):
        return [int(a.item()) if isinstance(a, torch.Tensor) else int(a) for a in v]
    return int(v)


def reshape(x, shape, name=None):
    shp = list(_shape(shape))
    if 0 in shp:                            # 0 keeps the input's dim at that position (reference: manipulation.py reshape)
        shp = [x.shape[i] if d == 0 else d for i, d in enumerate(shp)]
    return x.reshape(shp)


def reshape_(x, shape, name=None):
    x.data = x.data.reshape(_shape(shape))
    return x


def transpose(x, perm, name=None):
    return x.permute(*_ints(perm))


def squeeze(x, axis=None, name=None):
    if axis is None:
        return x.squeeze()
    axis = _ints(axis)
    if isinstance(axis, list):
        axis = tuple(a for a in axis if x.shape[a] == 1)
        return x.squeeze(axis) if axis else x
    return x.squeeze(axis) if x.shape[axis] == 1 else x


def unsqueeze(x, axis, name=None):
    axis = _ints(axis)
    if isinstance(axis, list):
        for a in axis:
            x = x.unsqueeze(a)
        return x
    return x.unsqueeze(axis)


def flatten(x, start_axis=0, stop_axis=-1, name=None):
    if x.dim() == 0:
        return x.reshape(1)
    return torch.flatten(x, start_axis, stop_axis)


def concat(x, axis=0, name=None):
    return torch.cat(list(x), dim=_ints(axis))


def stack(x, axis=0, name=None):
    return torch.stack(list(x), dim=axis)


def split(x, num_or_sections, axis=0, name=None):
    """``num_or_sections`` int = number of equal parts (NOT the part size, unlike torch.split)."""
    axis = _ints(axis)
    n = x.shape[axis]
    if isinstance(num_or_sections, int):
        if n % num_or_sections != 0:
            raise ValueError(f"The input's size along the split dimension ({n}) must be evenly divisible by "
                             f"num_or_sections ({num_or_sections})")
        return list(torch.split(x, n // num_or_sections, dim=axis))
    secs = _ints(num_or_sections)
    if -1 in secs:
        known = sum(s for s in secs if s != -1)
        secs = [n - known if s == -1 else s for s in secs]
    return list(torch.split(x, secs, dim=axis))


def chunk(x, chunks, axis=0, name=None):
    return split(x, chunks, axis)


def unbind(input, axis=0):
    return list(torch.unbind(input, dim=axis))


def unstack(x, axis=0, num=None):
    return unbind(x, axis)


def slice(input, axes, starts, ends):
    idx = [builtins_slice(None)] * input.dim()
    for a, s, e in zip(_ints(axes), _ints(starts), _ints(ends)):
        idx[a] = builtins_slice(s, e)
    return input[tuple(idx)]


import builtins as _b  # noqa: E402

builtins_slice = _b.slice


def strided_slice(x, axes, starts, ends, strides, name=None):
    idx = [builtins_slice(None)] * x.dim()
    for a, s, e, st in zip(_ints(axes), _ints(starts), _ints(ends), _ints(strides)):
        idx[a] = builtins_slice(s, e, st)
    return x[tuple(idx)]


def expand(x, shape, name=None):
    shape = _shape(shape)
    return x.expand(*shape)


def expand_as(x, y, name=None):
    return x.expand_as(y)


def broadcast_to(x, shape, name=None):
    return x.expand(*_shape(shape))


def broadcast_tensors(input, name=None):
    return list(torch.broadcast_tensors(*input))


def tile(x, repeat_times, name=None):
    return x.repeat(*_ints(repeat_times)) if len(_ints(repeat_times)) >= x.dim() else x.repeat(
        *([1] * (x.dim() - len(_ints(repeat_times))) + _ints(repeat_times)))


def repeat_interleave(x, repeats, axis=None, name=None):
    return torch.repeat_interleave(x, repeats, dim=axis)


def flip(x, axis, name=None):
    axis = _ints(axis)
    return torch.flip(x, axis if isinstance(axis, list) else [axis])


def roll(x, shifts, axis=None, name=None):
    return torch.roll(x, _ints(shifts), _ints(axis) if axis is not None else None)


def gather(x, index, axis=None, name=None):
    axis = 0 if axis is None else _ints(axis)
    return torch.index_select(x, axis, index.flatten().long())


def gather_nd(x, index, name=None):
    idx = index.long()
    return x[tuple(idx[..., i] for i in range(idx.shape[-1]))]


def take_along_axis(arr, indices, axis, broadcast=True):
    return torch.take_along_dim(arr, indices.long(), dim=axis)


def put_along_axis(arr, indices, values, axis, reduce="assign", include_self=True, broadcast=True):
    values = values if isinstance(values, torch.Tensor) else torch.as_tensor(values, dtype=arr.dtype, device=arr.device)
    values = values.expand_as(indices) if values.dim() == 0 or values.shape != indices.shape else values
    if reduce == "assign":
        return arr.scatter(axis, indices.long(), values)
    return arr.scatter_reduce(axis, indices.long(), values,
                              reduce={"add": "sum", "mul": "prod", "multiply": "prod", "amin": "amin", "amax": "amax",
                                      "mean": "mean"}[reduce], include_self=include_self)


def scatter(x, index, updates, overwrite=True, name=None):
    """Row scatter along axis 0 (paddle.scatter)."""
    out = x.clone()
    idx = index.flatten().long()
    if overwrite:
        out[idx] = updates
    else:
        out[idx] = 0
        out.index_add_(0, idx, updates)
    return out


def scatter_(x, index, updates, overwrite=True, name=None):
    idx = index.flatten().long()
    with torch.no_grad():
        if overwrite:
            x[idx] = updates
        else:
            x[idx] = 0
            x.index_add_(0, idx, updates)
    return x


def scatter_nd_add(x, index, updates, name=None):
    out = x.clone()
    idx = index.long()
    out.index_put_(tuple(idx[..., i] for i in range(idx.shape[-1])), updates, accumulate=True)
    return out


def scatter_nd(index, updates, shape, name=None):
    return scatter_nd_add(torch.zeros(_shape(shape), dtype=updates.dtype, device=updates.device), index, updates)


def index_add(x, index, axis, value, name=None):
    return torch.index_add(x, axis, index.long(), value)


def index_put(x, indices, value, accumulate=False, name=None):
    return torch.index_put(x, tuple(indices), value, accumulate)


def shard_index(input, index_num, nshards, shard_id, ignore_value=-1):
    size = (index_num + nshards - 1) // nshards
    lo, hi = shard_id * size, (shard_id + 1) * size
    inside = (input >= lo) & (input < hi)
    return torch.where(inside, input - lo, torch.full_like(input, ignore_value))


def moveaxis(x, source, destination, name=None):
    return torch.movedim(x, source, destination)


def swapaxes(x, axis0, axis1):
    return x.transpose(axis0, axis1)


def as_real(x, name=None):
    return torch.view_as_real(x)


def as_complex(x, name=None):
    return torch.view_as_complex(x.contiguous())


def shape(x):
    return torch.tensor(list(x.shape), dtype=torch.int32)


def rank(x):
    return torch.tensor(x.dim(), dtype=torch.int32)


def one_hot(x, num_classes, name=None):
    return torch.nn.functional.one_hot(x.long(), num_classes).to(_dt.default_torch_dtype())


def pad_nd(x, pad, mode="constant", value=0.0):
    return torch.nn.functional.pad(x, pad, mode=mode, value=value)


def tensordot(x, y, axes=2, name=None):
    return torch.tensordot(x, y, dims=axes)


def view(x, shape_or_dtype, name=None):
    if isinstance(shape_or_dtype, (list, tuple)):
        return x.view(*shape_or_dtype)
    return x.view(_dt.to_torch_dtype(shape_or_dtype))


def crop(x, shape=None, offsets=None, name=None):
    offsets = offsets or [0] * x.dim()
    idx = tuple(builtins_slice(o, o + (s if s != -1 else x.shape[i] - o)) for i, (o, s) in enumerate(zip(offsets, shape)))
    return x[idx]


# ------------------------------------------------------------------ random
def _rand_dtype(dtype):
    return _dt.to_torch_dtype(dtype) if dtype is not None else _dt.default_torch_dtype()


def rand(shape, dtype=None, name=None):
    return torch.rand(_shape(shape), dtype=_rand_dtype(dtype), device=current_torch_device())


def randn(shape, dtype=None, name=None):
    return torch.randn(_shape(shape), dtype=_rand_dtype(dtype), device=current_torch_device())


def standard_normal(shape, dtype=None, name=None):
    return randn(shape, dtype)


def normal(mean=0.0, std=1.0, shape=None, name=None):
    if isinstance(mean, torch.Tensor) or isinstance(std, torch.Tensor):
        return torch.normal(mean if isinstance(mean, torch.Tensor) else float(mean),
                            std if isinstance(std, torch.Tensor) else float(std))
    return torch.randn(_shape(shape), dtype=_dt.default_torch_dtype(), device=current_torch_device()) * std + mean


def uniform(shape, dtype=None, min=-1.0, max=1.0, seed=0, name=None):
    return torch.rand(_shape(shape), dtype=_rand_dtype(dtype), device=current_torch_device()) * (max - min) + min


def randint(low=0, high=None, shape=(1,), dtype=None, name=None):
    if high is None:
        low, high = 0, low
    return torch.randint(int(low), int(high), _shape(shape), dtype=_dt.to_torch_dtype(dtype or "int64"),
                         device=current_torch_device())


def randint_like(x, low=0, high=None, dtype=None, name=None):
    if high is None:
        low, high = 0, low
    return torch.randint(int(low), int(high), tuple(x.shape), dtype=_dt.to_torch_dtype(dtype) or x.dtype, device=x.device)


def randperm(n, dtype="int64", name=None):
    return torch.randperm(int(n), dtype=_dt.to_torch_dtype(dtype), device=current_torch_device())


def bernoulli(x, name=None):
    return torch.bernoulli(x)


def multinomial(x, num_samples=1, replacement=False, name=None):
    return torch.multinomial(x, num_samples, replacement)


def poisson(x, name=None):
    return torch.poisson(x)


def put_along_axis_(arr, indices, values, axis, reduce="assign", include_self=True, broadcast=True):
    """In-place ``put_along_axis``."""
    out = put_along_axis(arr, indices, values, axis, reduce, include_self, broadcast)
    with torch.no_grad():
        arr.copy_(out)
    return arr
