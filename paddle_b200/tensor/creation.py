"""Tensor creation (reference: python/paddle/tensor/creation.py)."""
from __future__ import annotations

import numpy as np
import torch

from ..framework import dtype as _dt
from ..framework.place import current_torch_device, to_torch_device


def _shape(shape):
    if isinstance(shape, torch.Tensor):
        return [int(v) for v in shape.tolist()]
    if isinstance(shape, (int, np.integer)):
        return [int(shape)]
    return [int(s.item()) if isinstance(s, torch.Tensor) else int(s) for s in shape]


def _dtype_or_default(dtype):
    return _dt.to_torch_dtype(dtype) if dtype is not None else _dt.default_torch_dtype()


def to_tensor(data, dtype=None, place=None, stop_gradient=True):
    dev = to_torch_device(place)
    td = _dt.to_torch_dtype(dtype)
    if isinstance(data, torch.Tensor):
        t = data.detach().to(device=dev, dtype=td if td is not None else data.dtype)
        if t is data or t.data_ptr() == data.data_ptr():
            t = t.clone()
    else:
        if isinstance(data, np.generic):
            data = np.asarray(data)         # numpy scalars keep their dtype (python scalars take the default one)
        if isinstance(data, np.ndarray):
            if data.dtype == np.uint16:  # bf16 carried as uint16
                t = torch.from_numpy(data.copy()).view(torch.bfloat16)
            else:
                t = torch.from_numpy(np.array(data, order="C"))      # a copy: the tensor never aliases the caller's array
        else:
            arr = np.asarray(data)
            if arr.dtype == np.float64 and td is None:
                # python floats default to the default dtype, like the reference
                t = torch.from_numpy(arr).to(_dt.default_torch_dtype())
            elif arr.dtype == object:
                t = torch.stack([to_tensor(d, dtype, "cpu") for d in data])
            else:
                t = torch.from_numpy(np.ascontiguousarray(arr))
        t = t.to(device=dev, dtype=td if td is not None else t.dtype)
    t.stop_gradient = stop_gradient
    return t


def zeros(shape, dtype=None, name=None):
    return torch.zeros(_shape(shape), dtype=_dtype_or_default(dtype), device=current_torch_device())


def ones(shape, dtype=None, name=None):
    return torch.ones(_shape(shape), dtype=_dtype_or_default(dtype), device=current_torch_device())


def full(shape, fill_value, dtype=None, name=None):
    if isinstance(fill_value, torch.Tensor):
        fill_value = fill_value.item()
    if dtype is None:
        dtype = "bool" if isinstance(fill_value, bool) else ("int64" if isinstance(fill_value, int) else None)
    return torch.full(_shape(shape), fill_value, dtype=_dtype_or_default(dtype), device=current_torch_device())


def empty(shape, dtype=None, name=None):
    return torch.empty(_shape(shape), dtype=_dtype_or_default(dtype), device=current_torch_device())


def zeros_like(x, dtype=None, name=None):
    return torch.zeros_like(x, dtype=_dt.to_torch_dtype(dtype))


def ones_like(x, dtype=None, name=None):
    return torch.ones_like(x, dtype=_dt.to_torch_dtype(dtype))


def full_like(x, fill_value, dtype=None, name=None):
    return torch.full_like(x, fill_value, dtype=_dt.to_torch_dtype(dtype))


def empty_like(x, dtype=None, name=None):
    return torch.empty_like(x, dtype=_dt.to_torch_dtype(dtype))


def arange(start=0, end=None, step=1, dtype=None, name=None):
    if end is None:
        start, end = 0, start
    vals = [v.item() if isinstance(v, torch.Tensor) else v for v in (start, end, step)]
    if dtype is None:
        dtype = "int64" if all(isinstance(v, (int, np.integer)) for v in vals) else None
    return torch.arange(*vals, dtype=_dtype_or_default(dtype), device=current_torch_device())


def linspace(start, stop, num, dtype=None, name=None):
    return torch.linspace(float(start), float(stop), int(num), dtype=_dtype_or_default(dtype),
                          device=current_torch_device())


def eye(num_rows, num_columns=None, dtype=None, name=None):
    return torch.eye(num_rows, num_columns if num_columns is not None else num_rows,
                     dtype=_dtype_or_default(dtype), device=current_torch_device())


def tril(x, diagonal=0, name=None):
    return torch.tril(x, diagonal)


def triu(x, diagonal=0, name=None):
    return torch.triu(x, diagonal)


def diag(x, offset=0, padding_value=0, name=None):
    out = torch.diag(x, offset)
    if padding_value != 0 and x.dim() == 1:
        mask = torch.diag(torch.ones_like(x, dtype=torch.bool), offset)
        out = torch.where(mask, out, torch.full_like(out, padding_value))
    return out


def meshgrid(*args, **kwargs):
    if len(args) == 1 and isinstance(args[0], (list, tuple)):
        args = args[0]
    return list(torch.meshgrid(*args, indexing="ij"))


def assign(x, output=None):
    src = x if isinstance(x, torch.Tensor) else to_tensor(x)
    if output is None:
        return src.clone()
    with torch.no_grad():
        output.copy_(src)
    return output


def clone(x, name=None):
    return x.clone()


def complex(real, imag, name=None):
    return torch.complex(real, imag)


def tolist(x):
    return x.tolist()


def create_tensor(dtype, name=None, persistable=False):
    """An empty (0-element) tensor of ``dtype`` to be filled later (reference: tensor/creation.py create_tensor)."""
    from ..framework import dtype as _dtm
    t = torch.empty(0, dtype=_dtm.to_torch_dtype(dtype))
    if name:
        t.name = name
    t.persistable = persistable
    return t
