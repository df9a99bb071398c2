"""Further paddle.* tensor functions (reference: python/paddle/tensor/{math,manipulation,random,search,stat}.py)."""
from __future__ import annotations

import torch

from ..framework import dtype as _dt


def diagonal(x, offset=0, axis1=0, axis2=1, name=None):
    return torch.diagonal(x, offset, axis1, axis2)


def bucketize(x, sorted_sequence, out_int32=False, right=False, name=None):
    return torch.bucketize(x, sorted_sequence, out_int32=out_int32, right=right)


def erfinv(x, name=None):
    return torch.erfinv(x)


def frac(x, name=None):
    return torch.frac(x)


def logcumsumexp(x, axis=None, dtype=None, name=None):
    if axis is None:
        x, axis = x.reshape(-1), 0
    return torch.logcumsumexp(x if dtype is None else x.to(_dt.to_torch_dtype(dtype)), axis)


def nansum(x, axis=None, dtype=None, keepdim=False, name=None):
    return torch.nansum(x, dim=axis, keepdim=keepdim) if axis is not None else torch.nansum(x)


def nanmean(x, axis=None, keepdim=False, name=None):
    return torch.nanmean(x, dim=axis, keepdim=keepdim) if axis is not None else torch.nanmean(x)


def diff(x, n=1, axis=-1, prepend=None, append=None, name=None):
    return torch.diff(x, n=n, dim=axis, prepend=prepend, append=append)


def quantile(x, q, axis=None, keepdim=False, interpolation="linear", name=None):
    qt = torch.as_tensor(q, dtype=x.dtype, device=x.device)
    return torch.quantile(x, qt, dim=axis, keepdim=keepdim, interpolation=interpolation)


def nanquantile(x, q, axis=None, keepdim=False, interpolation="linear", name=None):
    qt = torch.as_tensor(q, dtype=x.dtype, device=x.device)
    return torch.nanquantile(x, qt, dim=axis, keepdim=keepdim, interpolation=interpolation)


def unflatten(x, axis, shape, name=None):
    return torch.unflatten(x, axis, tuple(shape))


def as_strided(x, shape, stride, offset=0, name=None):
    return torch.as_strided(x, tuple(shape), tuple(stride), offset)


def rand_like(x, dtype=None, name=None):
    return torch.rand_like(x, dtype=_dt.to_torch_dtype(dtype) if dtype is not None else None)


def randn_like(x, dtype=None, name=None):
    return torch.randn_like(x, dtype=_dt.to_torch_dtype(dtype) if dtype is not None else None)


def heaviside(x, y, name=None):
    return torch.heaviside(x, y)


def gcd(x, y, name=None):
    return torch.gcd(x, y)


def lcm(x, y, name=None):
    return torch.lcm(x, y)


def inner(x, y, name=None):
    return torch.inner(x, y)


def renorm(x, p, axis, max_norm):
    return torch.renorm(x, p, axis, max_norm)


def vander(x, n=None, increasing=False, name=None):
    return torch.vander(x, N=n, increasing=increasing)


def cdist(x, y, p=2.0, compute_mode="use_mm_for_euclid_dist_if_necessary", name=None):
    return torch.cdist(x, y, p=p, compute_mode=compute_mode)


def shape_(x):
    return list(x.shape)


def cummax(x, axis=None, dtype="int64", name=None):
    if axis is None:
        x, axis = x.reshape(-1), 0
    v, i = torch.cummax(x, axis)
    return v, i.to(_dt.to_torch_dtype(dtype))


def cummin(x, axis=None, dtype="int64", name=None):
    if axis is None:
        x, axis = x.reshape(-1), 0
    v, i = torch.cummin(x, axis)
    return v, i.to(_dt.to_torch_dtype(dtype))


def trapezoid(y, x=None, dx=None, axis=-1, name=None):
    return torch.trapezoid(y, x=x, dim=axis) if x is not None else torch.trapezoid(y, dx=dx if dx is not None else 1.0, dim=axis)


def hypot(x, y, name=None):
    return torch.hypot(x, y)


def copysign(x, y, name=None):
    return torch.copysign(x, y)


def ldexp(x, y, name=None):
    return torch.ldexp(x, y)


def nextafter(x, y, name=None):
    return torch.nextafter(x, y)


def polygamma(x, n, name=None):
    return torch.polygamma(n, x)


def i0(x, name=None):
    return torch.i0(x)


def sinc(x, name=None):
    return torch.sinc(x)


def block_diag(inputs, name=None):
    return torch.block_diag(*inputs)


def cartesian_prod(x, name=None):
    return torch.cartesian_prod(*x)


def combinations(x, r=2, with_replacement=False, name=None):
    return torch.combinations(x, r, with_replacement)


def isin(x, test_x, assume_unique=False, invert=False, name=None):
    return torch.isin(x, test_x, assume_unique=assume_unique, invert=invert)


def signbit(x, name=None):
    return torch.signbit(x)
