"""Math, reduction, comparison, linalg and search functions with Paddle signatures
(reference: python/paddle/tensor/{math,logic,linalg,search,stat}.py).  Thin maps onto torch ops; ``axis`` replaces
``dim``, reductions return values only, matmul takes transpose flags."""
from __future__ import annotations

import builtins
import math as _math

import numpy as np
import torch

from ..framework import dtype as _dt


def _ax(axis):
    if isinstance(axis, torch.Tensor):
        axis = axis.tolist()
    if isinstance(axis, (list, tuple)):
        return tuple(int(a) for a in axis)
    return axis


def _t(x, like=None):
    if isinstance(x, torch.Tensor):
        return x
    if like is not None:
        return torch.as_tensor(x, dtype=like.dtype if not isinstance(x, bool) else torch.bool, device=like.device)
    return torch.as_tensor(x)


# ---- elementwise binary
def add(x, y, name=None):
    return torch.add(x, y)


def subtract(x, y, name=None):
    return torch.sub(x, y)


def multiply(x, y, name=None):
    return torch.mul(x, y)


def divide(x, y, name=None):
    return torch.div(x, y)


def floor_divide(x, y, name=None):
    return torch.div(x, y, rounding_mode="floor")


def remainder(x, y, name=None):
    return torch.remainder(x, y)


mod = remainder
floor_mod = remainder


def pow(x, y, name=None):
    return torch.pow(x, y)


def maximum(x, y, name=None):
    return torch.maximum(x, _t(y, x))


def minimum(x, y, name=None):
    return torch.minimum(x, _t(y, x))


fmax = lambda x, y, name=None: torch.fmax(x, y)  # noqa: E731
fmin = lambda x, y, name=None: torch.fmin(x, y)  # noqa: E731


def scale(x, scale=1.0, bias=0.0, bias_after_scale=True, act=None, name=None):
    out = x * scale + bias if bias_after_scale else (x + bias) * scale
    return out


def add_n(inputs, name=None):
    if isinstance(inputs, torch.Tensor):
        return inputs.clone()
    out = inputs[0]
    for t in inputs[1:]:
        out = out + t
    return out


def lerp(x, y, weight, name=None):
    return torch.lerp(x, y, _t(weight, x))


def addmm(input, x, y, beta=1.0, alpha=1.0, name=None):
    return torch.addmm(input, x, y, beta=beta, alpha=alpha)


# ---- elementwise unary
def _unary(fn):
    def f(x, name=None):
        return fn(x)
    return f


abs = _unary(torch.abs)
exp = _unary(torch.exp)
expm1 = _unary(torch.expm1)
log = _unary(torch.log)
log2 = _unary(torch.log2)
log10 = _unary(torch.log10)
log1p = _unary(torch.log1p)
sqrt = _unary(torch.sqrt)
rsqrt = _unary(torch.rsqrt)
square = _unary(torch.square)
sin = _unary(torch.sin)
cos = _unary(torch.cos)
tan = _unary(torch.tan)
asin = _unary(torch.asin)
acos = _unary(torch.acos)
atan = _unary(torch.atan)
sinh = _unary(torch.sinh)
cosh = _unary(torch.cosh)
tanh = _unary(torch.tanh)
sigmoid = _unary(torch.sigmoid)
floor = _unary(torch.floor)
ceil = _unary(torch.ceil)
round = _unary(torch.round)
trunc = _unary(torch.trunc)
sign = _unary(torch.sign)
neg = _unary(torch.neg)
reciprocal = _unary(torch.reciprocal)
erf = _unary(torch.erf)
isnan = _unary(torch.isnan)
isinf = _unary(torch.isinf)
isfinite = _unary(torch.isfinite)
logical_not = _unary(torch.logical_not)
bitwise_not = _unary(torch.bitwise_not)
conj = _unary(torch.conj)
real = _unary(torch.real)
imag = _unary(torch.imag)
angle = _unary(torch.angle)
digamma = _unary(torch.digamma)
lgamma = _unary(torch.lgamma)


def atan2(x, y, name=None):
    return torch.atan2(x, y)


def clip(x, min=None, max=None, name=None):
    if isinstance(min, torch.Tensor):
        min = min.item()
    if isinstance(max, torch.Tensor):
        max = max.item()
    return torch.clamp(x, min, max)


def cast(x, dtype):
    return x.to(_dt.to_torch_dtype(dtype))


def stanh(x, scale_a=0.67, scale_b=1.7159, name=None):
    return scale_b * torch.tanh(scale_a * x)


# ---- reductions
def sum(x, axis=None, dtype=None, keepdim=False, name=None):
    td = _dt.to_torch_dtype(dtype)
    if axis is None or (isinstance(axis, (list, tuple)) and len(axis) == 0):
        out = torch.sum(x, dtype=td)
        return out.reshape([1] * x.dim()) if keepdim else out
    return torch.sum(x, dim=_ax(axis), keepdim=keepdim, dtype=td)


def mean(x, axis=None, keepdim=False, name=None):
    if axis is None or (isinstance(axis, (list, tuple)) and len(axis) == 0):
        out = torch.mean(x)
        return out.reshape([1] * x.dim()) if keepdim else out
    return torch.mean(x, dim=_ax(axis), keepdim=keepdim)


def prod(x, axis=None, keepdim=False, dtype=None, name=None):
    td = _dt.to_torch_dtype(dtype)
    if axis is None:
        return torch.prod(x, dtype=td)
    if isinstance(axis, (list, tuple)):
        out = x
        for a in sorted([a % x.dim() for a in axis], reverse=True):
            out = torch.prod(out, dim=a, keepdim=keepdim, dtype=td)
        return out
    return torch.prod(x, dim=axis, keepdim=keepdim, dtype=td)


def max(x, axis=None, keepdim=False, name=None):
    if axis is None:
        out = torch.max(x)
        return out.reshape([1] * x.dim()) if keepdim else out
    return torch.amax(x, dim=_ax(axis), keepdim=keepdim)


def min(x, axis=None, keepdim=False, name=None):
    if axis is None:
        out = torch.min(x)
        return out.reshape([1] * x.dim()) if keepdim else out
    return torch.amin(x, dim=_ax(axis), keepdim=keepdim)


amax = max
amin = min


def all(x, axis=None, keepdim=False, name=None):
    if axis is None:
        return torch.all(x)
    return torch.all(x, dim=_ax(axis), keepdim=keepdim)


def any(x, axis=None, keepdim=False, name=None):
    if axis is None:
        return torch.any(x)
    return torch.any(x, dim=_ax(axis), keepdim=keepdim)


def logsumexp(x, axis=None, keepdim=False, name=None):
    if axis is None:
        axis = list(range(x.dim()))
    return torch.logsumexp(x, dim=_ax(axis), keepdim=keepdim)


def std(x, axis=None, unbiased=True, keepdim=False, name=None):
    return torch.std(x, dim=_ax(axis), unbiased=unbiased, keepdim=keepdim)


def var(x, axis=None, unbiased=True, keepdim=False, name=None):
    return torch.var(x, dim=_ax(axis), unbiased=unbiased, keepdim=keepdim)


def median(x, axis=None, keepdim=False, name=None):
    if axis is None:
        return torch.quantile(x.flatten().float(), 0.5).to(x.dtype)
    return torch.quantile(x.float(), 0.5, dim=axis, keepdim=keepdim).to(x.dtype)


def cumsum(x, axis=None, dtype=None, name=None):
    if axis is None:
        x, axis = x.flatten(), 0
    return torch.cumsum(x, dim=axis, dtype=_dt.to_torch_dtype(dtype))


def cumprod(x, dim=None, dtype=None, name=None):
    return torch.cumprod(x, dim=dim, dtype=_dt.to_torch_dtype(dtype))


def count_nonzero(x, axis=None, keepdim=False, name=None):
    out = torch.count_nonzero(x, dim=_ax(axis))
    return out.unsqueeze(axis) if keepdim and axis is not None and not isinstance(axis, (list, tuple)) else out


def numel(x, name=None):
    return torch.tensor(x.numel(), dtype=torch.int64)


def trace(x, offset=0, axis1=0, axis2=1, name=None):
    return torch.diagonal(x, offset, axis1, axis2).sum(-1)


# ---- linalg
def matmul(x, y, transpose_x=False, transpose_y=False, name=None):
    if transpose_x:
        x = x.transpose(-1, -2) if x.dim() > 1 else x
    if transpose_y:
        y = y.transpose(-1, -2) if y.dim() > 1 else y
    return torch.matmul(x, y)


def mm(input, mat2, name=None):
    return torch.matmul(input, mat2)


def bmm(x, y, name=None):
    return torch.bmm(x, y)


def dot(x, y, name=None):
    return (x * y).sum(-1)


def outer(x, y, name=None):
    return torch.outer(x.flatten(), y.flatten())


def t(input, name=None):
    return input.t() if input.dim() == 2 else input


def norm(x, p=None, axis=None, keepdim=False, name=None):
    if p is None or p == "fro":
        p = 2 if (axis is None or isinstance(axis, int)) else "fro"
    if axis is None:
        return torch.linalg.vector_norm(x.flatten(), ord=p if p != "fro" else 2, keepdim=False)
    if isinstance(axis, (list, tuple)) and len(axis) == 2 and p == "fro":
        return torch.linalg.matrix_norm(x, ord="fro", dim=tuple(axis), keepdim=keepdim)
    if p == float("inf") or p == "inf":
        p = _math.inf
    return torch.linalg.vector_norm(x, ord=p, dim=_ax(axis), keepdim=keepdim)


def einsum(equation, *operands):
    if len(operands) == 1 and isinstance(operands[0], (list, tuple)):
        operands = operands[0]
    return torch.einsum(equation, *operands)


def cross(x, y, axis=9, name=None):
    if axis == 9:
        axis = next(i for i, s in enumerate(x.shape) if s == 3)
    return torch.cross(x, y, dim=axis)


def inverse(x, name=None):
    return torch.linalg.inv(x)


def kron(x, y, name=None):
    return torch.kron(x, y)


def multiplex(inputs, index, name=None):
    stacked = torch.stack(inputs, 0)
    idx = index.flatten().long()
    return stacked[idx, torch.arange(idx.numel(), device=idx.device)]


# ---- comparison / logic
def equal(x, y, name=None):
    return torch.eq(x, _t(y, x))


def not_equal(x, y, name=None):
    return torch.ne(x, _t(y, x))


def greater_than(x, y, name=None):
    return torch.gt(x, _t(y, x))


def greater_equal(x, y, name=None):
    return torch.ge(x, _t(y, x))


def less_than(x, y, name=None):
    return torch.lt(x, _t(y, x))


def less_equal(x, y, name=None):
    return torch.le(x, _t(y, x))


def equal_all(x, y, name=None):
    return torch.tensor(x.shape == y.shape and bool(torch.equal(x, y)))


def allclose(x, y, rtol=1e-05, atol=1e-08, equal_nan=False, name=None):
    return torch.tensor(torch.allclose(x, y, rtol=rtol, atol=atol, equal_nan=equal_nan))


def isclose(x, y, rtol=1e-05, atol=1e-08, equal_nan=False, name=None):
    return torch.isclose(x, y, rtol=rtol, atol=atol, equal_nan=equal_nan)


def logical_and(x, y, out=None, name=None):
    return torch.logical_and(x, y)


def logical_or(x, y, out=None, name=None):
    return torch.logical_or(x, y)


def logical_xor(x, y, out=None, name=None):
    return torch.logical_xor(x, y)


def bitwise_and(x, y, out=None, name=None):
    return torch.bitwise_and(x, y)


def bitwise_or(x, y, out=None, name=None):
    return torch.bitwise_or(x, y)


def bitwise_xor(x, y, out=None, name=None):
    return torch.bitwise_xor(x, y)


def is_tensor(x):
    return isinstance(x, torch.Tensor)


def is_empty(x, name=None):
    return torch.tensor(x.numel() == 0)


# ---- search / sort
def argmax(x, axis=None, keepdim=False, dtype="int64", name=None):
    out = torch.argmax(x, dim=axis, keepdim=keepdim)
    return out.to(_dt.to_torch_dtype(dtype))


def argmin(x, axis=None, keepdim=False, dtype="int64", name=None):
    out = torch.argmin(x, dim=axis, keepdim=keepdim)
    return out.to(_dt.to_torch_dtype(dtype))


def argsort(x, axis=-1, descending=False, stable=False, name=None):
    return torch.argsort(x, dim=axis, descending=descending, stable=stable)


def sort(x, axis=-1, descending=False, stable=False, name=None):
    return torch.sort(x, dim=axis, descending=descending, stable=stable).values


def topk(x, k, axis=None, largest=True, sorted=True, name=None):
    if isinstance(k, torch.Tensor):
        k = int(k.item())
    v, i = torch.topk(x, k, dim=-1 if axis is None else axis, largest=largest, sorted=sorted)
    return v, i


def where(condition, x=None, y=None, name=None):
    if x is None and y is None:
        return tuple(i for i in torch.nonzero(condition, as_tuple=True))
    if not isinstance(x, torch.Tensor):
        x = torch.as_tensor(x, dtype=y.dtype if isinstance(y, torch.Tensor) else None, device=condition.device)
    if not isinstance(y, torch.Tensor):
        y = torch.as_tensor(y, dtype=x.dtype, device=condition.device)
    return torch.where(condition, x, y)


def nonzero(x, as_tuple=False):
    return torch.nonzero(x, as_tuple=as_tuple)


def masked_select(x, mask, name=None):
    return torch.masked_select(x, mask)


def masked_fill(x, mask, value, name=None):
    return torch.masked_fill(x, mask, value.item() if isinstance(value, torch.Tensor) and value.numel() == 1 else value)


def index_select(x, index, axis=0, name=None):
    return torch.index_select(x, axis, index.long() if index.dtype != torch.int64 and index.dtype != torch.int32 else index)


def index_sample(x, index):
    return torch.gather(x, 1, index.long())


def unique(x, return_index=False, return_inverse=False, return_counts=False, axis=None, dtype="int64", name=None):
    res = torch.unique(x, return_inverse=return_inverse or return_index, return_counts=return_counts, dim=axis)
    if not (return_index or return_inverse or return_counts):
        return res
    res = list(res) if isinstance(res, tuple) else [res]
    out = [res[0]]
    inv = res[1] if (return_inverse or return_index) else None
    if return_index:
        n = inv.numel()
        first = torch.full((res[0].shape[0] if axis is not None else res[0].numel(),), n, dtype=torch.int64, device=x.device)
        first.scatter_reduce_(0, inv.flatten(), torch.arange(n, device=x.device), reduce="amin")
        out.append(first)
    if return_inverse:
        out.append(inv)
    if return_counts:
        out.append(res[-1])
    return tuple(out)


def bincount(x, weights=None, minlength=0, name=None):
    return torch.bincount(x, weights=weights, minlength=minlength)


def histogram(input, bins=100, min=0, max=0, name=None):
    return torch.histc(input.float(), bins=bins, min=min, max=max).to(torch.int64)


def searchsorted(sorted_sequence, values, out_int32=False, right=False, name=None):
    return torch.searchsorted(sorted_sequence, values, out_int32=out_int32, right=right)


def kthvalue(x, k, axis=-1, keepdim=False, name=None):
    r = torch.kthvalue(x, k, dim=axis, keepdim=keepdim)
    return r.values, r.indices


def mode(x, axis=-1, keepdim=False, name=None):
    r = torch.mode(x, dim=axis, keepdim=keepdim)
    return r.values, r.indices


def increment(x, value=1.0, name=None):
    with torch.no_grad():
        x.add_(value)
    return x


def top_p_sampling(x, ps, threshold=None, seed=None, name=None):
    """Nucleus sampling (reference: python/paddle/tensor/search.py top_p_sampling, kernel
    paddle/phi/kernels/gpu/top_p_sampling_kernel.cu).  ``x`` [batch, vocab] probabilities, ``ps`` [batch]: per row, keep
    the most probable tokens until their cumulative probability reaches ``ps`` (the token that crosses it included, the
    top token always), drop tokens below ``threshold`` [batch] if given, renormalise and draw one.
    Returns (probability of the drawn token [batch, 1], its id [batch, 1] int64)."""
    if x.dim() != 2:
        raise ValueError("top_p_sampling expects x of shape [batch, vocab]")
    probs = x.float()
    sorted_p, order = probs.sort(-1, descending=True)
    cum = sorted_p.cumsum(-1)
    p = ps.reshape(-1, 1).float().to(probs.device)
    keep = (cum - sorted_p) < p                       # mass BEFORE the token is still below p
    keep[:, 0] = True
    if threshold is not None:
        keep = keep & ((sorted_p >= threshold.reshape(-1, 1).float().to(probs.device)) | (torch.arange(probs.shape[1], device=probs.device) == 0))
    kept = torch.where(keep, sorted_p, torch.zeros_like(sorted_p))
    gen = None
    if seed is not None and int(seed) >= 0:
        gen = torch.Generator(device=probs.device).manual_seed(int(seed))
    pick = torch.multinomial(kept / kept.sum(-1, keepdim=True), 1, generator=gen)
    ids = order.gather(1, pick)
    return x.gather(1, ids), ids
