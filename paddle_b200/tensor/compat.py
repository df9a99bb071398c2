"""Long tail of the top-level ``paddle.*`` tensor API (reference: python/paddle/__init__.py ``__all__``): functions with
a direct torch equivalent, and the in-place ``name_`` variants, which are generated from their out-of-place versions
(``x.copy_(name(x, ...))``) so that they exist both as functions and as Tensor methods."""
from __future__ import annotations

import math as _math

import torch

from ..framework import dtype as _dt


def diagflat(x, offset=0, name=None):
    return torch.diagflat(x, offset)


def diag_embed(input, offset=0, dim1=-2, dim2=-1):
    return torch.diag_embed(input, offset, dim1, dim2)


def logaddexp(x, y, name=None):
    return torch.logaddexp(x, y)


def logit(x, eps=None, name=None):
    return torch.logit(x, eps)


def is_complex(x):
    return x.is_complex()


def is_integer(x):
    return not x.is_floating_point() and not x.is_complex() and x.dtype != torch.bool


def is_floating_point(x):
    return x.is_floating_point()


def mv(x, vec, name=None):
    return torch.mv(x, vec)


def slice_scatter(x, value, axes, starts, ends, strides, name=None):
    out = x.clone()
    idx = [slice(None)] * x.dim()
    for a, s, e, st in zip(axes, starts, ends, strides):
        idx[a] = slice(s, e, st)
    out[tuple(idx)] = value
    return out


def select_scatter(x, values, axis, index, name=None):
    return torch.select_scatter(x, values, axis, index)


def diagonal_scatter(x, y, offset=0, axis1=0, axis2=1, name=None):
    return torch.diagonal_scatter(x, y, offset, axis1, axis2)


def masked_scatter(x, mask, value, name=None):
    return torch.masked_scatter(x, mask.expand_as(x) if mask.shape != x.shape else mask, value)


def index_fill(x, index, axis, value, name=None):
    return torch.index_fill(x, axis, index.long(), value)


def log_normal(mean=1.0, std=2.0, shape=None, name=None):
    return torch.empty(tuple(shape or [1])).log_normal_(mean, std)


def tensor_split(x, num_or_indices, axis=0, name=None):
    return list(torch.tensor_split(x, num_or_indices, dim=axis))


def hsplit(x, num_or_indices, name=None):
    return list(torch.hsplit(x, num_or_indices))


def vsplit(x, num_or_indices, name=None):
    return list(torch.vsplit(x, num_or_indices))


def dsplit(x, num_or_indices, name=None):
    return list(torch.dsplit(x, num_or_indices))


def hstack(x, name=None):
    return torch.hstack(list(x))


def vstack(x, name=None):
    return torch.vstack(list(x))


def dstack(x, name=None):
    return torch.dstack(list(x))


def column_stack(x, name=None):
    return torch.column_stack(list(x))


row_stack = vstack


def atleast_1d(*inputs, name=None):
    r = torch.atleast_1d(*inputs)
    return list(r) if isinstance(r, tuple) else r


def atleast_2d(*inputs, name=None):
    r = torch.atleast_2d(*inputs)
    return list(r) if isinstance(r, tuple) else r


def atleast_3d(*inputs, name=None):
    r = torch.atleast_3d(*inputs)
    return list(r) if isinstance(r, tuple) else r


def isneginf(x, name=None):
    return torch.isneginf(x)


def isposinf(x, name=None):
    return torch.isposinf(x)


def isreal(x, name=None):
    return torch.isreal(x)


def rot90(x, k=1, axes=(0, 1), name=None):
    return torch.rot90(x, k, list(axes))


def reverse(x, axis, name=None):
    return torch.flip(x, [axis] if isinstance(axis, int) else list(axis))


def histogramdd(x, bins=10, ranges=None, density=False, weights=None, name=None):
    r = None if ranges is None else [float(v) for v in ranges]
    h = torch.histogramdd(x, bins=bins, range=r, weight=weights, density=density)
    return h.hist, list(h.bin_edges)


def dist(x, y, p=2, name=None):
    return torch.dist(x, y, p)


def pdist(x, p=2.0, name=None):
    return torch.pdist(x, p)


def nanmedian(x, axis=None, keepdim=False, mode="avg", name=None):
    if axis is None:
        return torch.nanmedian(x) if mode == "min" else torch.nanquantile(x.float(), 0.5).to(x.dtype)
    return torch.nanmedian(x, dim=axis, keepdim=keepdim).values if mode == "min" else \
        torch.nanquantile(x.float(), 0.5, dim=axis, keepdim=keepdim).to(x.dtype)


def binomial(count, prob, name=None):
    return torch.binomial(count.float(), prob.float()).to(torch.int64)


def standard_gamma(x, name=None):
    return torch._standard_gamma(x)


def broadcast_shape(x_shape, y_shape):
    return list(torch.broadcast_shapes(tuple(x_shape), tuple(y_shape)))


def gammaln(x, name=None):
    return torch.lgamma(x)


def gammainc(x, y, name=None):
    return torch.special.gammainc(x, y)


def gammaincc(x, y, name=None):
    return torch.special.gammaincc(x, y)


def multigammaln(x, p, name=None):
    return torch.special.multigammaln(x, p)


def rad2deg(x, name=None):
    return torch.rad2deg(x.float() if not x.is_floating_point() else x)


def deg2rad(x, name=None):
    return torch.deg2rad(x.float() if not x.is_floating_point() else x)


def reduce_as(x, target, name=None):
    """Sum ``x`` over the broadcast dimensions so that it gets the shape of ``target``."""
    return x.sum_to_size(target.shape)


def create_parameter(shape, dtype, name=None, attr=None, is_bias=False, default_initializer=None):
    from ..nn.layer.layers import Layer
    holder = Layer()
    return holder.create_parameter(list(shape), attr=attr, dtype=dtype, is_bias=is_bias, default_initializer=default_initializer)


def unique_consecutive(x, return_inverse=False, return_counts=False, axis=None, dtype="int64", name=None):
    return torch.unique_consecutive(x, return_inverse=return_inverse, return_counts=return_counts, dim=axis)


def logspace(start, stop, num, base=10.0, dtype=None, name=None):
    return torch.logspace(float(start), float(stop), int(num), base=float(base), dtype=_dt.to_torch_dtype(dtype) if dtype else None)


def check_shape(shape, op_name=None, expected_shape_type=(list, tuple, torch.Tensor), expected_element_type=(int, torch.Tensor),
                expected_tensor_dtype=("int32", "int64")):
    if not isinstance(shape, expected_shape_type):
        raise TypeError(f"{op_name}: shape must be one of {expected_shape_type}, got {type(shape)}")


def asinh(x, name=None):
    return torch.asinh(x)


def acosh(x, name=None):
    return torch.acosh(x)


def atanh(x, name=None):
    return torch.atanh(x)


def nan_to_num(x, nan=0.0, posinf=None, neginf=None, name=None):
    return torch.nan_to_num(x, nan, posinf, neginf)


def tril_indices(row, col=None, offset=0, dtype="int64"):
    return torch.tril_indices(row, col if col is not None else row, offset, dtype=_dt.to_torch_dtype(dtype))


def triu_indices(row, col=None, offset=0, dtype="int64"):
    return torch.triu_indices(row, col if col is not None else row, offset, dtype=_dt.to_torch_dtype(dtype))


def sgn(x, name=None):
    return torch.sgn(x)


def take(x, index, mode="raise", name=None):
    flat, n = x.reshape(-1), x.numel()
    idx = index.long()
    if mode == "wrap":
        idx = idx % n
    elif mode == "clip":
        idx = idx.clamp(0, n - 1)
    else:
        idx = torch.where(idx < 0, idx + n, idx)
    return flat[idx]


def frexp(x, name=None):
    m, e = torch.frexp(x)
    return m, e.to(x.dtype)


def cumulative_trapezoid(y, x=None, dx=None, axis=-1, name=None):
    return torch.cumulative_trapezoid(y, x=x, dim=axis) if x is not None else \
        torch.cumulative_trapezoid(y, dx=dx if dx is not None else 1.0, dim=axis)


def polar(abs, angle, name=None):  # noqa: A002
    return torch.polar(abs, angle)


def view_as(x, other, name=None):
    return x.view(other.shape)


def unfold(x, axis, size, step, name=None):
    return x.unfold(axis, size, step)


def i0e(x, name=None):
    return torch.special.i0e(x)


def i1(x, name=None):
    return torch.special.i1(x)


def i1e(x, name=None):
    return torch.special.i1e(x)


def bitwise_left_shift(x, y, is_arithmetic=True, name=None):
    return torch.bitwise_left_shift(x, y)


def bitwise_right_shift(x, y, is_arithmetic=True, name=None):
    return torch.bitwise_right_shift(x, y)


def batch(reader, batch_size, drop_last=False):
    """Legacy reader decorator: groups the samples of ``reader()`` into lists of ``batch_size``."""
    def gen():
        buf = []
        for s in reader():
            buf.append(s)
            if len(buf) == batch_size:
                yield buf
                buf = []
        if buf and not drop_last:
            yield buf
    return gen


def flops(net, input_size, custom_ops=None, print_detail=False):
    """Multiply-accumulate based FLOPs of a Layer for one input of ``input_size`` (Linear / Conv layers counted)."""
    total = [0]
    hooks = []

    def lin(m, i, o):
        total[0] += 2 * i[0].numel() // i[0].shape[-1] * m.weight.numel()

    def conv(m, i, o):
        w = m.weight
        total[0] += 2 * o.numel() * (w.numel() // w.shape[0])
    for m in net.sublayers(include_self=True) if hasattr(net, "sublayers") else net.modules():
        n = type(m).__name__
        if n == "Linear":
            hooks.append(m.register_forward_hook(lin))
        elif n.startswith("Conv"):
            hooks.append(m.register_forward_hook(conv))
    was = net.training
    net.eval()
    with torch.no_grad():
        net(torch.zeros(tuple(input_size)))
    net.train(was)
    for h in hooks:
        h.remove()
    return total[0]


class LazyGuard:
    """Parameters created under the guard are materialised on first use in the reference; here creation is already
    cheap (direct device allocation), so the guard only exists for API compatibility."""

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def _make_inplace(base_fn):
    def fn(x, *args, **kwargs):
        with torch.no_grad():
            res = base_fn(x, *args, **kwargs)
            x.copy_(res.to(x.dtype) if res.dtype != x.dtype else res)
        return x
    fn.__name__ = base_fn.__name__ + "_"
    fn.__doc__ = f"In-place version of ``{base_fn.__name__}`` (writes the result into the first argument)."
    return fn


def cauchy_(x, loc=0, scale=1, name=None):
    return x.cauchy_(loc, scale)


def log_normal_(x, mean=1.0, std=2.0, name=None):
    return x.log_normal_(mean, std)


def install_inplace_variants(namespace, names):
    """Create ``name_`` for every ``name`` in ``names`` that the namespace defines but whose in-place form is missing,
    and attach both forms to torch.Tensor when the method does not exist yet."""
    made = []
    for n in names:
        if not n.endswith("_") or n in namespace:
            continue
        base = namespace.get(n[:-1])
        if base is None or not callable(base):
            continue
        namespace[n] = _make_inplace(base)
        made.append(n)
        if not hasattr(torch.Tensor, n):
            setattr(torch.Tensor, n, namespace[n])
    return made


def geometric_(x, probs, name=None):
    return x.geometric_(probs)
