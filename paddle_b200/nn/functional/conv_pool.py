"""Convolution, pooling, resampling, padding and normalisation front ends with the reference's argument algebra
(reference: python/paddle/nn/functional/{conv,pooling,common,norm,vision}.py; padding rules
paddle/phi/kernels/cpu/conv_util.h:23-72 UpdatePaddingAndDilation; pooling output sizes
paddle/phi/kernels/funcs/pooling.h PoolOutputSize).

What this layer does, the math itself being torch's (cuDNN) kernels - none of it is on the LLM hot path:

* ``data_format``: every op takes the channel-last spellings (NLC / NHWC / NDHWC).  The tensor is viewed channel-first
  (a stride permutation, no copy) for the kernel and viewed back.
* ``padding``: "SAME" / "VALID", one int, one int per spatial dim, (before, after) per spatial dim as a flat list, or
  one pair per tensor dim with zero pairs for batch and channel.  "SAME" gives ``ceil(in / stride)`` outputs, puts the
  odd pixel after, and - as the reference does - resets dilation to 1.  Asymmetric pads are applied as an explicit
  pad in front of the kernel (zeros for conv / average, -inf for max).
* transposed convolutions take ``output_size`` (turned into the output padding it implies) and asymmetric pads (the
  full result is cropped).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as TF

_CONV = {1: TF.conv1d, 2: TF.conv2d, 3: TF.conv3d}
_CONVT = {1: TF.conv_transpose1d, 2: TF.conv_transpose2d, 3: TF.conv_transpose3d}
_MAXP = {1: TF.max_pool1d, 2: TF.max_pool2d, 3: TF.max_pool3d}
_AVGP = {1: TF.avg_pool1d, 2: TF.avg_pool2d, 3: TF.avg_pool3d}
_FORMATS = {1: ("NCL", "NLC"), 2: ("NCHW", "NHWC"), 3: ("NCDHW", "NDHWC")}


# ------------------------------------------------------------------------------------------------------------ helpers
def _ntuple(v, n, name="value"):
    if isinstance(v, torch.Tensor):
        v = v.tolist()
    if isinstance(v, (list, tuple)):
        if len(v) == 1:
            v = list(v) * n
        if len(v) != n:
            raise ValueError(f"{name} should have {n} elements, got {list(v)}")
        return tuple(int(e) for e in v)
    return (int(v),) * n


def _channel_last(data_format, nd):
    first, last = _FORMATS[nd]
    if data_format == first or (nd == 1 and data_format == "NCW"):
        return False
    if data_format == last or (nd == 1 and data_format == "NWC"):
        return True
    raise ValueError(f"data_format should be '{first}' or '{last}', got {data_format!r}")


def _to_cf(x):
    """channel-last -> channel-first view."""
    return x.permute(0, x.dim() - 1, *range(1, x.dim() - 1))


def _to_cl(x):
    return x.permute(0, *range(2, x.dim()), 1)


def _is_seq(v):
    return isinstance(v, (list, tuple))


def _resolve_padding(padding, nd, channel_last):
    """-> (algorithm, [(before, after)] * nd)."""
    if isinstance(padding, str):
        algo = padding.upper()
        if algo not in ("SAME", "VALID"):
            raise ValueError(f"Unknown padding: '{padding}'. It can only be 'SAME' or 'VALID'.")
        return algo, [(0, 0)] * nd
    if _is_seq(padding):
        padding = list(padding)
        if len(padding) == nd + 2 and _is_seq(padding[0]):
            pairs = [tuple(int(v) for v in p) for p in padding]
            outer = [pairs[0], pairs[-1] if channel_last else pairs[1]]
            if any(p != (0, 0) for p in outer):
                raise ValueError(f"Non-zero padding({padding}) in the batch or channel dimensions is not supported.")
            pads = pairs[1:-1] if channel_last else pairs[2:]
        elif len(padding) == 2 * nd and not _is_seq(padding[0]):
            pads = [(int(padding[2 * i]), int(padding[2 * i + 1])) for i in range(nd)]
        elif len(padding) == nd and not _is_seq(padding[0]):
            pads = [(int(p), int(p)) for p in padding]
        elif len(padding) == 1 and not _is_seq(padding[0]):
            pads = [(int(padding[0]), int(padding[0]))] * nd
        else:
            raise ValueError(f"In valid padding: {padding}")
    else:
        pads = [(int(padding), int(padding))] * nd
    if any(b < 0 or a < 0 for b, a in pads):
        raise ValueError(f"Invalid padding, all value should be larger than or equal to 0, but received: {padding}")
    return "EXPLICIT", pads


def _same_pads(sizes, ksize, stride):
    pads = []
    for n, k, s in zip(sizes, ksize, stride):
        out = (n + s - 1) // s
        total = max((out - 1) * s + k - n, 0)
        pads.append((total // 2, total - total // 2))
    return pads


def _flat_pad(pads):
    """[(b, a)] per spatial dim, first dim first -> torch's last-dim-first flat list."""
    flat = []
    for b, a in reversed(pads):
        flat += [b, a]
    return flat


def _symmetric(pads):
    return all(b == a for b, a in pads)


# --------------------------------------------------------------------------------------------------------------- conv
def _conv(nd, x, weight, bias, stride, padding, dilation, groups, data_format):
    cl = _channel_last(data_format, nd)
    if x.dim() != nd + 2:
        raise ValueError(f"conv{nd}d expects a {nd + 2}-D input, got shape {tuple(x.shape)}")
    if cl:
        x = _to_cf(x)
    if x.shape[1] % groups != 0 or x.shape[1] // groups != weight.shape[1]:
        raise ValueError(f"the channel of input must be divisible by groups and match the filter: input channels "
                         f"{x.shape[1]}, groups {groups}, filter shape {tuple(weight.shape)}")
    stride, dilation = _ntuple(stride, nd, "stride"), _ntuple(dilation, nd, "dilation")
    algo, pads = _resolve_padding(padding, nd, cl)
    if algo == "SAME":
        dilation = (1,) * nd
        pads = _same_pads(x.shape[2:], weight.shape[2:], stride)
    if _symmetric(pads):
        pad_arg = tuple(b for b, _ in pads)
    else:
        x = TF.pad(x, _flat_pad(pads))
        pad_arg = 0
    y = _CONV[nd](x, weight, bias, stride, pad_arg, dilation, groups)
    return _to_cl(y) if cl else y


def conv1d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, data_format="NCL", name=None):
    return _conv(1, x, weight, bias, stride, padding, dilation, groups, data_format)


def conv2d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, data_format="NCHW", name=None):
    return _conv(2, x, weight, bias, stride, padding, dilation, groups, data_format)


def conv3d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, data_format="NCDHW", name=None):
    return _conv(3, x, weight, bias, stride, padding, dilation, groups, data_format)


def _conv_transpose(nd, x, weight, bias, stride, padding, output_padding, dilation, groups, output_size, data_format):
    """out = (in - 1) * stride - pad_before - pad_after + dilation * (k - 1) + 1 + output_padding; ``output_size``
    chooses among the ``stride`` sizes that map back to the same input (conv.py conv2d_transpose)."""
    cl = _channel_last(data_format, nd)
    if x.dim() != nd + 2:
        raise ValueError(f"conv{nd}d_transpose expects a {nd + 2}-D input, got shape {tuple(x.shape)}")
    if cl:
        x = _to_cf(x)
    if x.shape[1] != weight.shape[0]:
        raise ValueError(f"input channels {x.shape[1]} do not match the filter's {weight.shape[0]}")
    stride, dilation = _ntuple(stride, nd, "stride"), _ntuple(dilation, nd, "dilation")
    algo, pads = _resolve_padding(padding, nd, cl)
    k = tuple(weight.shape[2:])
    if algo == "SAME":
        dilation = (1,) * nd
        pads = _same_pads(x.shape[2:], k, stride)
    base = [(n - 1) * s - b - a + d * (kk - 1) + 1 for n, s, (b, a), d, kk in zip(x.shape[2:], stride, pads, dilation, k)]
    if output_size is not None:
        if isinstance(output_size, torch.Tensor):
            output_size = output_size.tolist()
        if _is_seq(output_size):
            output_size = [int(v) for v in output_size]
            if len(output_size) == nd + 2:
                output_size = output_size[1:-1] if cl else output_size[2:]
        output_size = _ntuple(output_size, nd, "output_size")
        if output_padding not in (0, [0] * nd, (0,) * nd):
            raise ValueError("output_padding option is mutually exclusive with output_size")
        opad = tuple(o - b for o, b in zip(output_size, base))
        for o, s, d, want, b in zip(opad, stride, dilation, output_size, base):
            if o < 0 or o >= max(s, d):
                raise ValueError(f"output_size {want} is not reachable: valid sizes are [{b}, {b + max(s, d)})")
    else:
        opad = _ntuple(output_padding, nd, "output_padding")
    if _symmetric(pads):
        y = _CONVT[nd](x, weight, bias, stride, tuple(b for b, _ in pads), opad, groups, dilation)
    else:
        y = _CONVT[nd](x, weight, bias, stride, 0, opad, groups, dilation)
        index = [slice(None), slice(None)]
        for (b, a), length, o in zip(pads, base, opad):
            index.append(slice(b, b + length + o))
        y = y[tuple(index)]
    return _to_cl(y) if cl else y


def conv1d_transpose(x, weight, bias=None, stride=1, padding=0, output_padding=0, groups=1, dilation=1,
                     output_size=None, data_format="NCL", name=None):
    return _conv_transpose(1, x, weight, bias, stride, padding, output_padding, dilation, groups, output_size, data_format)


def conv2d_transpose(x, weight, bias=None, stride=1, padding=0, output_padding=0, dilation=1, groups=1,
                     output_size=None, data_format="NCHW", name=None):
    return _conv_transpose(2, x, weight, bias, stride, padding, output_padding, dilation, groups, output_size, data_format)


def conv3d_transpose(x, weight, bias=None, stride=1, padding=0, output_padding=0, groups=1, dilation=1,
                     output_size=None, data_format="NCDHW", name=None):
    return _conv_transpose(3, x, weight, bias, stride, padding, output_padding, dilation, groups, output_size, data_format)


# ------------------------------------------------------------------------------------------------------------ pooling
def _pool_setup(nd, x, kernel_size, stride, padding, data_format):
    cl = _channel_last(data_format, nd)
    if x.dim() != nd + 2:
        raise ValueError(f"{nd}-D pooling expects a {nd + 2}-D input, got shape {tuple(x.shape)}")
    if cl:
        x = _to_cf(x)
    k = _ntuple(kernel_size, nd, "kernel_size")
    s = k if stride is None else _ntuple(stride, nd, "stride")
    algo, pads = _resolve_padding(padding, nd, cl)
    if algo == "SAME":
        pads = _same_pads(x.shape[2:], k, s)
    return cl, x, k, s, pads, algo


def _max_pool(nd, x, kernel_size, stride, padding, return_mask, ceil_mode, data_format):
    cl, x, k, s, pads, algo = _pool_setup(nd, x, kernel_size, stride, padding, data_format)
    if algo == "SAME":
        ceil_mode = False
    if _symmetric(pads):
        out = _MAXP[nd](x, k, s, tuple(b for b, _ in pads), 1, ceil_mode, return_mask)
        y, mask = out if return_mask else (out, None)
    else:
        xp = TF.pad(x, _flat_pad(pads), value=float("-inf") if x.dtype.is_floating_point else torch.iinfo(x.dtype).min)
        out = _MAXP[nd](xp, k, s, 0, 1, ceil_mode, return_mask)
        y, mask = out if return_mask else (out, None)
        if mask is not None:                      # flat index in the padded map -> flat index in the input map
            padded = xp.shape[2:]
            coords, rem = [], mask
            for n in reversed(padded):
                coords.append(rem % n)
                rem = rem // n
            coords = coords[::-1]
            flat = torch.zeros_like(mask)
            for c, (b, _), n in zip(coords, pads, x.shape[2:]):
                flat = flat * n + (c - b)
            mask = flat
    if cl:
        y = _to_cl(y)
        mask = _to_cl(mask) if mask is not None else None
    return (y, mask) if return_mask else y


def _avg_pool(nd, x, kernel_size, stride, padding, ceil_mode, exclusive, divisor_override, data_format):
    cl, x, k, s, pads, algo = _pool_setup(nd, x, kernel_size, stride, padding, data_format)
    if algo == "SAME":
        ceil_mode = False
    extra = {} if nd == 1 else {"divisor_override": divisor_override}
    if nd == 1 and divisor_override is not None:
        raise ValueError("avg_pool1d has no divisor_override")
    if _symmetric(pads):
        y = _AVGP[nd](x, k, s, tuple(b for b, _ in pads), ceil_mode, not exclusive, **extra)
    else:
        xp = TF.pad(x, _flat_pad(pads))
        if exclusive and divisor_override is None:
            # sum over the window / number of real (un-padded) elements in it
            area = float(math.prod(k))
            total = _AVGP[nd](xp, k, s, 0, ceil_mode, True) * area
            ones = TF.pad(torch.ones((1, 1) + tuple(x.shape[2:]), dtype=x.dtype, device=x.device), _flat_pad(pads))
            count = _AVGP[nd](ones, k, s, 0, ceil_mode, True) * area
            y = total / count.clamp(min=1.0)
        else:
            y = _AVGP[nd](xp, k, s, 0, ceil_mode, True, **extra)
    return _to_cl(y) if cl else y


def max_pool1d(x, kernel_size, stride=None, padding=0, return_mask=False, ceil_mode=False, name=None):
    return _max_pool(1, x, kernel_size, stride, padding, return_mask, ceil_mode, "NCL")


def max_pool2d(x, kernel_size, stride=None, padding=0, return_mask=False, ceil_mode=False, data_format="NCHW", name=None):
    if return_mask and data_format != "NCHW":
        raise ValueError("When setting return_mask to true, data_format must be set to NCHW in API:max_pool2d")
    return _max_pool(2, x, kernel_size, stride, padding, return_mask, ceil_mode, data_format)


def max_pool3d(x, kernel_size, stride=None, padding=0, return_mask=False, ceil_mode=False, data_format="NCDHW", name=None):
    if return_mask and data_format != "NCDHW":
        raise ValueError("When setting return_mask to true, data_format must be set to NCDHW in API:max_pool3d")
    return _max_pool(3, x, kernel_size, stride, padding, return_mask, ceil_mode, data_format)


def avg_pool1d(x, kernel_size, stride=None, padding=0, exclusive=True, ceil_mode=False, name=None):
    return _avg_pool(1, x, kernel_size, stride, padding, ceil_mode, exclusive, None, "NCL")


def avg_pool2d(x, kernel_size, stride=None, padding=0, ceil_mode=False, exclusive=True, divisor_override=None,
               data_format="NCHW", name=None):
    return _avg_pool(2, x, kernel_size, stride, padding, ceil_mode, exclusive, divisor_override, data_format)


def avg_pool3d(x, kernel_size, stride=None, padding=0, ceil_mode=False, exclusive=True, divisor_override=None,
               data_format="NCDHW", name=None):
    return _avg_pool(3, x, kernel_size, stride, padding, ceil_mode, exclusive, divisor_override, data_format)


def _lp_pool(nd, x, norm_type, kernel_size, stride, padding, ceil_mode, data_format):
    """(sum over the window of x ** p) ** (1 / p); p = inf is max pooling (pooling.py lp_pool2d)."""
    p = float(norm_type)
    if p == 0:
        raise ValueError("norm_type should not be 0")
    if math.isinf(p):
        return _max_pool(nd, x, kernel_size, stride, padding, False, ceil_mode, data_format)
    k = _ntuple(kernel_size, nd, "kernel_size")
    avg = _avg_pool(nd, x.pow(p), kernel_size, stride, padding, ceil_mode, False, None, data_format)
    total = avg * float(math.prod(k))
    return (torch.sign(total) * torch.relu(torch.abs(total))).pow(1.0 / p)


def lp_pool1d(x, norm_type, kernel_size, stride=None, padding=0, ceil_mode=False, data_format="NCL", name=None):
    return _lp_pool(1, x, norm_type, kernel_size, stride, padding, ceil_mode, data_format)


def lp_pool2d(x, norm_type, kernel_size, stride=None, padding=0, ceil_mode=False, data_format="NCHW", name=None):
    return _lp_pool(2, x, norm_type, kernel_size, stride, padding, ceil_mode, data_format)


def _adaptive_size(output_size, in_sizes, nd):
    if isinstance(output_size, int):
        return (output_size,) * nd
    out = list(output_size)
    if len(out) != nd:
        raise ValueError(f"output_size should have {nd} elements, got {out}")
    return tuple(int(n) if o is None else int(o) for o, n in zip(out, in_sizes))


_AAVG = {1: TF.adaptive_avg_pool1d, 2: TF.adaptive_avg_pool2d, 3: TF.adaptive_avg_pool3d}
_AMAX = {1: TF.adaptive_max_pool1d, 2: TF.adaptive_max_pool2d, 3: TF.adaptive_max_pool3d}


def _adaptive(nd, table, x, output_size, data_format, return_mask=None):
    cl = _channel_last(data_format, nd)
    if cl:
        x = _to_cf(x)
    size = _adaptive_size(output_size, x.shape[2:], nd)
    if return_mask is None:
        y = table[nd](x, size)
        return _to_cl(y) if cl else y
    out = table[nd](x, size, return_indices=return_mask)
    return out


def adaptive_avg_pool1d(x, output_size, name=None):
    return _adaptive(1, _AAVG, x, output_size, "NCL")


def adaptive_avg_pool2d(x, output_size, data_format="NCHW", name=None):
    return _adaptive(2, _AAVG, x, output_size, data_format)


def adaptive_avg_pool3d(x, output_size, data_format="NCDHW", name=None):
    return _adaptive(3, _AAVG, x, output_size, data_format)


def adaptive_max_pool1d(x, output_size, return_mask=False, name=None):
    return _adaptive(1, _AMAX, x, output_size, "NCL", return_mask)


def adaptive_max_pool2d(x, output_size, return_mask=False, name=None):
    return _adaptive(2, _AMAX, x, output_size, "NCHW", return_mask)


def adaptive_max_pool3d(x, output_size, return_mask=False, name=None):
    return _adaptive(3, _AMAX, x, output_size, "NCDHW", return_mask)


def _max_unpool(nd, x, indices, kernel_size, stride, padding, data_format, output_size):
    """Scatter every value to the flat position its index names; default output size
    (in - 1) * stride - 2 * padding + kernel (pooling.py max_unpool2d)."""
    if _channel_last(data_format, nd):
        raise ValueError(f"max_unpool{nd}d only supports the channel-first data_format")
    k = _ntuple(kernel_size, nd, "kernel_size")
    s = k if stride is None else _ntuple(stride, nd, "stride")
    p = _ntuple(padding, nd, "padding")
    default = [(n - 1) * ss - 2 * pp + kk for n, ss, pp, kk in zip(x.shape[2:], s, p, k)]
    if output_size is None:
        out = default
    else:
        out = [int(v) for v in output_size]
        if len(out) == nd + 2:
            out = out[2:]
        if len(out) != nd:
            raise ValueError(f"output_size should be a sequence containing {nd} or {nd + 2} elements")
        for o, d, ss in zip(out, default, s):
            if not d - ss < o < d + ss:
                raise ValueError(f"invalid output_size {output_size}: each size must be within a stride of {default}")
    n, c = x.shape[:2]
    flat = torch.zeros((n, c, math.prod(out)), dtype=x.dtype, device=x.device)
    flat = flat.scatter(2, indices.reshape(n, c, -1).long(), x.reshape(n, c, -1))
    return flat.reshape((n, c) + tuple(out))


def max_unpool1d(x, indices, kernel_size, stride=None, padding=0, data_format="NCL", output_size=None, name=None):
    return _max_unpool(1, x, indices, kernel_size, stride, padding, data_format, output_size)


def max_unpool2d(x, indices, kernel_size, stride=None, padding=0, data_format="NCHW", output_size=None, name=None):
    return _max_unpool(2, x, indices, kernel_size, stride, padding, data_format, output_size)


def max_unpool3d(x, indices, kernel_size, stride=None, padding=0, data_format="NCDHW", output_size=None, name=None):
    return _max_unpool(3, x, indices, kernel_size, stride, padding, data_format, output_size)


# -------------------------------------------------------------------------------------------------------- interpolate
_INTERP_ND = {"linear": (1,), "bilinear": (2,), "trilinear": (3,), "bicubic": (2,), "nearest": (1, 2, 3), "area": (1, 2, 3)}


def _linear_along(x, dim, out_size, scale):
    """Linear resampling of one dim with source coordinate ``dst * scale`` (the reference's align_mode=1)."""
    n = x.shape[dim]
    src = torch.arange(out_size, device=x.device, dtype=torch.float64) * scale
    i0 = src.floor().clamp(0, n - 1).long()
    i1 = (i0 + 1).clamp(max=n - 1)
    lam = (src - i0.to(src.dtype)).clamp(0, 1).to(x.dtype)
    shape = [1] * x.dim()
    shape[dim] = out_size
    lam = lam.reshape(shape)
    return x.index_select(dim, i0) * (1 - lam) + x.index_select(dim, i1) * lam


def interpolate(x, size=None, scale_factor=None, mode="nearest", align_corners=False, align_mode=0,
                data_format=None, recompute_scale_factor=None, name=None):
    """Reference: python/paddle/nn/functional/common.py interpolate.  ``align_mode`` matters for (bi/tri)linear with
    ``align_corners=False``: 0 samples at half-pixel centres ((dst + 0.5) * scale - 0.5), 1 at ``dst * scale``."""
    mode = mode.lower()
    if mode not in _INTERP_ND:
        raise ValueError("The 'resample' of image_resize can only be 'area', 'linear', 'bilinear', 'trilinear', "
                         "'bicubic' or 'nearest' currently.")
    nd = x.dim() - 2
    if nd not in (1, 2, 3):
        raise ValueError(f"interpolate expects a 3-D, 4-D or 5-D input, got {x.dim()}-D")
    if nd not in _INTERP_ND[mode]:
        raise ValueError(f"'{mode}' only supports {[n + 2 for n in _INTERP_ND[mode]]}-D input, got {x.dim()}-D")
    if data_format is None:
        data_format = _FORMATS[nd][0]
    cl = _channel_last(data_format, nd)
    if size is None and scale_factor is None:
        raise ValueError("One of size and scale_factor must not be None.")
    if size is not None and scale_factor is not None:
        raise ValueError("Only one of size or scale_factor should be defined.")
    if align_corners and mode in ("nearest", "area"):
        raise ValueError("align_corners option can only be set with the interpolating modes: linear | bilinear | "
                         "bicubic | trilinear")
    if align_mode not in (0, 1):
        raise ValueError("align_mode can only be 0 or 1")
    if cl:
        x = _to_cf(x)
    in_sizes = tuple(x.shape[2:])
    if size is not None:
        if isinstance(size, torch.Tensor):
            size = size.tolist()
        if not _is_seq(size):
            size = [size] * nd
        out_sizes = tuple(int(v.item()) if isinstance(v, torch.Tensor) else int(v) for v in size)
        if len(out_sizes) != nd:
            raise ValueError(f"size should have {nd} elements for a {x.dim()}-D input, got {list(size)}")
        scales = None
    else:
        if isinstance(scale_factor, torch.Tensor):
            scale_factor = scale_factor.tolist()
        sf = list(scale_factor) if _is_seq(scale_factor) else [scale_factor] * nd
        if len(sf) != nd:
            raise ValueError(f"scale_factor should have {nd} elements for a {x.dim()}-D input, got {sf}")
        if any(float(s) <= 0 for s in sf):
            raise ValueError("Attr(scale) should be greater than zero.")
        scales = tuple(float(s) for s in sf)
        out_sizes = tuple(int(math.floor(n * s)) for n, s in zip(in_sizes, scales))
    if any(o <= 0 for o in out_sizes):
        raise ValueError(f"output sizes must be positive, got {out_sizes}")
    if mode == "area":
        y = _AAVG[nd](x, out_sizes)
    elif mode == "nearest":
        y = None
        if scales is not None:
            y = TF.interpolate(x, scale_factor=list(scales), mode="nearest", recompute_scale_factor=False)
        if y is None or tuple(y.shape[2:]) != out_sizes:
            y = TF.interpolate(x, size=out_sizes, mode="nearest")
    elif mode in ("linear", "bilinear", "trilinear") and not align_corners and align_mode == 1:
        y = x
        for d, (n, o) in enumerate(zip(in_sizes, out_sizes)):
            ratio = (1.0 / scales[d]) if scales is not None else n / o
            y = _linear_along(y, d + 2, o, ratio)
    else:
        kw = {}
        if scales is not None and not recompute_scale_factor and not align_corners:
            # the given scale (not out / in) drives the source coordinates, as in the reference's kernels
            y = TF.interpolate(x, scale_factor=list(scales), mode=mode, align_corners=align_corners,
                               recompute_scale_factor=False)
            if tuple(y.shape[2:]) != out_sizes:
                y = TF.interpolate(x, size=out_sizes, mode=mode, align_corners=align_corners)
        else:
            y = TF.interpolate(x, size=out_sizes, mode=mode, align_corners=align_corners, **kw)
    return _to_cl(y) if cl else y


def upsample(x, size=None, scale_factor=None, mode="nearest", align_corners=False, align_mode=0, data_format=None,
             name=None):
    return interpolate(x, size, scale_factor, mode, align_corners, align_mode, data_format)


# ---------------------------------------------------------------------------------------------------------------- pad
_PAD_MODES = {"constant": "constant", "reflect": "reflect", "replicate": "replicate", "circular": "circular"}


def pad(x, pad, mode="constant", value=0.0, data_format="NCHW", pad_from_left_axis=True, name=None):
    """Reference: python/paddle/nn/functional/common.py pad.  Two forms:

    * ``len(pad) == 2 * x.ndim`` with ``mode="constant"``: (before, after) for EVERY dim - from the first dim when
      ``pad_from_left_axis`` (default), from the last dim otherwise;
    * otherwise ``pad`` covers the spatial dims of a 3-D / 4-D / 5-D tensor in ``data_format``, innermost dim first:
      (left, right[, top, bottom[, front, back]]).
    """
    if isinstance(pad, torch.Tensor):
        pad = pad.tolist()
    pad = [int(p.item()) if isinstance(p, torch.Tensor) else int(p) for p in pad]
    if mode not in _PAD_MODES:
        raise ValueError(f"mode should be one of constant, reflect, replicate, circular, but got {mode}.")
    nd_total = x.dim()
    if mode == "constant" and len(pad) == 2 * nd_total:
        if pad_from_left_axis:
            flat = []
            for i in range(nd_total - 1, -1, -1):
                flat += [pad[2 * i], pad[2 * i + 1]]
        else:
            flat = pad
        return TF.pad(x, flat, mode="constant", value=value)
    nd = nd_total - 2
    if nd not in (1, 2, 3):
        raise ValueError(f"pad with {len(pad)} values expects a 3-D, 4-D or 5-D input, got {nd_total}-D")
    if len(pad) != 2 * nd:
        raise ValueError(f"pad should have {2 * nd} values for a {nd_total}-D input in mode '{mode}', got {len(pad)}")
    supported = {1: ("NCL", "NLC"), 2: ("NCHW", "NHWC"), 3: ("NCDHW", "NDHWC")}[nd]
    if data_format not in supported:
        if data_format == "NCHW":                 # the signature default, whatever the rank
            data_format = supported[0]
        else:
            raise ValueError(f"data_format should be in {supported} for a {nd_total}-D input, got {data_format}")
    cl = data_format == supported[1]
    if cl:
        x = _to_cf(x)
    if mode == "constant":
        y = TF.pad(x, pad, mode="constant", value=value)
    else:
        y = TF.pad(x, pad, mode=_PAD_MODES[mode])
    return _to_cl(y) if cl else y


def zeropad2d(x, padding, data_format="NCHW", name=None):
    return pad(x, padding, mode="constant", value=0.0, data_format=data_format)


# -------------------------------------------------------------------------------------------------------------- norms
def _norm_channel_last(x, data_format):
    """Normalisation layers take N...C spellings for any rank (NC, NLC, NHWC, NDHWC)."""
    if data_format in ("NC", "NCL", "NCW", "NCHW", "NCDHW"):
        return False
    if data_format in ("NLC", "NWC", "NHWC", "NDHWC"):
        return x.dim() > 2
    raise ValueError(f"unknown data_format {data_format!r}")


def batch_norm(x, running_mean, running_var, weight=None, bias=None, training=False, momentum=0.9, epsilon=1e-05,
               data_format="NCHW", use_global_stats=None, name=None):
    """Reference: python/paddle/nn/functional/norm.py batch_norm, paddle/phi/kernels/cpu/batch_norm_kernel.cc:100-157.
    The running statistics move as ``running = momentum * running + (1 - momentum) * batch`` and the batch variance
    that goes into them is the BIASED one (sum / N), the same value the normalisation uses."""
    cl = _norm_channel_last(x, data_format)
    if cl:
        x = _to_cf(x)
    use_batch = training and not use_global_stats
    if use_global_stats is None:
        use_batch = bool(training)
    if not use_batch:
        y = TF.batch_norm(x, running_mean, running_var, weight, bias, False, 0.0, epsilon)
        return _to_cl(y) if cl else y
    if running_mean is None or running_var is None:
        y = TF.batch_norm(x, None, None, weight, bias, True, 0.0, epsilon)
        return _to_cl(y) if cl else y
    # one pass: the library kernel updates the running statistics with the UNBIASED variance; the biased update the
    # reference makes follows from it on the C-sized vectors: (new - m old) is (1 - m) var_unbiased, times (n - 1) / n
    n = x.numel() // x.shape[1]
    # the kernel may keep the variance buffer it was given for its backward pass, so it gets a scratch copy and the
    # real buffer is written once, afterwards
    new_var = running_var.detach().clone()
    y = TF.batch_norm(x, running_mean, new_var, weight, bias, True, 1.0 - momentum, epsilon)
    with torch.no_grad():
        if n > 1:
            keep = running_var * momentum
            running_var.copy_(keep + (new_var - keep) * ((n - 1) / n))
        else:
            running_var.copy_(new_var)
    return _to_cl(y) if cl else y


def instance_norm(x, running_mean=None, running_var=None, weight=None, bias=None, use_input_stats=True, momentum=0.9,
                  eps=1e-05, data_format="NCHW", name=None):
    cl = _norm_channel_last(x, data_format)
    if cl:
        x = _to_cf(x)
    y = TF.instance_norm(x, running_mean, running_var, weight, bias, use_input_stats, 1.0 - momentum, eps)
    return _to_cl(y) if cl else y


def group_norm(x, num_groups, epsilon=1e-05, weight=None, bias=None, data_format="NCHW", name=None):
    cl = _norm_channel_last(x, data_format)
    if cl:
        x = _to_cf(x)
    y = TF.group_norm(x, num_groups, weight, bias, epsilon)
    return _to_cl(y) if cl else y


def local_response_norm(x, size, alpha=1e-4, beta=0.75, k=1.0, data_format="NCHW", name=None):
    """x / (k + alpha / size * sum over ``size`` neighbouring channels of x^2) ** beta (norm.py local_response_norm)."""
    cl = _norm_channel_last(x, data_format)
    if cl:
        x = _to_cf(x)
    y = TF.local_response_norm(x, size, alpha, beta, k)
    return _to_cl(y) if cl else y


# ----------------------------------------------------------------------------------------------------------- shuffles
def pixel_shuffle(x, upscale_factor, data_format="NCHW", name=None):
    cl = _channel_last(data_format, 2)
    if x.dim() != 4:
        raise ValueError(f"Input x should be 4D tensor, but received x with the shape of {tuple(x.shape)}")
    y = TF.pixel_shuffle(_to_cf(x) if cl else x, upscale_factor)
    return _to_cl(y) if cl else y


def pixel_unshuffle(x, downscale_factor, data_format="NCHW", name=None):
    cl = _channel_last(data_format, 2)
    if x.dim() != 4:
        raise ValueError(f"Input x should be 4D tensor, but received x with the shape of {tuple(x.shape)}")
    if downscale_factor <= 0:
        raise ValueError("Downscale factor must be positive")
    y = TF.pixel_unshuffle(_to_cf(x) if cl else x, downscale_factor)
    return _to_cl(y) if cl else y


def channel_shuffle(x, groups, data_format="NCHW", name=None):
    cl = _channel_last(data_format, 2)
    if x.dim() != 4:
        raise ValueError(f"Input x should be 4D tensor, but received x with the shape of {tuple(x.shape)}")
    if groups <= 0:
        raise ValueError("groups should be a positive integer")
    y = TF.channel_shuffle(_to_cf(x) if cl else x, groups)
    return _to_cl(y) if cl else y


def unfold(x, kernel_sizes, strides=1, paddings=0, dilations=1, name=None):
    """im2col: [N, C, H, W] -> [N, C * kh * kw, L].  ``paddings``: int, (ph, pw) or (top, left, bottom, right)
    (reference: python/paddle/nn/functional/common.py unfold)."""
    k = _ntuple(kernel_sizes, 2, "kernel_sizes")
    s = _ntuple(strides, 2, "strides")
    d = _ntuple(dilations, 2, "dilations")
    if _is_seq(paddings) and len(paddings) == 4:
        top, left, bottom, right = (int(p) for p in paddings)
        if top == bottom and left == right:
            return TF.unfold(x, k, d, (top, left), s)
        return TF.unfold(TF.pad(x, (left, right, top, bottom)), k, d, 0, s)
    return TF.unfold(x, k, d, _ntuple(paddings, 2, "paddings"), s)


def fold(x, output_sizes, kernel_sizes, strides=1, paddings=0, dilations=1, name=None):
    """col2im, the adjoint of ``unfold`` (common.py fold)."""
    out = _ntuple(output_sizes, 2, "output_sizes")
    k = _ntuple(kernel_sizes, 2, "kernel_sizes")
    s = _ntuple(strides, 2, "strides")
    d = _ntuple(dilations, 2, "dilations")
    if _is_seq(paddings) and len(paddings) == 4:
        top, left, bottom, right = (int(p) for p in paddings)
        if top == bottom and left == right:
            return TF.fold(x, out, k, d, (top, left), s)
        full = TF.fold(x, (out[0] + top + bottom, out[1] + left + right), k, d, 0, s)
        return full[:, :, top:top + out[0], left:left + out[1]]
    return TF.fold(x, out, k, d, _ntuple(paddings, 2, "paddings"), s)
