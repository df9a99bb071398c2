"""paddle.nn.quant (reference: python/paddle/nn/quant/{quantized_linear,stub}.py; kernels
paddle/phi/kernels/impl/weight_quantize_kernel_impl.h): weight-only and LLM.int8 linear layers.

Layouts of this build (the reference's GPU layouts are interleaved per SM architecture and are not portable between
architectures either; ``weight_quantize`` / ``weight_dequantize`` / the linears below agree with each other):
  * int8: ``[n, k]`` int8 - the transpose of the ``[k, n]`` float weight, one symmetric scale per output column
    (``scale[n] = max |w[:, n]| / 127``) or per group of ``group_size`` input rows (``scale [k / g, n]``);
  * int4: ``[n, k / 2]`` int8, two signed nibbles per byte (even k in the low nibble), bound 7.

The matmul dequantises the weight to the activation dtype and goes through ``F.linear`` - i.e. the tcgen05 GEMM on B200;
what is saved is weight memory and bandwidth of the stored model, not GEMM time.  ``llm_int8_linear`` does the integer
matmul exactly (int32 accumulation) for the non-outlier features and keeps outlier features in floating point.
"""
from __future__ import annotations

import torch

from ..layer.layers import Layer

__all__ = ["Stub", "weight_only_linear", "llm_int8_linear", "weight_quantize", "weight_dequantize"]

_BOUND = {"weight_only_int8": 127.0, "llm.int8": 127.0, "weight_only_int4": 7.0}


def _check(algo, group_size):
    if algo not in _BOUND:
        raise ValueError("The algo must be in ['weight_only_int8', 'weight_only_int4', 'llm.int8'], but got " + str(algo))
    assert group_size in (-1, 64, 128), f"Currently group_size only support -1/64/128. but got {group_size} "


def _pack_int4(q):
    """[n, k] values in [-8, 7] -> [n, k / 2] bytes."""
    if q.shape[1] % 2:
        raise ValueError("weight_only_int4 needs an even number of input features")
    lo, hi = q[:, 0::2] & 0x0F, q[:, 1::2] & 0x0F
    return (lo | (hi << 4)).to(torch.uint8).view(torch.int8)


def _unpack_int4(p):
    b = p.view(torch.uint8).to(torch.int16)
    lo, hi = b & 0x0F, (b >> 4) & 0x0F
    lo = torch.where(lo > 7, lo - 16, lo)
    hi = torch.where(hi > 7, hi - 16, hi)
    return torch.stack([lo, hi], -1).reshape(p.shape[0], -1)


def weight_quantize(x, algo="weight_only_int8", arch=None, group_size=-1):
    """``x`` [k, n] float -> (quantised weight, float32 scale).  See the module docstring for the layouts."""
    _check(algo, group_size)
    if x.dim() != 2:
        raise ValueError("weight_quantize expects a 2-D weight")
    k, n = x.shape
    bound = _BOUND[algo]
    w = x.float()
    if group_size == -1:
        scale = w.abs().amax(0) / bound                                          # [n]
        full = scale[None, :]
    else:
        if k % group_size:
            raise ValueError(f"the number of input features {k} must be a multiple of group_size {group_size}")
        scale = w.reshape(k // group_size, group_size, n).abs().amax(1) / bound  # [k / g, n]
        full = scale.repeat_interleave(group_size, 0)
    q = torch.round(w / full.clamp_min(1e-30)).clamp(-bound, bound).to(torch.int16).t().contiguous()   # [n, k]
    out = _pack_int4(q) if algo == "weight_only_int4" else q.to(torch.int8)
    return out, scale.to(torch.float32)


def weight_dequantize(x, scale, algo="weight_only_int8", out_dtype="float16", group_size=-1):
    """Back to the ``[k, n]`` float weight."""
    _check(algo, group_size)
    from ...framework import dtype as _dt
    td = out_dtype if isinstance(out_dtype, torch.dtype) else _dt.to_torch_dtype(out_dtype)
    q = _unpack_int4(x) if algo == "weight_only_int4" else x.to(torch.int16)
    w = q.t().float()                                                            # [k, n]
    sc = scale.float()
    full = sc[None, :] if sc.dim() == 1 else sc.repeat_interleave(w.shape[0] // sc.shape[0], 0)
    return (w * full).to(td)


def weight_only_linear(x, weight, bias=None, weight_scale=None, weight_dtype="int8", arch=None, group_size=-1):
    """``x @ dequantise(weight) + bias`` with ``weight`` as produced by ``weight_quantize``."""
    if weight_dtype not in ("int8", "int4"):
        raise ValueError("weight_dtype must be 'int8' or 'int4'")
    assert group_size in (-1, 64, 128), f"Currently weight_quantize only support group size of -1, 64 or 128. but got {group_size} "
    if weight_scale is None:
        raise ValueError("weight_scale is required")
    from .. import functional as F
    algo = "weight_only_int8" if weight_dtype == "int8" else "weight_only_int4"
    w = weight_dequantize(weight, weight_scale, algo, x.dtype, group_size)       # [k, n]
    return F.linear(x, w, bias)


def llm_int8_linear(x, weight, bias=None, weight_scale=None, threshold=6.0):
    """LLM.int8() (Dettmers et al.): input features where some activation exceeds ``threshold`` in magnitude are
    multiplied in floating point with the dequantised weight rows; the other features are quantised per token to int8
    (abs-max / 127), multiplied with the int8 weight in exact integer arithmetic and rescaled by
    (token scale x output-channel scale)."""
    if weight_scale is None:
        raise ValueError("weight_scale is required")
    shape = x.shape
    x2 = x.reshape(-1, shape[-1])
    xf = x2.float()
    outlier = (xf.abs() > threshold).any(0)                                       # [k]
    wq = weight.to(torch.int16)                                                   # [n, k]
    sc = weight_scale.float()                                                     # [n]
    regular = ~outlier
    xr = xf[:, regular]
    row_scale = xr.abs().amax(1, keepdim=True).clamp_min(1e-30) / 127.0
    xq = torch.round(xr / row_scale).clamp(-127, 127)
    wr = wq[:, regular]
    if x.is_cuda:
        acc = (xq.double() @ wr.double().t()).float()                             # exact: |sum| < 2^53
    else:
        acc = (xq.to(torch.int32) @ wr.to(torch.int32).t()).float()
    out = acc * row_scale * sc[None, :]
    if outlier.any():
        out = out + xf[:, outlier] @ (wq[:, outlier].float() * sc[:, None]).t()
    if bias is not None:
        out = out + bias.float()
    return out.to(x.dtype).reshape(*shape[:-1], weight.shape[0])


def apply_per_channel_scale(x, scales):
    """Pre-quantisation smoothing: ``x * scales`` along the feature dim."""
    return x * scales.to(x.dtype)


class Stub(Layer):
    """A place holder in a network's ``forward`` where quantisation should observe a functional call's input.  With no
    observer it is the identity; ``QAT.quantize`` / ``PTQ.quantize`` replace it by the configured quanter
    (reference: python/paddle/nn/quant/stub.py)."""

    def __init__(self, observer=None):
        super().__init__()
        self._observer = observer

    def forward(self, input):
        return input


class QuanterStub(Layer):
    """What a ``Stub`` becomes after ``quantize``: the stub's own observer factory, or the layer's activation config."""

    def __init__(self, layer, q_config):
        super().__init__()
        self._observer = None
        factory = layer._observer if layer._observer is not None else q_config.activation
        if factory is not None:
            self._observer = factory._instance(layer)

    def forward(self, input):
        return self._observer(input) if self._observer is not None else input
