"""Linear op with Paddle's [in, out] weight layout.

Forward  Y = X W (+ b);  backward  dX = dY W^T,  dW = X^T dY (+ into fp32 ``main_grad`` when the parameter carries
one - the reference's fused_linear_param_grad_add, paddle/phi/kernels/fusion/gpu/fused_linear_param_grad_add_kernel.cu:35-289,
K13).  GEMMs go to the tcgen05/TMA kernel (ops.gemm, K23) when FLAGS_use_tcgen05_gemm is on and the shape fits its
tiling, otherwise to cuBLAS through torch.matmul (plain library GEMM).
"""
from __future__ import annotations

import torch

from ..framework.flags import flag
from . import gemm as _gemm


def _mm(a, b, trans_a=False, trans_b=False, out_dtype=None, accum_into=None, out=None):
    """2-D GEMM dispatcher; accum_into (fp32) += result when given; ``out``: preallocated contiguous result."""
    if a.is_cuda and flag("use_tcgen05_gemm") and _gemm.supported(a, b, trans_a, trans_b, accum_into):
        return _gemm.gemm(a, b, trans_a, trans_b, out_dtype=out_dtype, accum_into=accum_into, out=out)
    A = a.t() if trans_a else a
    B = b.t() if trans_b else b
    if out is not None:
        return torch.matmul(A, B, out=out)
    if accum_into is not None:
        if accum_into.dtype == A.dtype:
            accum_into.addmm_(A, B)
        else:
            accum_into.add_(torch.matmul(A, B))
        return accum_into
    out = torch.matmul(A, B)
    return out if out_dtype is None or out.dtype == out_dtype else out.to(out_dtype)


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        x2 = x.reshape(-1, x.shape[-1])
        y = _mm(x2, w)
        if b is not None:
            y = y + b
        ctx.save_for_backward(x2, w)
        ctx.has_b = b is not None
        ctx.x_shape = x.shape
        ctx.wref = w
        return y.reshape(*x.shape[:-1], w.shape[1])

    @staticmethod
    def backward(ctx, dy):
        x2, w = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        dx = _mm(dy2, w, trans_b=True).reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = None
        if ctx.needs_input_grad[1]:
            mg = getattr(ctx.wref, "main_grad", None)
            if mg is not None and getattr(ctx.wref, "_fuse_main_grad", False):
                _mm(x2, dy2, trans_a=True, accum_into=mg)
                dw = None
            else:
                # a sharding wrapper may name the landing buffer of this gradient (ZeRO-3 gradient stage in symmetric
                # memory): the GEMM writes there and autograd adopts a fresh alias of it - no staging copy
                dst = getattr(ctx.wref, "_grad_dst", None)
                if dst is not None and dst.dtype == dy2.dtype and tuple(dst.shape) == (x2.shape[1], dy2.shape[1]):
                    dw = _mm(x2, dy2, trans_a=True, out=dst).view(dst.shape)
                else:
                    dw = _mm(x2, dy2, trans_a=True)
        db = dy2.sum(0) if (ctx.has_b and ctx.needs_input_grad[2]) else None
        return dx, dw, db


def _amp_black_listed():
    """``matmul`` / ``linear`` put on auto_cast's custom black list: the product stays in the operands' precision."""
    from ..amp import _state
    black = _state.get("black")
    return bool(black) and bool(black & {"matmul", "matmul_v2", "linear", "mul"})


def linear(x, weight, bias=None):
    if hasattr(x, "placements") or hasattr(weight, "placements"):
        # distributed tensors (auto_parallel): the matmul SPMD rule decides resharding, the local product comes back here
        from ..distributed.auto_parallel import dtensor as D
        return D.dist_linear(x, weight, bias)
    if torch.is_autocast_enabled(x.device.type) and not _amp_black_listed():
        dt = torch.get_autocast_dtype(x.device.type)
        x = x.to(dt)
        weight_c = weight.to(dt)
        bias = bias.to(dt) if bias is not None else None
    else:
        weight_c = weight
    needs_custom = x.is_cuda and (flag("use_tcgen05_gemm") or getattr(weight, "_fuse_main_grad", False)) \
        and x.dtype in (torch.bfloat16, torch.float16) and weight_c is weight
    if needs_custom:
        return _LinearFn.apply(x, weight, bias)
    y = torch.matmul(x, weight_c)
    return y if bias is None else y + bias
