"""RMSNorm / LayerNorm ops (K18/K19).  Reference API: python/paddle/incubate/nn/functional/fused_rms_norm.py:20-72,
fused_layer_norm.py, nn/functional/norm.py:layer_norm."""
from __future__ import annotations

import torch

from ._util import K, check, dtype_code, ptr, stream_ptr, use_native


def _norm_ref(x, w, b, eps, layernorm, residual=None):
    xf = x.float()
    if residual is not None:
        xf = (x + residual).float()
        res_out = (x + residual)
    if layernorm:
        mean = xf.mean(-1, keepdim=True)
        var = xf.var(-1, unbiased=False, keepdim=True)
        y = (xf - mean) * torch.rsqrt(var + eps)
    else:
        y = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    if w is not None:
        y = y * w.float()
    if b is not None:
        y = y + b.float()
    y = y.to(x.dtype)
    return (y, res_out) if residual is not None else y


class _NormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, residual, eps, layernorm):
        cols = x.shape[-1]
        xc = x.contiguous()
        rows = xc.numel() // cols
        y = torch.empty_like(xc)
        res_c = residual.contiguous() if residual is not None else None
        res_out = torch.empty_like(xc) if residual is not None else None
        mean = torch.empty(rows, dtype=torch.float32, device=x.device) if layernorm else None
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        w_is_f32 = int(w is not None and w.dtype == torch.float32)
        if w is not None and b is not None and b.dtype != w.dtype:
            b = b.to(w.dtype)
        wc = w.contiguous() if w is not None else None
        bc = b.contiguous() if b is not None else None
        check(K().pdb200_norm_fwd(ptr(xc), ptr(res_c), ptr(wc), ptr(bc), ptr(y), ptr(res_out), ptr(mean), ptr(rstd),
                                  rows, cols, float(eps), dtype_code(xc), w_is_f32, int(layernorm),
                                  stream_ptr(x.device)), "norm_fwd")
        ctx.save_for_backward(res_out if residual is not None else xc, wc, mean, rstd)
        ctx.layernorm, ctx.has_b, ctx.has_res = layernorm, b is not None, residual is not None
        ctx.w_is_f32 = w_is_f32
        if residual is not None:
            return y, res_out
        return y

    @staticmethod
    def backward(ctx, dy, *dres):
        xin, w, mean, rstd = ctx.saved_tensors
        cols = xin.shape[-1]
        rows = xin.numel() // cols
        dyc = dy.contiguous()
        if ctx.has_res and dres and dres[0] is not None:
            pass  # gradient flowing into residual_out is added after the norm backward below
        dx = torch.empty_like(xin)
        need_w = w is not None and ctx.needs_input_grad[1]
        dw = torch.empty_like(w) if need_w else None
        db = torch.empty_like(w) if (ctx.has_b and ctx.needs_input_grad[2]) else None
        need_param = dw is not None or db is not None
        if need_param and dw is None:
            dw = torch.empty_like(w)
        scratch_bytes = K().pdb200_norm_bwd_scratch_bytes(rows, cols, int(ctx.layernorm)) if need_param else 0
        scratch = torch.empty(max(scratch_bytes, 4) // 4, dtype=torch.float32, device=xin.device) if need_param else None
        check(K().pdb200_norm_bwd(ptr(dyc), ptr(xin), ptr(w), ptr(mean), ptr(rstd), ptr(dx), ptr(dw), ptr(db),
                                  ptr(scratch), scratch_bytes, rows, cols, dtype_code(xin), ctx.w_is_f32,
                                  int(ctx.layernorm), stream_ptr(xin.device)), "norm_bwd")
        if ctx.has_res and dres and dres[0] is not None:
            dx = dx + dres[0]
        return dx, (dw if need_w else None), db, (dx if ctx.has_res else None), None, None


def rms_norm(x, weight=None, epsilon=1e-6, bias=None, residual=None):
    if hasattr(x, "placements") or hasattr(weight, "placements"):
        from ..distributed.auto_parallel import dtensor as D
        if residual is not None:
            x = x + residual                      # element-wise rule; the fused form is a local optimisation only
            y = D.run_rule("rms_norm", lambda a, w, b: (rms_norm(a, w, epsilon, b),), [x, weight, bias], {"begin_norm_axis": -1})[0]
            return y, x
        return D.run_rule("rms_norm", lambda a, w, b: (rms_norm(a, w, epsilon, b),), [x, weight, bias], {"begin_norm_axis": -1})[0]
    if use_native(x):
        return _NormFn.apply(x, weight, bias, residual, epsilon, False)
    return _norm_ref(x, weight, bias, epsilon, False, residual)


def layer_norm(x, normalized_shape, weight=None, bias=None, epsilon=1e-5, residual=None):
    n = 1
    for s in normalized_shape:
        n *= int(s)
    shp = x.shape
    x2 = x.reshape(-1, n)
    w2 = weight.reshape(-1) if weight is not None else None
    b2 = bias.reshape(-1) if bias is not None else None
    if use_native(x):
        if w2 is None and b2 is not None:
            w2 = torch.ones_like(b2)
        out = _NormFn.apply(x2, w2, b2, residual.reshape(-1, n) if residual is not None else None, epsilon, True)
    else:
        out = _norm_ref(x2, w2, b2, epsilon, True, residual.reshape(-1, n) if residual is not None else None)
    if residual is not None:
        return out[0].reshape(shp), out[1].reshape(shp)
    return out.reshape(shp)
