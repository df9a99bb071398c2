"""SwiGLU (K21) and fused bias+activation (K22).  Reference API: python/paddle/incubate/nn/functional/swiglu.py,
op fused_bias_act (paddle/phi/ops/yaml/fused_ops.yaml:190-199)."""
from __future__ import annotations

import torch
import torch.nn.functional as TF

from ._util import K, check, dtype_code, ptr, stream_ptr, use_native

_ACT = {"gelu": 0, "relu": 1, "swiglu": 2, "geglu": 3, "silu": 4, "swish": 4, "identity": 5, "none": 5}


def _rows_view(t):
    """(tensor usable by the kernel, rows, H, row stride in elements): last dim contiguous, uniform row stride."""
    H = t.shape[-1]
    if t.stride(-1) == 1:
        t2 = t.reshape(-1, H) if t.is_contiguous() else None
        if t2 is not None:
            return t2, t2.shape[0], H, H
        if t.dim() == 2:
            return t, t.shape[0], H, t.stride(0)
        # e.g. a chunk of the last dim of a contiguous [..., 2H] tensor
        lead = t.shape[:-1]
        st = t.stride()
        ok = all(st[i] == st[i + 1] * t.shape[i + 1] for i in range(len(lead) - 1))
        if ok:
            rows = 1
            for s in lead:
                rows *= s
            return t, rows, H, st[-2]
    tc = t.contiguous()
    return tc, tc.numel() // H, H, H


class _SwigluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y):
        xv, rows, H, ldx = _rows_view(x)
        yv, _, _, ldy = _rows_view(y)
        out = torch.empty(x.shape, dtype=x.dtype, device=x.device)
        check(K().pdb200_swiglu_fwd(ptr(xv), ptr(yv), ptr(out), rows, H, ldx, ldy, dtype_code(x), stream_ptr(x.device)),
              "swiglu_fwd")
        ctx.save_for_backward(x, y)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, y = ctx.saved_tensors
        xv, rows, H, ldx = _rows_view(x)
        yv, _, _, ldy = _rows_view(y)
        dout = dout.contiguous()
        # when x and y are the two halves of one packed tensor, write both grads into one packed buffer so that
        # autograd's cat-backward (narrow) does not copy
        packed = (x._base is not None and x._base is y._base and x.dim() >= 1 and
                  x.storage_offset() + H == y.storage_offset() and ldx == 2 * H and ldy == 2 * H)
        if packed:
            buf = torch.empty(*x.shape[:-1], 2 * H, dtype=x.dtype, device=x.device)
            dx, dy = buf[..., :H], buf[..., H:]
            lddx = lddy = 2 * H
        else:
            dx, dy = torch.empty(x.shape, dtype=x.dtype, device=x.device), torch.empty(y.shape, dtype=y.dtype, device=y.device)
            lddx = lddy = H
        check(K().pdb200_swiglu_bwd(ptr(xv), ptr(yv), ptr(dout), ptr(dx), ptr(dy), rows, H, ldx, ldy, lddx, lddy,
                                    dtype_code(x), stream_ptr(x.device)), "swiglu_bwd")
        return dx, dy


class _SwigluPackedFn(torch.autograd.Function):
    """silu(xy[..., :H]) * xy[..., H:] on the packed gate|up projection: the halves are addressed by stride inside the
    kernel and the gradient is produced packed, so neither a split copy nor a cat appears in the graph."""

    @staticmethod
    def forward(ctx, xy):
        xy = xy.contiguous()
        H = xy.shape[-1] // 2
        rows = xy.numel() // (2 * H)
        out = torch.empty(*xy.shape[:-1], H, dtype=xy.dtype, device=xy.device)
        base = xy.data_ptr()
        check(K().pdb200_swiglu_fwd(base, base + H * xy.element_size(), ptr(out), rows, H, 2 * H, 2 * H, dtype_code(xy),
                                    stream_ptr(xy.device)), "swiglu_fwd")
        ctx.save_for_backward(xy)
        return out

    @staticmethod
    def backward(ctx, dout):
        (xy,) = ctx.saved_tensors
        H = xy.shape[-1] // 2
        rows = xy.numel() // (2 * H)
        dout = dout.contiguous()
        dxy = torch.empty_like(xy)
        es = xy.element_size()
        check(K().pdb200_swiglu_bwd(xy.data_ptr(), xy.data_ptr() + H * es, ptr(dout), dxy.data_ptr(), dxy.data_ptr() + H * es,
                                    rows, H, 2 * H, 2 * H, 2 * H, 2 * H, dtype_code(xy), stream_ptr(xy.device)), "swiglu_bwd")
        return dxy


def swiglu(x, y=None):
    if hasattr(x, "placements") or hasattr(y, "placements"):
        from ..distributed.auto_parallel import dtensor as D
        return D.run_rule("swiglu", (lambda a: swiglu(a)) if y is None else (lambda a, b: swiglu(a, b)), [x] if y is None else [x, y])
    if y is None:
        if use_native(x) and x.shape[-1] % 2 == 0 and (x.shape[-1] // 2 * x.element_size()) % 16 == 0:
            return _SwigluPackedFn.apply(x)
        x, y = x.chunk(2, dim=-1)
    if use_native(x, y):
        return _SwigluFn.apply(x, y)
    return TF.silu(x) * y


def fused_bias_act(x, bias=None, act_method="gelu"):
    act = act_method.lower()
    if act not in _ACT:
        raise ValueError(f"unsupported act_method {act_method}")
    glu = act in ("swiglu", "geglu")
    cols_in = x.shape[-1]
    cols_out = cols_in // 2 if glu else cols_in
    if use_native(x) and not torch.is_grad_enabled() or (use_native(x) and not x.requires_grad):
        xc = x.contiguous()
        out = torch.empty(*x.shape[:-1], cols_out, dtype=x.dtype, device=x.device)
        bc = bias.to(x.dtype).contiguous() if bias is not None else None
        check(K().pdb200_bias_act(ptr(xc), ptr(bc), ptr(out), xc.numel() // cols_in, cols_out, _ACT[act],
                                  dtype_code(x), stream_ptr(x.device)), "bias_act")
        return out
    v = x if bias is None else x + bias
    if act == "gelu":
        return TF.gelu(v)
    if act == "relu":
        return TF.relu(v)
    if act in ("silu", "swish"):
        return TF.silu(v)
    if act in ("identity", "none"):
        return v
    a, b = v.chunk(2, dim=-1)
    return (TF.silu(a) if act == "swiglu" else TF.gelu(a)) * b
