"""paddle.incubate.nn.functional fused ops (reference: python/paddle/incubate/nn/functional/*.py).  Each maps onto a
hand-written sm_100a kernel from ``paddle_b200.ops`` (K18-K23)."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.nn import functional as F


def fused_rms_norm(x, norm_weight, norm_bias, epsilon, begin_norm_axis, bias=None, residual=None, quant_scale=-1,
                   quant_round_type=0, quant_max_bound=0, quant_min_bound=0):
    """Returns ``out`` or ``(out, residual_out)`` when residual is given (fused_rms_norm.py:20-72)."""
    if quant_scale > 0:
        raise NotImplementedError("int8 quantised output is an inference feature outside this framework's scope")
    shp = x.shape
    cols = 1
    for s in shp[begin_norm_axis:]:
        cols *= s
    x2 = x.reshape(-1, cols)
    if bias is not None:
        x2 = x2 + bias.reshape(-1)
    res2 = residual.reshape(-1, cols) if residual is not None else None
    w = norm_weight.reshape(-1) if norm_weight is not None else None
    b = norm_bias.reshape(-1) if norm_bias is not None else None
    out = ops.rms_norm(x2, w, epsilon, bias=b, residual=res2)
    if residual is not None:
        return out[0].reshape(shp), out[1].reshape(shp)
    return out.reshape(shp)


def fused_layer_norm(x, norm_weight, norm_bias, epsilon, residual_alpha=1.0, begin_norm_axis=1, bias=None, residual=None,
                     quant_scale=-1, quant_round_type=0, quant_max_bound=0, quant_min_bound=0):
    shp = x.shape
    norm_shape = list(shp[begin_norm_axis:])
    xin = x if bias is None else x + bias
    if residual is not None:
        res_in = residual * residual_alpha if residual_alpha != 1.0 else residual
        out, res_out = ops.layer_norm(xin, norm_shape, norm_weight, norm_bias, epsilon, residual=res_in)
        return out, res_out
    return ops.layer_norm(xin, norm_shape, norm_weight, norm_bias, epsilon)


def fused_rotary_position_embedding(q, k=None, v=None, sin=None, cos=None, position_ids=None, use_neox_rotary_style=True,
                                    time_major=False, rotary_emb_base=10000.0):
    if (sin is None) or (cos is None):
        assert position_ids is None, "position_ids without sin/cos is not correctly supported now."
        assert use_neox_rotary_style, "rotate_half without sin/cos is not correctly supported now."
    return ops.fused_rope(q, k, v, sin, cos, position_ids, use_neox_rotary_style, time_major, rotary_emb_base)


def swiglu(x, y=None, name=None):
    return ops.swiglu(x, y)


def fused_linear(x, weight, bias=None, transpose_weight=False, name=None):
    return F.linear(x, weight.t() if transpose_weight else weight, bias)


def fused_matmul_bias(x, y, bias=None, transpose_x=False, transpose_y=False, name=None):
    a = x.transpose(-1, -2) if transpose_x else x
    b = y.transpose(-1, -2) if transpose_y else y
    out = torch.matmul(a, b)
    return out if bias is None else out + bias


def fused_linear_activation(x, y, bias, trans_x=False, trans_y=False, activation=None):
    out = fused_matmul_bias(x, y, bias, trans_x, trans_y)
    if activation in (None, "none"):
        return out
    return getattr(F, activation)(out)


def fused_bias_act(x, bias=None, dequant_scales=None, shift=None, smooth=None, act_method="gelu", compute_dtype="default",
                   quant_scale=-1, quant_round_type=1, quant_max_bound=127.0, quant_min_bound=-127.0):
    if dequant_scales is not None or quant_scale > 0 or shift is not None or smooth is not None:
        raise NotImplementedError("int8 (de)quantisation variants are inference-only and out of scope")
    return ops.fused_bias_act(x, bias, act_method)


def fused_dropout_add(x, y, p=0.5, training=True, mode="upscale_in_train", name=None):
    return F.dropout(x, p, training=training, mode=mode) + y


def fused_linear_param_grad_add(x, dout, dweight=None, dbias=None, multi_precision=True, has_bias=True):
    """dW (+)= X^T dY, db (+)= sum dY (fused_linear_param_grad_add_kernel.cu:35-289; K13)."""
    x2, d2 = x.reshape(-1, x.shape[-1]), dout.reshape(-1, dout.shape[-1])
    acc_dtype = torch.float32 if multi_precision else x.dtype
    if dweight is None:
        dweight = torch.zeros(x2.shape[1], d2.shape[1], dtype=acc_dtype, device=x.device)
    from paddle_b200.ops.linear import _mm
    _mm(x2, d2, trans_a=True, accum_into=dweight)
    if has_bias:
        if dbias is None:
            dbias = torch.zeros(d2.shape[1], dtype=acc_dtype, device=x.device)
        dbias.add_(d2.sum(0).to(dbias.dtype))
        return dweight, dbias
    return dweight, None


def fused_multi_head_attention(*a, **k):
    raise NotImplementedError("fused_multi_head_attention (legacy fused op): use nn.functional.flash_attention")


def fused_feedforward(*a, **k):
    raise NotImplementedError("fused_feedforward (legacy fused op): compose Linear + swiglu/fused_bias_act")
