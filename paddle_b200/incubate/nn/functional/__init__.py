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


def fused_bias_dropout_residual_layer_norm(x, residual, bias=None, ln_scale=None, ln_bias=None, dropout_rate=0.5,
                                           ln_epsilon=1e-5, training=True, mode="upscale_in_train", name=None):
    """layer_norm(residual + dropout(x + bias)) (reference: incubate/nn/functional/fused_transformer.py:236)."""
    y = x if bias is None else x + bias
    y = torch.nn.functional.dropout(y, dropout_rate, training) if mode == "upscale_in_train" else \
        (torch.nn.functional.dropout(y, dropout_rate, training) * (1 - dropout_rate) if training else y * (1 - dropout_rate))
    return F.layer_norm(residual + y, [x.shape[-1]], ln_scale, ln_bias, ln_epsilon)


def fused_ec_moe(x, gate, bmm0_weight, bmm0_bias, bmm1_weight, bmm1_bias, act_type="gelu"):
    """Expert-choice style dense MoE: every expert FFN is evaluated with batched GEMMs and mixed by softmax(gate)
    (reference: incubate/nn/functional/fused_ec_moe.py).  x [b,s,d], gate [b,s,e], weights [e,d,f] / [e,f,d]."""
    b, s, d = x.shape
    probs = torch.softmax(gate.float(), -1).to(x.dtype)                       # [b, s, e]
    xe = x.reshape(1, b * s, d).expand(bmm0_weight.shape[0], -1, -1)           # [e, tokens, d]
    h = torch.baddbmm(bmm0_bias, xe, bmm0_weight)
    h = {"gelu": torch.nn.functional.gelu, "relu": torch.relu}[act_type](h)
    y = torch.baddbmm(bmm1_bias, h, bmm1_weight)                               # [e, tokens, d]
    return torch.einsum("etd,te->td", y, probs.reshape(b * s, -1)).reshape(b, s, d)


def variable_length_memory_efficient_attention(query, key, value, seq_lens, kv_seq_lens, mask=None, scale=None,
                                               causal=False, pre_cache_length=0):
    """Attention over padded [b, h, s, d] batches with per-sample valid lengths (reference: incubate/nn/functional/
    variable_length_memory_efficient_attention.py); padded keys are masked, padded query rows return zeros."""
    b, h, sq, d = query.shape
    sk = key.shape[2]
    scale = scale if scale is not None else d ** -0.5
    kq = torch.arange(sq, device=query.device)[None, :] < seq_lens.reshape(-1, 1)
    kk = torch.arange(sk, device=query.device)[None, :] < kv_seq_lens.reshape(-1, 1)
    m = (kq[:, None, :, None] & kk[:, None, None, :])
    if causal:
        m = m & torch.ones(sq, sk, dtype=torch.bool, device=query.device).tril(sk - sq)
    bias = torch.zeros(b, 1, sq, sk, dtype=query.dtype, device=query.device).masked_fill(~m, float("-inf"))
    if mask is not None:
        bias = bias + mask
    sc = torch.matmul(query, key.transpose(-1, -2)) * scale + bias
    p = torch.nan_to_num(torch.softmax(sc.float(), -1), nan=0.0).to(query.dtype)
    return torch.matmul(p, value)


def _inference_only(name):
    def f(*a, **k):
        raise NotImplementedError(f"{name} is a decode-time inference kernel of the reference (fused_multi_transformer / "
                                  "block attention family); this framework targets training (SURVEY 2.6.b, out of scope)")
    f.__name__ = name
    return f


fused_multi_transformer = _inference_only("fused_multi_transformer")
masked_multihead_attention = _inference_only("masked_multihead_attention")
block_multihead_attention = _inference_only("block_multihead_attention")
blha_get_max_len = _inference_only("blha_get_max_len")
