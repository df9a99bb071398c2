"""paddle.incubate.nn.functional fused ops (reference: python/paddle/incubate/nn/functional/*.py).  Each maps onto a
hand-written sm_100a kernel from ``paddle_b200.ops`` (K18-K23)."""
from __future__ import annotations

import torch
import torch.nn.functional as TF

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


def _mp_sum(t, ring_id):
    """Tensor-parallel partial sums (``ring_id >= 0`` in the reference's fused ops): all-reduce over the fleet mp group."""
    if ring_id is None or ring_id < 0:
        return t
    try:
        from paddle_b200.distributed import fleet
        hcg = fleet.get_hybrid_communicate_group()
        if hcg is None or hcg.get_model_parallel_world_size() <= 1:
            return t
        from paddle_b200.distributed.fleet.layers.mpu import mp_ops
        return mp_ops._mp_allreduce(t, group=hcg.get_model_parallel_group())
    except Exception:  # noqa: BLE001  (fleet not initialised: single group)
        return t


def _dropout(t, p, training, mode):
    if not training or p == 0.0:
        return t if (mode == "upscale_in_train" or not p or training) else t * (1.0 - p)
    out = TF.dropout(t, p, True)
    return out if mode == "upscale_in_train" else out * (1.0 - p)


def fused_multi_head_attention(x, qkv_weight, linear_weight, pre_layer_norm=False, pre_ln_scale=None, pre_ln_bias=None,
                               ln_scale=None, ln_bias=None, pre_ln_epsilon=1e-05, qkv_bias=None, linear_bias=None,
                               cache_kv=None, attn_mask=None, dropout_rate=0.5, attn_dropout_rate=0.5, ln_epsilon=1e-05,
                               training=True, mode="upscale_in_train", ring_id=-1, add_residual=True, num_heads=-1,
                               transpose_qkv_wb=False, name=None):
    """Self-attention block of the reference's fused op (fused_transformer.py:502-800): [pre-LN] -> packed QKV
    projection -> attention (+ additive mask, + KV cache) -> output projection -> dropout -> [+ residual] -> [post-LN].
    ``qkv_weight`` is [3, heads, head_dim, embed] (or [embed, 3 * embed] with ``transpose_qkv_wb``), ``linear_weight``
    [embed, embed].  The attention itself runs on this framework's flash-attention op when no mask / cache is given.
    Returns ``out`` or ``(out, cache_kv_out)`` when a cache [2, b, heads, s_cache, head_dim] is passed."""
    from paddle_b200 import ops
    b, s, e = x.shape
    residual = x
    h_in = TF.layer_norm(x, (e,), pre_ln_scale, pre_ln_bias, pre_ln_epsilon) if pre_layer_norm else x
    if transpose_qkv_wb:
        assert num_heads > 0, "num_heads is required with transpose_qkv_wb"
        nh = num_heads
        hd = qkv_weight.shape[1] // 3 // nh
        qkv = torch.matmul(h_in, qkv_weight)
        if qkv_bias is not None:
            qkv = qkv + qkv_bias
        qkv = qkv.reshape(b, s, 3, nh, hd)
    else:
        _, nh, hd, _ = qkv_weight.shape
        qkv = torch.matmul(h_in, qkv_weight.reshape(3 * nh * hd, e).t())
        if qkv_bias is not None:
            qkv = qkv + qkv_bias.reshape(-1)
        qkv = qkv.reshape(b, s, 3, nh, hd)
    q, k, v = qkv.unbind(2)                                   # [b, s, h, d]
    cache_out = None
    if cache_kv is not None:
        k = torch.cat([cache_kv[0].transpose(1, 2).to(k.dtype), k], dim=1)
        v = torch.cat([cache_kv[1].transpose(1, 2).to(v.dtype), v], dim=1)
        cache_out = torch.stack([k.transpose(1, 2), v.transpose(1, 2)])
    p_attn = attn_dropout_rate if training else 0.0
    if attn_mask is None and cache_kv is None:
        o = ops.flash_attention(q, k, v, dropout=p_attn, causal=False, training=training)
    else:
        sc = torch.matmul(q.transpose(1, 2) * (hd ** -0.5), k.transpose(1, 2).transpose(-1, -2))      # [b, h, s, sk]
        if attn_mask is not None:
            sc = sc + attn_mask.to(sc.dtype) if attn_mask.dtype != torch.bool else sc.masked_fill(~attn_mask, float("-inf"))
        pr = _dropout(torch.softmax(sc, -1), attn_dropout_rate, training, mode)
        o = torch.matmul(pr, v.transpose(1, 2)).transpose(1, 2)
    o = torch.matmul(o.reshape(b, s, nh * hd), linear_weight)
    o = _mp_sum(o, ring_id)
    if linear_bias is not None:
        o = o + linear_bias
    o = _dropout(o, dropout_rate, training, mode)
    if add_residual:
        o = residual + o
    if not pre_layer_norm:
        o = TF.layer_norm(o, (e,), ln_scale, ln_bias, ln_epsilon)
    return o if cache_out is None else (o, cache_out)


def fused_feedforward(x, linear1_weight, linear2_weight, linear1_bias=None, linear2_bias=None, ln1_scale=None, ln1_bias=None,
                      ln2_scale=None, ln2_bias=None, dropout1_rate=0.5, dropout2_rate=0.5, activation="relu",
                      ln1_epsilon=1e-5, ln2_epsilon=1e-5, pre_layer_norm=False, training=True, mode="upscale_in_train",
                      ring_id=-1, add_residual=True, name=None):
    """Feed-forward block of the reference's fused op (fused_transformer.py:41-300): [pre-LN] -> linear1 -> activation ->
    dropout1 -> linear2 -> dropout2 -> [+ residual] -> [post-LN]; bias + activation go through this framework's fused
    bias-act kernel."""
    e = x.shape[-1]
    residual = x
    h = TF.layer_norm(x, (e,), ln1_scale, ln1_bias, ln1_epsilon) if pre_layer_norm else x
    h = torch.matmul(h, linear1_weight)
    h = fused_bias_act(h, linear1_bias, act_method=activation)
    h = _dropout(h, dropout1_rate, training, mode)
    h = _mp_sum(torch.matmul(h, linear2_weight), ring_id)
    if linear2_bias is not None:
        h = h + linear2_bias
    h = _dropout(h, dropout2_rate, training, mode)
    if add_residual:
        h = residual + h
    if not pre_layer_norm:
        h = TF.layer_norm(h, (e,), ln2_scale, ln2_bias, ln2_epsilon)
    return h


def masked_multihead_attention(x, cache_kv=None, bias=None, src_mask=None, cum_offsets=None, sequence_lengths=None,
                               rotary_tensor=None, beam_cache_offset=None, qkv_out_scale=None, out_shift=None,
                               out_smooth=None, seq_len=1, rotary_emb_dims=0, use_neox_rotary_style=False,
                               compute_dtype="default", out_scale=-1, quant_round_type=1, quant_max_bound=127.0,
                               quant_min_bound=-127.0):
    """One decode step against a KV cache (reference: masked_multihead_attention.py:19-140).  ``x`` [b, 3 * heads * d] is
    the packed projection of the new token, ``cache_kv`` [2, b, heads, max_seq, d]; the token is written at position
    ``sequence_lengths[b]`` (or ``src_mask.shape[-1] - 1``) and attends to positions <= its own.  Returns
    ``(out [b, heads * d], cache_kv)`` with the cache updated in place.  Quantised variants and the beam-search offset are
    not part of this framework."""
    if qkv_out_scale is not None or out_scale > 0 or beam_cache_offset is not None:
        raise NotImplementedError("masked_multihead_attention: int8 / beam-offset variants are outside this framework's scope")
    _, b, nh, max_seq, d = cache_kv.shape
    qkv = x.reshape(b, 3, nh, d)
    if bias is not None:
        qkv = qkv + bias.reshape(1, 3, nh, d).to(qkv.dtype)
    q, k, v = qkv.unbind(1)                                  # [b, h, d]
    if sequence_lengths is not None:
        pos = sequence_lengths.reshape(b).long()
    else:
        t = (src_mask.shape[-1] - 1) if src_mask is not None else 0
        pos = torch.full((b,), t, dtype=torch.long, device=x.device)
    if rotary_tensor is not None and rotary_emb_dims > 0:
        # rotary_tensor: [2, b, 1, 1, d] (cos, sin) of the current position
        cos, sin = rotary_tensor[0].reshape(b, 1, d).to(q.dtype), rotary_tensor[1].reshape(b, 1, d).to(q.dtype)

        def rot(t_):
            if use_neox_rotary_style:
                t1, t2 = t_[..., : d // 2], t_[..., d // 2:]
                return torch.cat([-t2, t1], -1)
            t1, t2 = t_[..., 0::2], t_[..., 1::2]
            return torch.stack([-t2, t1], -1).reshape(t_.shape)
        q, k = q * cos + rot(q) * sin, k * cos + rot(k) * sin
    idx = pos.view(b, 1, 1, 1).expand(b, nh, 1, d)
    cache_kv[0].scatter_(2, idx, k.unsqueeze(2).to(cache_kv.dtype))
    cache_kv[1].scatter_(2, idx, v.unsqueeze(2).to(cache_kv.dtype))
    scores = torch.einsum("bhd,bhsd->bhs", q.float(), cache_kv[0].float()) * (d ** -0.5)
    valid = torch.arange(max_seq, device=x.device)[None, None, :] <= pos.view(b, 1, 1)
    if src_mask is not None:
        m = src_mask.reshape(b, 1, -1).float()
        scores[..., : m.shape[-1]] = scores[..., : m.shape[-1]] + m
    scores = scores.masked_fill(~valid, float("-inf"))
    out = torch.einsum("bhs,bhsd->bhd", torch.softmax(scores, -1), cache_kv[1].float()).to(x.dtype)
    return out.reshape(b, nh * d), cache_kv


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
block_multihead_attention = _inference_only("block_multihead_attention")
blha_get_max_len = _inference_only("blha_get_max_len")
