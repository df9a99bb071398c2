"""Attention op (K16/K17).  Layout is Paddle's [batch, seq, heads, head_dim].

Dispatch: the hand-written tcgen05 flash-attention kernels (csrc/kernels/flash_attn.cu, exported as
``pdb200_flash_attn_fwd/bwd`` + ``_dropout`` + ``_varlen_`` variants) when present and the shape fits (bf16/fp16,
head_dim 64/128, any sequence lengths, dense or packed, dropout, no explicit mask); otherwise torch's
scaled_dot_product_attention (library path, also the numerical oracle).
Reference: paddle/phi/kernels/gpu/flash_attn_kernel.cu:305-447, flash_attn_grad_kernel.cu,
python/paddle/nn/functional/flash_attention.py:147-960."""
from __future__ import annotations

import math

import torch
import torch.nn.functional as TF

from .. import lib
from ..framework.flags import flag
from ._util import check, dtype_code, ptr, stream_ptr


def _native_ok(q, k, v, dropout, mask):
    if not (q.is_cuda and flag("use_native_kernels") and lib.have_kernels()):
        return False
    K = lib.kernels.require()
    if not hasattr(K, "pdb200_flash_attn_fwd"):
        return False
    if q.dtype not in (torch.bfloat16, torch.float16) or k.dtype != q.dtype or v.dtype != q.dtype or mask is not None:
        return False
    if dropout > 0.0 and not hasattr(K, "pdb200_flash_attn_fwd_dropout"):
        return False
    if not hasattr(K, "pdb200_flash_attn_varlen_fwd") and (q.shape[-3] % 128 or k.shape[-3] % 128):
        return False                                     # a library built before tails were handled in-kernel
    return q.shape[-1] in (64, 128) and q.shape[-2] % k.shape[-2] == 0 and 0.0 <= dropout < 1.0


def _seed_offset(device, fixed, numel):
    """(seed, offset) of this call's dropout mask.  Taken from the device's default generator, which is what
    ``paddle.seed`` and the TP RNG tracker (fleet/layers/mpu/random.py) switch, and advanced past this call so the next
    one draws a different mask; ``fixed_seed_offset`` pins both (reference: flash_attention.py ``fixed_seed_offset``)."""
    if fixed is not None:
        so = fixed.tolist() if isinstance(fixed, torch.Tensor) else list(fixed)
        return int(so[0]), int(so[1])
    gen = torch.cuda.default_generators[device.index if device.index is not None else torch.cuda.current_device()]
    seed, off = gen.initial_seed(), gen.get_offset()
    gen.set_offset(off + 4 * ((numel + 3) // 4 // 1024 + 1))
    return seed & ((1 << 63) - 1), off


def _bwd_ws(q, rows_hq):
    return torch.empty(max(rows_hq * 4, 16), dtype=torch.uint8, device=q.device)


class _FlashFn(torch.autograd.Function):
    """Dense [b, s, h, d] attention on the tcgen05 kernels; any sequence lengths, bf16 / fp16, optional dropout."""

    @staticmethod
    def forward(ctx, q, k, v, causal, scale, p_drop, seed, offset):
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        b, sq, hq, d = q.shape
        sk, hk = k.shape[1], k.shape[2]
        o = torch.empty_like(q)
        lse = torch.empty(b, hq, sq, dtype=torch.float32, device=q.device)
        K = lib.kernels
        if p_drop > 0.0:
            check(K.pdb200_flash_attn_fwd_dropout(ptr(q), ptr(k), ptr(v), ptr(o), ptr(lse), b, sq, sk, hq, hk, d, float(scale),
                                                  int(causal), float(p_drop), seed, offset, dtype_code(q),
                                                  stream_ptr(q.device)), "flash_attn_fwd_dropout")
        else:
            check(K.pdb200_flash_attn_fwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(lse), b, sq, sk, hq, hk, d, float(scale),
                                          int(causal), dtype_code(q), stream_ptr(q.device)), "flash_attn_fwd")
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.cfg = (causal, scale, p_drop, seed, offset)
        return o, lse

    @staticmethod
    def backward(ctx, do, _dlse):
        q, k, v, o, lse = ctx.saved_tensors
        causal, scale, p_drop, seed, offset = ctx.cfg
        b, sq, hq, d = q.shape
        sk, hk = k.shape[1], k.shape[2]
        do = do.contiguous()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        ws = _bwd_ws(q, b * sq * hq)
        K = lib.kernels
        if p_drop > 0.0:
            check(K.pdb200_flash_attn_bwd_dropout(ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(dq), ptr(dk), ptr(dv),
                                                  ptr(ws), ws.numel(), b, sq, sk, hq, hk, d, float(scale), int(causal),
                                                  float(p_drop), seed, offset, dtype_code(q), stream_ptr(q.device)),
                  "flash_attn_bwd_dropout")
        else:
            check(K.pdb200_flash_attn_bwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(dq), ptr(dk), ptr(dv),
                                          ptr(ws), ws.numel(), b, sq, sk, hq, hk, d, float(scale), int(causal),
                                          dtype_code(q), stream_ptr(q.device)), "flash_attn_bwd")
        return dq, dk, dv, None, None, None, None, None


class _FlashVarlenFn(torch.autograd.Function):
    """Packed [total, h, d] attention: the kernels read cu_seqlens on the device, one launch for the whole batch."""

    @staticmethod
    def forward(ctx, q, k, v, cu_q, cu_k, max_q, max_k, causal, scale, p_drop, seed, offset):
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        tq, hq, d = q.shape
        tk, hk = k.shape[0], k.shape[1]
        nb = cu_q.numel() - 1
        o = torch.empty_like(q)
        lse = torch.empty(hq, tq, dtype=torch.float32, device=q.device)
        check(lib.kernels.pdb200_flash_attn_varlen_fwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(lse), ptr(cu_q), ptr(cu_k), nb, tq,
                                                       tk, int(max_q), int(max_k), hq, hk, d, float(scale), int(causal),
                                                       float(p_drop), seed, offset, dtype_code(q), stream_ptr(q.device)),
              "flash_attn_varlen_fwd")
        ctx.save_for_backward(q, k, v, o, lse, cu_q, cu_k)
        ctx.cfg = (int(max_q), int(max_k), causal, scale, p_drop, seed, offset)
        return o, lse

    @staticmethod
    def backward(ctx, do, _dlse):
        q, k, v, o, lse, cu_q, cu_k = ctx.saved_tensors
        max_q, max_k, causal, scale, p_drop, seed, offset = ctx.cfg
        tq, hq, d = q.shape
        tk, hk = k.shape[0], k.shape[1]
        do = do.contiguous()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        ws = _bwd_ws(q, tq * hq)
        check(lib.kernels.pdb200_flash_attn_varlen_bwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(dq), ptr(dk),
                                                       ptr(dv), ptr(ws), ws.numel(), ptr(cu_q), ptr(cu_k), cu_q.numel() - 1,
                                                       tq, tk, max_q, max_k, hq, hk, d, float(scale), int(causal),
                                                       float(p_drop), seed, offset, dtype_code(q), stream_ptr(q.device)),
              "flash_attn_varlen_bwd")
        return (dq, dk, dv) + (None,) * 9


def _sdpa(q, k, v, causal, scale, dropout, mask, training):
    qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
    gqa = qt.shape[1] != kt.shape[1]
    if causal and mask is None and q.shape[1] != k.shape[1]:
        # flash-attention convention (and the native kernel's): the causal diagonal is aligned to the LAST key, i.e.
        # query i sees keys <= i + (sk - sq); torch's is_causal aligns it to the first key instead
        sq, sk = q.shape[1], k.shape[1]
        mask = torch.ones(sq, sk, dtype=torch.bool, device=q.device).tril(sk - sq)
    out = TF.scaled_dot_product_attention(qt, kt, vt, attn_mask=mask, dropout_p=dropout if training else 0.0,
                                          is_causal=causal and mask is None, scale=scale, enable_gqa=gqa)
    return out.transpose(1, 2)


def flash_attention(q, k, v, dropout=0.0, causal=False, softmax_scale=None, attn_mask=None, training=True,
                    return_lse=False, fixed_seed_offset=None):
    if hasattr(q, "placements") or hasattr(k, "placements") or hasattr(v, "placements"):
        # distributed tensors (auto_parallel): batch and heads may stay sharded, the rule gathers the rest
        from ..distributed.auto_parallel import dtensor as D
        if attn_mask is not None and hasattr(attn_mask, "placements"):
            attn_mask = D.replicate_all((attn_mask,), {})[0][0]
        out = D.run_rule("flash_attention",
                         lambda a, b, c: flash_attention(a, b, c, dropout, causal, softmax_scale, attn_mask, training, True,
                                                         fixed_seed_offset), [q, k, v])
        return out if return_lse else out[0]
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(q.shape[-1])
    p_drop = float(dropout) if training else 0.0
    if _native_ok(q, k, v, p_drop, attn_mask):
        seed, offset = _seed_offset(q.device, fixed_seed_offset, q.shape[0] * q.shape[2] * q.shape[1] * k.shape[1]) \
            if p_drop > 0.0 else (0, 0)
        o, lse = _FlashFn.apply(q, k, v, bool(causal), float(scale), p_drop, seed, offset)
        return (o, lse) if return_lse else o
    o = _sdpa(q, k, v, causal, scale, dropout, attn_mask, training)
    return (o, None) if return_lse else o


def flash_attn_varlen(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale, dropout=0.0,
                      causal=False, training=True, fixed_seed_offset=None):
    """Packed variable-length attention: q [total_q, h, d], cu_seqlens int32 [b + 1] on the device.  One kernel launch
    for the whole batch, no host read of the lengths (reference: paddle/phi/kernels/gpu/flash_attn_kernel.cu:36-170).
    Without the native library: sequence by sequence through the dense op."""
    p_drop = float(dropout) if training else 0.0
    if _native_ok(q, k, v, p_drop, None) and hasattr(lib.kernels.require(), "pdb200_flash_attn_varlen_fwd"):
        cu_q = cu_seqlens_q.to(device=q.device, dtype=torch.int32).contiguous()
        cu_k = cu_seqlens_k.to(device=q.device, dtype=torch.int32).contiguous()
        seed, offset = _seed_offset(q.device, fixed_seed_offset, q.shape[0] * q.shape[1] * int(max_seqlen_k)) \
            if p_drop > 0.0 else (0, 0)
        return _FlashVarlenFn.apply(q, k, v, cu_q, cu_k, int(max_seqlen_q), int(max_seqlen_k), bool(causal), float(scale),
                                    p_drop, seed, offset)[0]
    outs = []
    cq, ck = cu_seqlens_q.tolist(), cu_seqlens_k.tolist()
    for i in range(len(cq) - 1):
        qi, ki, vi = q[cq[i]:cq[i + 1]][None], k[ck[i]:ck[i + 1]][None], v[ck[i]:ck[i + 1]][None]
        outs.append(flash_attention(qi, ki, vi, dropout, causal, scale, None, training)[0])
    return torch.cat(outs, 0)
