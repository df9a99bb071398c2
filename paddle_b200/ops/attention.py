"""Attention op (K16/K17).  Layout is Paddle's [batch, seq, heads, head_dim].

Dispatch: the hand-written tcgen05 flash-attention kernels (csrc/kernels/flash_attn.cu, exported as
``pdb200_flash_attn_fwd/bwd``) when present and the shape fits (bf16/fp16, head_dim 64/128, no dropout); otherwise
torch's scaled_dot_product_attention (library path, also the numerical oracle).
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
    if not hasattr(lib.kernels.require(), "pdb200_flash_attn_fwd"):
        return False
    if q.dtype != torch.bfloat16 or dropout > 0.0 or mask is not None:
        return False
    d = q.shape[-1]
    return d in (64, 128) and q.shape[1] % 128 == 0 and k.shape[1] % 128 == 0 and q.shape[2] % k.shape[2] == 0


class _FlashFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, causal, scale):
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        b, sq, hq, d = q.shape
        sk, hk = k.shape[1], k.shape[2]
        o = torch.empty_like(q)
        lse = torch.empty(b, hq, sq, dtype=torch.float32, device=q.device)
        check(lib.kernels.pdb200_flash_attn_fwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(lse), b, sq, sk, hq, hk, d,
                                                float(scale), int(causal), dtype_code(q), stream_ptr(q.device)),
              "flash_attn_fwd")
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.cfg = (causal, scale)
        return o, lse

    @staticmethod
    def backward(ctx, do, _dlse):
        q, k, v, o, lse = ctx.saved_tensors
        causal, scale = ctx.cfg
        b, sq, hq, d = q.shape
        sk, hk = k.shape[1], k.shape[2]
        do = do.contiguous()
        if hasattr(lib.kernels.require(), "pdb200_flash_attn_bwd"):
            dq = torch.empty_like(q)
            dk = torch.empty_like(k)
            dv = torch.empty_like(v)
            ws_bytes = lib.kernels.pdb200_flash_attn_bwd_workspace_bytes(b, sq, sk, hq, hk, d)
            ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=q.device)
            check(lib.kernels.pdb200_flash_attn_bwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(dq), ptr(dk),
                                                    ptr(dv), ptr(ws), ws_bytes, b, sq, sk, hq, hk, d, float(scale),
                                                    int(causal), dtype_code(q), stream_ptr(q.device)), "flash_attn_bwd")
            return dq, dk, dv, None, None
        # no native backward yet: recompute through the library path
        with torch.enable_grad():
            qq, kk, vv = (t.detach().requires_grad_(True) for t in (q, k, v))
            out = _sdpa(qq, kk, vv, causal, scale, 0.0, None, True)
            dq, dk, dv = torch.autograd.grad(out, (qq, kk, vv), do)
        return dq, dk, dv, None, None


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
                    return_lse=False):
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if _native_ok(q, k, v, dropout if training else 0.0, attn_mask):
        o, lse = _FlashFn.apply(q, k, v, bool(causal), float(scale))
        return (o, lse) if return_lse else o
    o = _sdpa(q, k, v, causal, scale, dropout, attn_mask, training)
    return (o, None) if return_lse else o


def flash_attn_varlen(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale, dropout=0.0,
                      causal=False, training=True):
    """Packed variable-length attention: q [total_q, h, d].  Sequences are processed one by one through the dense
    op (correct for any lengths; the dense kernel handles each)."""
    outs = []
    cq, ck = cu_seqlens_q.tolist(), cu_seqlens_k.tolist()
    for i in range(len(cq) - 1):
        qi, ki, vi = q[cq[i]:cq[i + 1]][None], k[ck[i]:ck[i + 1]][None], v[ck[i]:ck[i + 1]][None]
        outs.append(flash_attention(qi, ki, vi, dropout, causal, scale, None, training)[0])
    return torch.cat(outs, 0)
