"""Attention front-ends (reference: python/paddle/nn/functional/flash_attention.py:147-960)."""
from __future__ import annotations

import contextlib

import torch


@contextlib.contextmanager
def sdp_kernel(enable_math=False, enable_flash=True, enable_mem_efficient=True):
    yield


def flash_attention(query, key, value, dropout=0.0, causal=False, return_softmax=False, *, fixed_seed_offset=None,
                    rng_name="", training=True, name=None):
    """[b, s, h, d] in/out; returns (out, softmax-or-None) like the reference."""
    from ... import ops
    if return_softmax:
        scale = query.shape[-1] ** -0.5
        s = torch.einsum("bqhd,bkhd->bhqk", query.float(), key.float()) * scale
        if causal:
            m = torch.ones(s.shape[-2:], dtype=torch.bool, device=s.device).tril(s.shape[-1] - s.shape[-2])
            s = s.masked_fill(~m, float("-inf"))
        p = torch.softmax(s, -1)
        return torch.einsum("bhqk,bkhd->bqhd", p, value.float()).to(query.dtype), p.to(query.dtype)
    return ops.flash_attention(query, key, value, dropout, causal, training=training, fixed_seed_offset=fixed_seed_offset), None


def flash_attn_qkvpacked(qkv, dropout=0.0, causal=False, return_softmax=False, *, fixed_seed_offset=None, rng_name="",
                         training=True, name=None):
    # qkv: [b, s, 2 + hq/hk, hk, d] (last two groups are k and v)
    q = qkv[:, :, :-2].reshape(qkv.shape[0], qkv.shape[1], -1, qkv.shape[-1])
    return flash_attention(q, qkv[:, :, -2], qkv[:, :, -1], dropout, causal, return_softmax, fixed_seed_offset=fixed_seed_offset,
                           training=training)


def flash_attn_unpadded(query, key, value, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale, dropout=0.0,
                        causal=False, return_softmax=False, fixed_seed_offset=None, rng_name="", training=True, name=None):
    from ... import ops
    return ops.flash_attn_varlen(query, key, value, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale,
                                 dropout, causal, training, fixed_seed_offset), None


def flash_attn_varlen_qkvpacked(qkv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale, dropout=0.0,
                                causal=False, return_softmax=False, fixed_seed_offset=None, rng_name="", varlen_padded=True,
                                training=True, name=None):
    q = qkv[:, :-2].reshape(qkv.shape[0], -1, qkv.shape[-1])
    return flash_attn_unpadded(q, qkv[:, -2], qkv[:, -1], cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, scale,
                               dropout, causal, fixed_seed_offset=fixed_seed_offset, training=training)


def scaled_dot_product_attention(query, key, value, attn_mask=None, dropout_p=0.0, is_causal=False, training=True,
                                 name=None):
    from ... import ops
    return ops.flash_attention(query, key, value, dropout_p, is_causal, attn_mask=attn_mask, training=training)


def flash_attention_with_sparse_mask(query, key, value, attn_mask_start_row_indices, attn_mask_start_row=0,
                                     dropout_p=0.0, is_causal=False, return_softmax=False, return_softmax_lse=False,
                                     return_seed_offset=False, training=True, name=None):
    """Causal attention where, for key column j, rows >= start_row_indices[b,h,j] are additionally masked."""
    from ... import ops
    b, sq, h, d = query.shape
    sk = key.shape[1]
    rows = torch.arange(sq, device=query.device)[None, None, :, None]
    start = attn_mask_start_row_indices[:, :, None, :].to(rows.dtype)
    allowed = rows < start
    if is_causal:
        allowed = allowed & (rows >= torch.arange(sk, device=query.device)[None, None, None, :])
    return ops.flash_attention(query, key, value, dropout_p, False, attn_mask=allowed, training=training)


# ``paddle.nn.functional.flash_attention`` is a MODULE in the reference (``...flash_attention.flash_attention(q, k, v)``);
# the package-level name must stay the module, but calling it directly is a convenient shorthand, so make it callable.
import sys as _sys
import types as _types


class _CallableModule(_types.ModuleType):
    def __call__(self, *args, **kwargs):
        return flash_attention(*args, **kwargs)


_sys.modules[__name__].__class__ = _CallableModule
