"""Fused rotary position embedding (K20).  Reference API:
python/paddle/incubate/nn/functional/fused_rotary_position_embedding.py (q,k,v [b,s,h,d]; sin/cos [1,s,1,d];
``use_neox_rotary_style=True`` rotates adjacent pairs, ``False`` rotates halves - fused_rope_kernel.cu:188-190)."""
from __future__ import annotations

import torch

from ._util import K, check, dtype_code, ptr, stream_ptr, use_native


def _rope_ref(x, sin, cos, pos_ids, pairs, time_major, base):
    if x is None:
        return None
    xs = x.transpose(0, 1) if time_major else x
    b, s, h, d = xs.shape
    if sin is None:
        if pairs:
            inv = base ** (-torch.arange(0, d, 2, device=x.device, dtype=torch.float32) / d)
            ang = torch.arange(s, device=x.device, dtype=torch.float32)[:, None] * inv[None]
            ang = ang.repeat_interleave(2, dim=-1)
        else:
            inv = base ** (-2.0 * torch.arange(0, d // 2, device=x.device, dtype=torch.float32) / d)
            ang = torch.arange(s, device=x.device, dtype=torch.float32)[:, None] * inv[None]
            ang = torch.cat([ang, ang], -1)
        sin_t, cos_t = ang.sin(), ang.cos()
    else:
        sin_t, cos_t = sin.reshape(-1, d).float(), cos.reshape(-1, d).float()
    if pos_ids is not None:
        sn, cs = sin_t[pos_ids.long()][:, :, None, :], cos_t[pos_ids.long()][:, :, None, :]
    else:
        sn, cs = sin_t[None, :s, None, :], cos_t[None, :s, None, :]
    xf = xs.float()
    if pairs:
        x0, x1 = xf[..., 0::2], xf[..., 1::2]
        rot = torch.stack([-x1, x0], -1).reshape(xf.shape)
    else:
        rot = torch.cat([-xf[..., d // 2:], xf[..., :d // 2]], -1)
    out = (xf * cs + rot * sn).to(x.dtype)
    return out.transpose(0, 1) if time_major else out


class _RopeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, sin, cos, pos_ids, pairs, time_major, base):
        outs = _RopeFn._run(q, k, v, sin, cos, pos_ids, pairs, time_major, base, 1.0)
        ctx.save_for_backward(sin, cos, pos_ids)
        ctx.cfg = (pairs, time_major, base, q is not None, k is not None, v is not None)
        return outs

    @staticmethod
    def _run(q, k, v, sin, cos, pos_ids, pairs, time_major, base, sign):
        ref = next(t for t in (q, k, v) if t is not None)
        ts = [t.contiguous() if t is not None else None for t in (q, k, v)]
        outs = [torch.empty_like(t) if t is not None else None for t in ts]
        if time_major:
            s, b = ref.shape[0], ref.shape[1]
        else:
            b, s = ref.shape[0], ref.shape[1]
        d = ref.shape[-1]
        hs = [t.shape[2] if t is not None else 0 for t in ts]
        check(K().pdb200_rope(ptr(ts[0]), ptr(ts[1]), ptr(ts[2]), ptr(outs[0]), ptr(outs[1]), ptr(outs[2]),
                              hs[0], hs[1], hs[2], ptr(sin), ptr(cos), ptr(pos_ids), b, s, d, int(not pairs),
                              int(time_major), float(base), float(sign), dtype_code(ref), stream_ptr(ref.device)), "rope")
        return tuple(outs)

    @staticmethod
    def backward(ctx, dq, dk, dv):
        sin, cos, pos_ids = ctx.saved_tensors
        pairs, time_major, base, hq, hk, hv = ctx.cfg
        g = _RopeFn._run(dq if hq else None, dk if hk else None, dv if hv else None, sin, cos, pos_ids, pairs,
                         time_major, base, -1.0)
        return g[0], g[1], g[2], None, None, None, None, None, None


def fused_rope(q, k=None, v=None, sin=None, cos=None, position_ids=None, use_neox_rotary_style=True,
               time_major=False, rotary_emb_base=10000.0):
    pairs = bool(use_neox_rotary_style)
    if use_native(q):
        d = q.shape[-1]
        sin_t = sin.reshape(-1, d).float().contiguous() if sin is not None else None
        cos_t = cos.reshape(-1, d).float().contiguous() if cos is not None else None
        pid = position_ids.long().contiguous() if position_ids is not None else None
        return _RopeFn.apply(q, k, v, sin_t, cos_t, pid, pairs, time_major, rotary_emb_base)
    return tuple(_rope_ref(t, sin, cos, position_ids, pairs, time_major, rotary_emb_base) for t in (q, k, v))


class _QKVRopeFn(torch.autograd.Function):
    """Fused split of the packed QKV projection + rotary embedding on q/k (K20b).  Backward merges dq/dk/dv back
    into one packed gradient with the inverse rotation - one pass each way, no strided copies, no zero/add kernels."""

    @staticmethod
    def forward(ctx, qkv, nq, nkv, sin, cos, pos_ids, pairs, base):
        b, s = qkv.shape[0], qkv.shape[1]
        d = qkv.shape[-1] // (nq + 2 * nkv)
        qkv = qkv.contiguous()
        q = torch.empty(b, s, nq, d, dtype=qkv.dtype, device=qkv.device)
        k = torch.empty(b, s, nkv, d, dtype=qkv.dtype, device=qkv.device)
        v = torch.empty(b, s, nkv, d, dtype=qkv.dtype, device=qkv.device)
        check(K().pdb200_qkv_rope(ptr(qkv), ptr(q), ptr(k), ptr(v), nq, nkv, ptr(sin), ptr(cos), ptr(pos_ids), b, s, d,
                                  int(not pairs), float(base), 1.0, 1, dtype_code(qkv), stream_ptr(qkv.device)), "qkv_rope")
        ctx.save_for_backward(sin, cos, pos_ids)
        ctx.cfg = (nq, nkv, pairs, base, d)
        return q, k, v

    @staticmethod
    def backward(ctx, dq, dk, dv):
        sin, cos, pos_ids = ctx.saved_tensors
        nq, nkv, pairs, base, d = ctx.cfg
        dq, dk, dv = dq.contiguous(), dk.contiguous(), dv.contiguous()
        b, s = dq.shape[0], dq.shape[1]
        dqkv = torch.empty(b, s, (nq + 2 * nkv) * d, dtype=dq.dtype, device=dq.device)
        check(K().pdb200_qkv_rope(ptr(dqkv), ptr(dq), ptr(dk), ptr(dv), nq, nkv, ptr(sin), ptr(cos), ptr(pos_ids), b, s, d,
                                  int(not pairs), float(base), -1.0, 0, dtype_code(dq), stream_ptr(dq.device)), "qkv_rope(bwd)")
        return dqkv, None, None, None, None, None, None, None


def qkv_split_rope(qkv, num_q_heads, num_kv_heads, sin=None, cos=None, position_ids=None, use_neox_rotary_style=True,
                   rotary_emb_base=10000.0):
    """qkv: [b, s, (hq + 2 hk) * d] packed projection output -> (q [b,s,hq,d], k, v [b,s,hk,d]) with q/k rotated."""
    if hasattr(qkv, "placements"):
        from ..distributed.auto_parallel import dtensor as D
        return D.run_rule("qkv_split_rope", lambda a: qkv_split_rope(a, num_q_heads, num_kv_heads, sin, cos, position_ids,
                                                                     use_neox_rotary_style, rotary_emb_base),
                          [qkv], {"num_q_heads": num_q_heads, "num_kv_heads": num_kv_heads})
    pairs = bool(use_neox_rotary_style)
    d = qkv.shape[-1] // (num_q_heads + 2 * num_kv_heads)
    if use_native(qkv):
        sin_t = sin.reshape(-1, d).float().contiguous() if sin is not None else None
        cos_t = cos.reshape(-1, d).float().contiguous() if cos is not None else None
        pid = position_ids.long().contiguous() if position_ids is not None else None
        return _QKVRopeFn.apply(qkv, num_q_heads, num_kv_heads, sin_t, cos_t, pid, pairs, rotary_emb_base)
    b, s = qkv.shape[0], qkv.shape[1]
    x = qkv.reshape(b, s, num_q_heads + 2 * num_kv_heads, d)
    q, k, v = x[:, :, :num_q_heads], x[:, :, num_q_heads:num_q_heads + num_kv_heads], x[:, :, num_q_heads + num_kv_heads:]
    q, k, _ = fused_rope(q, k, None, sin, cos, position_ids, use_neox_rotary_style, False, rotary_emb_base)
    return q, k, v
