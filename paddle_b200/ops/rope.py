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
