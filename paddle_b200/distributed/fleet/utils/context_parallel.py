"""Long-context attention over the ``sep`` axis: Ulysses, all-gather-KV and ring attention.

The reference only provides the ``sep`` group, the parameter broadcast and the fused dp x sep gradient all-reduce;
the model has to split / concatenate the sequence itself (python/paddle/distributed/fleet/base/topology.py:223-247,
meta_parallel/segment_parallel.py:26-40, test/collective/fleet/hybrid_parallel_sep_model.py:33-76).  These are the
net-new pieces SURVEY.md 5.7 asks for.  Every rank holds a contiguous sequence shard ``[b, s/n, h, d]``:

* ``ulysses_attention``       all-to-all seq-shard -> head-shard, full-sequence attention on h/n heads, all-to-all back.
* ``allgather_kv_attention``  K/V of the ranks up to mine are all-gathered; causal attention with the key offset
                              (rank r sees keys [0, (r+1) s/n)), the flash kernel's ``sk > sq`` causal alignment.
* ``ring_flash_attention``    K/V blocks travel around the ring (NCCL p2p), each step runs the flash kernel on one
                              block and the partial results are merged with their log-sum-exp; the backward pass sends
                              dK/dV along with the blocks and reuses the global LSE, so no block is ever re-normalised.
All three are differentiable and are validated against single-device attention in tests (CPU: gloo, torch math)."""
from __future__ import annotations

import math

import torch

from paddle_b200.distributed import communication as comm

__all__ = ["ulysses_attention", "allgather_kv_attention", "ring_flash_attention", "split_sequence", "gather_sequence"]


def _attn(q, k, v, causal, scale):
    from paddle_b200 import ops
    return ops.flash_attention(q, k, v, causal=causal, softmax_scale=scale)


# ------------------------------------------------------------------------------------------------ helpers
class _SplitSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group, dim):
        ctx.group, ctx.dim = group, dim
        return x.chunk(group.nranks, dim)[group.rank].contiguous()

    @staticmethod
    def backward(ctx, g):
        parts = []
        comm.all_gather(parts, g.contiguous(), ctx.group)
        return torch.cat(parts, ctx.dim), None, None


class _GatherSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group, dim):
        ctx.group, ctx.dim = group, dim
        parts = []
        comm.all_gather(parts, x.contiguous(), group)
        return torch.cat(parts, dim)

    @staticmethod
    def backward(ctx, g):
        return g.chunk(ctx.group.nranks, ctx.dim)[ctx.group.rank].contiguous(), None, None


def split_sequence(x, group, dim=1):
    """Keep this rank's contiguous shard of ``dim`` (backward all-gathers the gradient)."""
    return x if group is None or group.nranks == 1 else _SplitSeq.apply(x, group, dim)


def gather_sequence(x, group, dim=1):
    """Concatenate every rank's shard along ``dim`` (backward keeps the local slice)."""
    return x if group is None or group.nranks == 1 else _GatherSeq.apply(x, group, dim)


# ------------------------------------------------------------------------------------------------ Ulysses
class _SeqToHead(torch.autograd.Function):
    """[b, s/n, h, d] -> [b, s, h/n, d] (forward) and back (backward): one all-to-all each way."""

    @staticmethod
    def forward(ctx, x, group, inverse):
        ctx.group, ctx.inverse = group, inverse
        return _a2a(x, group, inverse)

    @staticmethod
    def backward(ctx, g):
        return _a2a(g.contiguous(), ctx.group, not ctx.inverse), None, None


def _a2a(x, group, inverse):
    n = group.nranks
    b, s, h, d = x.shape
    if not inverse:          # scatter heads, gather sequence
        assert h % n == 0, f"Ulysses needs heads ({h}) divisible by the sep degree ({n})"
        send = x.reshape(b, s, n, h // n, d).permute(2, 0, 1, 3, 4).contiguous()       # [n, b, s_loc, h/n, d]
        recv = torch.empty_like(send)
        comm.alltoall_single(recv, send, group=group)
        return recv.permute(1, 0, 2, 3, 4).reshape(b, n * s, h // n, d)               # sequence blocks in rank order
    assert s % n == 0
    send = x.reshape(b, n, s // n, h, d).permute(1, 0, 2, 3, 4).contiguous()            # [n, b, s_loc, h_loc, d]
    recv = torch.empty_like(send)
    comm.alltoall_single(recv, send, group=group)
    return recv.permute(1, 2, 0, 3, 4).reshape(b, s // n, n * h, d)


def ulysses_attention(q, k, v, group, causal=True, softmax_scale=None):
    """DeepSpeed-Ulysses: q, k, v [b, s/n, h, d] sequence shards -> attention output [b, s/n, h, d]."""
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if group is None or group.nranks == 1:
        return _attn(q, k, v, causal, scale)
    qh, kh, vh = (_SeqToHead.apply(t, group, False) for t in (q, k, v))
    oh = _attn(qh, kh, vh, causal, scale)
    return _SeqToHead.apply(oh, group, True)


# ------------------------------------------------------------------------------------------------ all-gather KV
class _AllGatherSum(torch.autograd.Function):
    """All-gather along the sequence whose backward SUMS every rank's gradient for my shard (reduce-scatter): each
    rank's attention uses my keys, so their gradients all flow back to me."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        parts = []
        comm.all_gather(parts, x.contiguous(), group)
        return torch.cat(parts, 1)

    @staticmethod
    def backward(ctx, g):
        n = ctx.group.nranks
        chunks = [c.contiguous() for c in g.chunk(n, 1)]
        out = torch.empty_like(chunks[0])
        comm.reduce_scatter(out, chunks, group=ctx.group)
        return out, None


def allgather_kv_attention(q, k, v, group, causal=True, softmax_scale=None):
    """K/V all-gathered along the sequence; causal ranks only attend to keys up to their own shard."""
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if group is None or group.nranks == 1:
        return _attn(q, k, v, causal, scale)
    kf, vf = _AllGatherSum.apply(k, group), _AllGatherSum.apply(v, group)
    if causal:
        end = (group.rank + 1) * k.shape[1]
        kf, vf = kf[:, :end], vf[:, :end]       # keys beyond my shard are masked for every query of mine anyway
    return _attn(q, kf.contiguous(), vf.contiguous(), causal, scale)


# ------------------------------------------------------------------------------------------------ ring attention
def _native_block_ok(q, k):
    from paddle_b200 import lib
    from paddle_b200.ops._util import use_native
    return (q.is_cuda and use_native(q) and q.dtype == torch.bfloat16 and q.shape[-1] in (64, 128) and
            q.shape[1] % 128 == 0 and k.shape[1] % 128 == 0 and hasattr(lib.kernels.require(), "pdb200_flash_attn_bwd"))


def _block_fwd(q, k, v, causal, scale):
    """(o [b,s,h,d] in q.dtype, lse [b,h,s] fp32) of one K/V block; tcgen05 kernel on GPU, torch math elsewhere."""
    if _native_block_ok(q, k):
        from paddle_b200.ops.attention import _FlashFn
        with torch.no_grad():
            return _FlashFn.apply(q, k, v, bool(causal), float(scale))
    g = q.shape[2] // k.shape[2]
    kf, vf = k.float().repeat_interleave(g, 2), v.float().repeat_interleave(g, 2)
    s = torch.einsum("bqhd,bkhd->bhqk", q.float(), kf) * scale
    if causal:
        sq, sk = q.shape[1], k.shape[1]
        s = s.masked_fill(~torch.ones(sq, sk, dtype=torch.bool, device=q.device).tril(sk - sq), float("-inf"))
    lse = torch.logsumexp(s, -1)
    o = torch.einsum("bhqk,bkhd->bqhd", torch.exp(s - lse[..., None]), vf)
    return o.to(q.dtype), lse


def _block_bwd(q, k, v, o, do, lse, causal, scale):
    """Gradient contributions of one K/V block given the GLOBAL output and log-sum-exp of the rows."""
    if _native_block_ok(q, k):
        from paddle_b200 import lib
        from paddle_b200.lib import check
        from paddle_b200.ops._util import dtype_code, ptr, stream_ptr
        b, sq, hq, d = q.shape
        sk, hk = k.shape[1], k.shape[2]
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        nb = lib.kernels.pdb200_flash_attn_bwd_workspace_bytes(b, sq, sk, hq, hk, d)
        ws = torch.empty(max(nb, 16), dtype=torch.uint8, device=q.device)
        check(lib.kernels.pdb200_flash_attn_bwd(ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(dq), ptr(dk), ptr(dv),
                                                ptr(ws), nb, b, sq, sk, hq, hk, d, float(scale), int(causal), dtype_code(q),
                                                stream_ptr(q.device)), "flash_attn_bwd")
        return dq.float(), dk.float(), dv.float()
    g = q.shape[2] // k.shape[2]
    kf, vf = k.float().repeat_interleave(g, 2), v.float().repeat_interleave(g, 2)
    s = torch.einsum("bqhd,bkhd->bhqk", q.float(), kf) * scale
    if causal:
        sq, sk = q.shape[1], k.shape[1]
        s = s.masked_fill(~torch.ones(sq, sk, dtype=torch.bool, device=q.device).tril(sk - sq), float("-inf"))
    p = torch.exp(s - lse[..., None])
    dof = do.float()
    dv = torch.einsum("bhqk,bqhd->bkhd", p, dof)
    dp = torch.einsum("bqhd,bkhd->bhqk", dof, vf)
    delta = (dof * o.float()).sum(-1).permute(0, 2, 1)                  # [b, h, q]
    ds = p * (dp - delta[..., None]) * scale
    dq = torch.einsum("bhqk,bkhd->bqhd", ds, kf)
    dk = torch.einsum("bhqk,bqhd->bkhd", ds, q.float())
    if g > 1:
        b, sk, hq, d = dk.shape
        dk = dk.reshape(b, sk, hq // g, g, d).sum(3)
        dv = dv.reshape(b, sk, hq // g, g, d).sum(3)
    return dq, dk, dv


def _rotate(tensors, group):
    """Send every tensor to the next rank of the ring and receive the previous rank's (batched p2p)."""
    n, r = group.nranks, group.rank
    nxt, prv = group.ranks[(r + 1) % n], group.ranks[(r - 1) % n]
    outs = [torch.empty_like(t) for t in tensors]
    ops = []
    for t, o in zip(tensors, outs):
        ops.append(comm.P2POp(comm.isend, t.contiguous(), nxt, group))
        ops.append(comm.P2POp(comm.irecv, o, prv, group))
    for task in comm.batch_isend_irecv(ops):
        task.wait()
    return outs


class _RingAttn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, group, causal, scale):
        n, r = group.nranks, group.rank
        o_acc, lse_acc = None, None
        kb, vb = k.contiguous(), v.contiguous()
        for t in range(n):
            src = (r - t) % n
            if not (causal and src > r):
                o_i, lse_i = _block_fwd(q, kb, vb, causal and src == r, scale)
                if o_acc is None:
                    o_acc, lse_acc = o_i.float(), lse_i
                else:
                    new = torch.logaddexp(lse_acc, lse_i)
                    wa = torch.exp(lse_acc - new).permute(0, 2, 1)[..., None]
                    wb = torch.exp(lse_i - new).permute(0, 2, 1)[..., None]
                    o_acc = o_acc * wa + o_i.float() * wb
                    lse_acc = new
            if t + 1 < n:
                kb, vb = _rotate([kb, vb], group)
        out = o_acc.to(q.dtype)
        ctx.save_for_backward(q, k, v, out, lse_acc.contiguous())
        ctx.cfg = (group, causal, scale)
        return out

    @staticmethod
    def backward(ctx, do):
        q, k, v, out, lse = ctx.saved_tensors
        group, causal, scale = ctx.cfg
        n, r = group.nranks, group.rank
        do = do.contiguous()
        dq = torch.zeros_like(q, dtype=torch.float32)
        kb, vb = k.contiguous(), v.contiguous()
        dkb, dvb = torch.zeros_like(k, dtype=torch.float32), torch.zeros_like(v, dtype=torch.float32)
        for t in range(n):
            src = (r - t) % n
            if not (causal and src > r):
                dq_i, dk_i, dv_i = _block_bwd(q, kb, vb, out, do, lse, causal and src == r, scale)
                dq += dq_i
                dkb += dk_i
                dvb += dv_i
            if t + 1 < n:
                kb, vb, dkb, dvb = _rotate([kb, vb, dkb, dvb], group)
            else:
                dkb, dvb = _rotate([dkb, dvb], group)       # the accumulators finish the lap at their owner
        return dq.to(q.dtype), dkb.to(k.dtype), dvb.to(v.dtype), None, None, None


def ring_flash_attention(q, k, v, group, causal=True, softmax_scale=None):
    """Ring attention over contiguous sequence shards: q, k, v [b, s/n, h, d] -> [b, s/n, h, d]."""
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if group is None or group.nranks == 1:
        return _attn(q, k, v, causal, scale)
    return _RingAttn.apply(q, k, v, group, bool(causal), float(scale))
