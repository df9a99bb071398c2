"""Fused softmax-cross-entropy (K8 local part).  Reference: python/paddle/nn/functional/loss.py cross_entropy,
c_softmax_with_cross_entropy (fleet/layers/mpu/mp_ops.py:406-471)."""
from __future__ import annotations

import torch
import torch.nn.functional as TF

from ._util import K, check, dtype_code, ptr, stream_ptr, use_native


class _SoftmaxCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        V = logits.shape[-1]
        lg = logits if logits.stride(-1) == 1 and logits.is_contiguous() else logits.contiguous()
        rows = lg.numel() // V
        lab = labels.reshape(-1).long().contiguous()
        loss = torch.empty(rows, dtype=torch.float32, device=lg.device)
        lse = torch.empty(rows, dtype=torch.float32, device=lg.device)
        check(K().pdb200_ce_fwd(ptr(lg), ptr(lab), None, ptr(loss), ptr(lse), rows, V, V, 0, int(ignore_index),
                                dtype_code(lg), stream_ptr(lg.device)), "ce_fwd")
        ctx.save_for_backward(lg, lab, lse)
        ctx.ignore_index = int(ignore_index)
        return loss.reshape(labels.shape)

    @staticmethod
    def backward(ctx, dloss):
        lg, lab, lse = ctx.saved_tensors
        V = lg.shape[-1]
        rows = lg.numel() // V
        dl = dloss.reshape(-1).float().contiguous()
        dlogits = torch.empty_like(lg)
        check(K().pdb200_ce_bwd(ptr(lg), ptr(lab), ptr(lse), ptr(dl), 0.0, ptr(dlogits), rows, V, V, 0,
                                ctx.ignore_index, dtype_code(lg), stream_ptr(lg.device)), "ce_bwd")
        return dlogits, None, None


def softmax_cross_entropy(logits, labels, ignore_index=-100):
    """Per-row loss (fp32), shape = labels.shape; rows whose label == ignore_index give 0."""
    if hasattr(logits, "placements") or hasattr(labels, "placements"):
        from ..distributed.auto_parallel import dtensor as D
        V = logits.shape[-1]
        flat = D._TABLE[TF.cross_entropy](logits.reshape(-1, V), labels.reshape(-1), ignore_index=ignore_index, reduction="none")
        return flat.reshape(*labels._global_shape) if hasattr(labels, "_global_shape") else flat.reshape(labels.shape)
    if use_native(logits):
        return _SoftmaxCEFn.apply(logits, labels, ignore_index)
    V = logits.shape[-1]
    return TF.cross_entropy(logits.reshape(-1, V).float(), labels.reshape(-1).long(), ignore_index=ignore_index,
                            reduction="none").reshape(labels.shape)


def ce_local_stats(logits, labels, vocab_start, ignore_index=-100):
    """Per-rank half of the vocab-parallel loss: [rows, 3] fp32 = (max, sum exp(x-max), target logit or 0)."""
    V = logits.shape[-1]
    lg = logits.contiguous()
    rows = lg.numel() // V
    lab = labels.reshape(-1).long().contiguous()
    if use_native(lg):
        stats = torch.empty(rows, 3, dtype=torch.float32, device=lg.device)
        check(K().pdb200_ce_fwd(ptr(lg), ptr(lab), ptr(stats), None, None, rows, V, V, int(vocab_start),
                                int(ignore_index), dtype_code(lg), stream_ptr(lg.device)), "ce_fwd(stats)")
        return stats
    x = lg.reshape(rows, V).float()
    m = x.max(-1).values
    s = torch.exp(x - m[:, None]).sum(-1)
    local = lab - vocab_start
    ok = (lab != ignore_index) & (local >= 0) & (local < V)
    tgt = torch.where(ok, x.gather(1, local.clamp(0, V - 1)[:, None])[:, 0], torch.zeros_like(m))
    return torch.stack([m, s, tgt], -1)


def ce_combine(all_stats, labels, ignore_index=-100):
    """all_stats [nranks, rows, 3] -> (loss[rows], lse[rows])."""
    nranks, rows, _ = all_stats.shape
    lab = labels.reshape(-1).long().contiguous()
    if use_native(all_stats):
        st = all_stats.contiguous()
        loss = torch.empty(rows, dtype=torch.float32, device=st.device)
        lse = torch.empty(rows, dtype=torch.float32, device=st.device)
        check(K().pdb200_ce_combine(ptr(st), ptr(lab), ptr(loss), ptr(lse), rows, nranks, int(ignore_index),
                                    stream_ptr(st.device)), "ce_combine")
        return loss, lse
    gm = all_stats[..., 0].max(0).values
    gs = (all_stats[..., 1] * torch.exp(all_stats[..., 0] - gm[None])).sum(0)
    tgt = all_stats[..., 2].sum(0)
    lse = gm + torch.log(gs)
    loss = torch.where(lab == ignore_index, torch.zeros_like(lse), lse - tgt)
    return loss, lse


def ce_local_backward(logits, labels, lse, dloss, vocab_start, ignore_index=-100):
    """dlogits for this rank's vocab shard given the global lse."""
    V = logits.shape[-1]
    lg = logits.contiguous()
    rows = lg.numel() // V
    lab = labels.reshape(-1).long().contiguous()
    dl = dloss.reshape(-1).float().contiguous()
    if use_native(lg):
        out = torch.empty_like(lg)
        check(K().pdb200_ce_bwd(ptr(lg), ptr(lab), ptr(lse), ptr(dl), 0.0, ptr(out), rows, V, V, int(vocab_start),
                                int(ignore_index), dtype_code(lg), stream_ptr(lg.device)), "ce_bwd")
        return out
    x = lg.reshape(rows, V).float()
    p = torch.exp(x - lse[:, None])
    local = lab - vocab_start
    ok = (local >= 0) & (local < V)
    p[torch.arange(rows, device=x.device)[ok], local[ok]] -= 1.0
    g = torch.where(lab == ignore_index, torch.zeros_like(dl), dl)
    return (p * g[:, None]).to(lg.dtype).reshape(lg.shape)
