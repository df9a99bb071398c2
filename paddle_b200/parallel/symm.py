"""Symmetric-memory communication layer.

Baseline implementations (NCCL through ``paddle.distributed``) live here behind the same call signatures the fused
NVLink kernels use, so every caller is written once and ``FLAGS_use_fused_comm`` switches the transport.
"""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.framework.flags import flag


def all_gather_small(t, group):
    """[...]-> [nranks, ...] for small payloads (vocab-parallel CE statistics)."""
    n = group.nranks
    out = torch.empty((n,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    from . import symm_fused
    if flag("use_fused_comm") and symm_fused.available(group, t):
        return symm_fused.all_gather_small(t, group, out)
    comm.all_gather(out, t.contiguous(), group)
    return out


def all_gather_seq(x, group):
    """[s/n, ...] -> [s, ...] along dim 0 (sequence-parallel all-gather)."""
    n = group.nranks
    out = torch.empty((x.shape[0] * n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    comm.all_gather(out, x.contiguous(), group)
    return out
