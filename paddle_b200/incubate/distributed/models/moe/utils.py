"""count_by_gate / limit_by_capacity (reference: python/paddle/incubate/distributed/models/moe/utils.py:58-93)."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.distributed import communication as comm


def _alltoall(in_tensor, group=None):
    if group is None or group.nranks == 1:
        return in_tensor.clone()
    out = torch.empty_like(in_tensor)
    comm.alltoall_single(out, in_tensor.contiguous(), group=group)
    return out


def count_by_gate(gate, num_expert, world_size, require_pos=True, group=None):
    total = num_expert * world_size
    with torch.no_grad():
        local_expert_count = ops.number_count(gate, total)
        global_expert_count = _alltoall(local_expert_count, group) if world_size > 1 else local_expert_count
        pos = None
        if require_pos:
            lec_cum = torch.cumsum(local_expert_count, 0)
            pos = ops.assign_pos(gate, lec_cum)
    return pos, local_expert_count, global_expert_count


def limit_by_capacity(topk_idx, num_expert, world_size, capacity, group=None):
    with torch.no_grad():
        cap = torch.full((num_expert,), int(capacity), dtype=torch.int64, device=topk_idx.device)
        _, lec, gec = count_by_gate(topk_idx, num_expert, world_size, require_pos=False, group=group)
        new_gec = ops.limit_by_capacity(gec, cap, world_size)
        new_lec = _alltoall(new_gec, group) if world_size > 1 else new_gec
        topk_idx = ops.prune_gate_by_capacity(topk_idx, new_lec, num_expert, world_size)
    return new_lec, new_gec, topk_idx
