"""ClipGradForMOEByGlobalNorm (reference: python/paddle/incubate/distributed/models/moe/grad_clip.py): expert
parameters are different on every rank of the moe group, so their squared norm is all-reduced over that group
before it joins the norm of the shared parameters."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.distributed import communication as comm
from paddle_b200.nn.clip import ClipGradBase


class ClipGradForMOEByGlobalNorm(ClipGradBase):
    def __init__(self, clip_norm, is_expert_param_func=None, moe_group=None, group_name="default_moe_group"):
        self.clip_norm = float(clip_norm)
        self.group_name = group_name
        self.moe_group = moe_group
        if moe_group is not None and moe_group.nranks > 1:
            assert is_expert_param_func is not None, "When moe group size > 1, a function for selecting expert params must be specified."
        self.is_expert_param_func = is_expert_param_func

    @torch.no_grad()
    def _dygraph_clip(self, params_grads):
        normal, expert = [], []
        for p, g in params_grads:
            if g is None or getattr(p, "need_clip", True) is False:
                continue
            is_e = self.moe_group is not None and self.moe_group.nranks > 1 and self.is_expert_param_func(p)
            (expert if is_e else normal).append(g)
        some = next((g for _, g in params_grads if g is not None), None)
        if some is None:
            return params_grads
        dev = some.device
        sq = ops.squared_l2_norm(normal).to(dev).float() if normal else torch.zeros(1, device=dev)
        if expert:
            esq = ops.squared_l2_norm(expert).to(dev).float()
            comm.all_reduce(esq, group=self.moe_group)
            sq = sq + esq
        gn = torch.sqrt(sq)
        mx = torch.full_like(gn, self.clip_norm)
        coef = mx / torch.maximum(gn, mx)
        for p, g in params_grads:
            if g is not None and getattr(p, "need_clip", True):
                g.mul_(coef.to(g.dtype))
        return params_grads
