"""GroupSharded helpers (reference: .../sharding/group_sharded_utils.py): sharded global-norm clip and the scaler
wrapper whose found_inf is all-reduced (MAX) over the sharding group."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp


class GroupShardedClipGrad:
    def __init__(self, clip, device, group):
        self._clip = clip
        self._device = device
        self._group = group

    @torch.no_grad()
    def _dygraph_clip(self, params_grads):
        """params_grads hold only this rank's slices/owned params except ``unslice`` ones (replicated)."""
        sliced, unsliced = [], []
        for p, g in params_grads:
            if g is None or getattr(p, "need_clip", True) is False:
                continue
            (unsliced if getattr(p, "_gs_unslice", False) else sliced).append(g)
        some = next((g for _, g in params_grads if g is not None), None)
        if some is None:
            return params_grads
        dev = some.device
        sq = ops.squared_l2_norm(sliced).to(dev).float() if sliced else torch.zeros(1, device=dev)
        comm.all_reduce(sq, ReduceOp.SUM, self._group)
        if unsliced:
            sq = sq + ops.squared_l2_norm(unsliced).to(dev).float()
        gn = torch.sqrt(sq)
        mx = torch.full_like(gn, self._clip.clip_norm)
        coef = mx / torch.maximum(gn, mx)
        for p, g in params_grads:
            if g is not None and getattr(p, "need_clip", True):
                g.mul_(coef.to(g.dtype))
        return params_grads

    def __call__(self, params_grads):
        return self._dygraph_clip(params_grads)

    def __getattr__(self, item):
        return getattr(self.__dict__["_clip"], item)


def GroupShardedScaler(scaler):
    def _sync(optimizer):
        group = getattr(optimizer, "_group", None)
        if scaler._found_inf is not None and group is not None and group.nranks > 1:
            comm.all_reduce(scaler._found_inf, ReduceOp.MAX, group)
    scaler._sync_found_inf = _sync
    return scaler
