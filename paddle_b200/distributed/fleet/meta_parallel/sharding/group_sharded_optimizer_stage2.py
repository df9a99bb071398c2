"""GroupShardedOptimizerStage2 (reference: .../sharding/group_sharded_optimizer_stage2.py:53-700): ZeRO-2 optimizer
side - parameters are assigned to ranks greedily by size per dtype (:347-366), each rank updates only the ones it
owns (optimizer state exists only there), then owners broadcast the updated parameters (:630-644)."""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.nn.clip import ClipGradByGlobalNorm

from .group_sharded_utils import GroupShardedClipGrad


class GroupShardedOptimizerStage2:
    def __init__(self, params, optim, group=None, offload=False, device="gpu", pertrain_sync_models=True, dp_group=None,
                 **kw):
        from paddle_b200.distributed import collective as C
        self._optim = optim
        self._group = group if group is not None else C.new_group(list(range(C.get_world_size())))
        self._dp_group = dp_group
        self.world_size = self._group.nranks
        self._rank = self._group.rank
        self._global_root_rank = self._group.ranks[0]
        if offload:
            raise NotImplementedError("offload=True is not implemented in this build")
        self._local_params = list(params)
        self._all_params = list(params)
        self.__segment_params = None
        self._param2rank = None
        self._segment()
        if pertrain_sync_models:
            self._sync_params_and_buffers()
        local = self.local_params
        base = dict(optim._param_groups[0]) if getattr(optim, "_param_groups", None) else {}
        base["params"] = local
        optim._parameter_list = local
        optim._param_groups = [base]
        optim._group = self._group
        if isinstance(optim._grad_clip, ClipGradByGlobalNorm):
            optim._grad_clip = GroupShardedClipGrad(optim._grad_clip, None, self._group)

    def _segment(self):
        segs = [[] for _ in range(self.world_size)]
        sizes = [0] * self.world_size
        for p in self._all_params:
            r = sizes.index(min(sizes))
            segs[r].append(p)
            if p.requires_grad:
                sizes[r] += p.numel()
        self.__segment_params = segs
        self._param2rank = {p.name: r for r, ps in enumerate(segs) for p in ps}

    def _segment_params(self):
        return self.__segment_params

    @property
    def local_params(self):
        return self.__segment_params[self._rank]

    @property
    def param2rank(self):
        return self._param2rank

    @torch.no_grad()
    def _sync_params_and_buffers(self):
        for p in self._all_params:
            comm.broadcast(p, self._global_root_rank, self._group)

    @torch.no_grad()
    def step(self):
        self._optim.step()
        self._broadcast_params()

    def minimize(self, *a, **k):
        self.step()

    @torch.no_grad()
    def _broadcast_params(self):
        tasks = []
        for r, ps in enumerate(self.__segment_params):
            for p in ps:
                tasks.append(comm.broadcast(p, self._group.ranks[r], self._group, sync_op=False))
        for t in tasks:
            t.wait()

    def clear_grad(self, set_to_zero=True):
        for p in self._all_params:
            if p.grad is not None:
                if set_to_zero:
                    p.grad.zero_()
                else:
                    p.grad = None

    clear_gradients = clear_grad

    def state_dict(self):
        return self._optim.state_dict()

    def set_state_dict(self, sd):
        return self._optim.set_state_dict(sd)

    def __getattr__(self, item):
        return getattr(self.__dict__["_optim"], item)
