"""GroupShardedOptimizerStage2 (reference: .../sharding/group_sharded_optimizer_stage2.py:53-700): ZeRO-2 optimizer
side - parameters are assigned to ranks greedily by size per dtype (:347-366), each rank updates only the ones it
owns (optimizer state exists only there), then owners broadcast the updated parameters (:630-644).  Every owner's
parameters are packed per dtype into a flat ``ParamStorage`` (reference :368-409), so that exchange is one broadcast
per (owner, dtype); ``offload=True`` keeps fp32 master copies and moments of the owned parameters in pinned host
memory and steps on the CPU (reference :120-160, 580-620)."""
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
        self.offload = bool(offload)
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
        self._param_storages = {}      # (owner index, dtype) -> ParamStorage
        self._model = None
        self._build_param_storages()
        if self.offload:
            from .group_sharded_utils import offload_optimizer_state
            offload_optimizer_state(optim, [p for p in local if p.requires_grad], wrap_step=False)

    def _attach_model(self, model):
        self._model = model

    @torch.no_grad()
    def _build_param_storages(self):
        """Pack every owner's trainable parameters into flat per-dtype buffers (all ranks build all of them: stage 2
        keeps full parameter replicas).  Views of non-contiguous / already-viewed parameters are left alone."""
        from .group_sharded_storage import ParamStorage, _aligned_numel
        for r, ps in enumerate(self.__segment_params):
            by_dtype = {}
            for p in ps:
                if p.requires_grad and p.is_contiguous() and not p.data._is_view():
                    by_dtype.setdefault(p.dtype, []).append(p)
            for dtype, group_ps in by_dtype.items():
                size = sum(_aligned_numel(p) for p in group_ps)
                st = ParamStorage(size, dtype, group_ps[0].device)
                st.add_rank_params(group_ps)
                self._param_storages[(r, dtype)] = st

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
        if getattr(self._optim, "_offload_pre", None) is not None:
            self._optim._offload_pre()
            self._optim.step()
            self._optim._offload_post()
        else:
            self._optim.step()
        self._broadcast_params()

    def minimize(self, *a, **k):
        self.step()

    @torch.no_grad()
    def _broadcast_params(self):
        tasks, packed = [], set()
        for (r, _dtype), st in self._param_storages.items():
            tasks.append(comm.broadcast(st.buffer, self._group.ranks[r], self._group, sync_op=False))
            packed.update(id(p) for p in st._params)
        for r, ps in enumerate(self.__segment_params):
            for p in ps:
                if id(p) not in packed:
                    tasks.append(comm.broadcast(p, self._group.ranks[r], self._group, sync_op=False))
        for t in tasks:
            t.wait()

    def clear_grad(self, set_to_zero=True):
        if self._model is not None:
            return self._model._clear_grads(set_to_zero)
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
