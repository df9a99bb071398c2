"""GroupShardedStage2 (reference: .../sharding/group_sharded_stage2.py:46-700): ZeRO-2 model side - as backward
produces gradients they are reduced (AVG) to the rank that owns the parameter; non-owners drop them.

Small gradients travel in per-destination fused buckets of up to ``buffer_max_size`` elements (``GradStorage``,
reference :411-527): ``param.grad`` is a view of its bucket, autograd accumulates in place, and when the last
gradient of a bucket has been produced the whole bucket leaves in one asynchronous ``reduce``.  Parameters larger than
a bucket are reduced on their own.  ``no_sync()`` accumulates locally; the next synchronised backward reduces the sums.
"""
from __future__ import annotations

import contextlib

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.nn.layer.layers import Layer

from .group_sharded_storage import GradStorage, _aligned_numel


class GroupShardedStage2(Layer):
    def __init__(self, layer, sharding_optimizer, group=None, sync_buffers=False, buffer_max_size=2 ** 23,
                 auto_refresh_trainable=True, device="gpu", dp_group=None):
        super().__init__()
        self._layer = layer
        self._sharding_optimizers = sharding_optimizer if isinstance(sharding_optimizer, list) else [sharding_optimizer]
        opt = self._sharding_optimizers[0]
        self._group = group if group is not None else opt._group
        self._dp_group = dp_group
        self._rank = self._group.rank
        self._world = self._group.nranks
        self._sync_buffers = sync_buffers
        self._buffer_max_size = int(buffer_max_size)
        self._param2rank = opt.param2rank
        self._trainable = [p for p in layer.parameters() if p.requires_grad]
        self._tasks = []
        self._callback_queued = False
        self._need_sync = True
        self._grad_storages = {}       # dtype -> {destination rank index: GradStorage}
        self._grad_storage_list = []
        self._setup_grad_storages()
        for p in self._trainable:
            p.register_post_accumulate_grad_hook(self._make_hook(p))
        for o in self._sharding_optimizers:
            o._attach_model(self)

    # ------------------------------------------------------------------ buckets
    def _setup_grad_storages(self):
        """One bucket per (dtype, owner): its size is what the owner's small parameters need, capped at
        ``buffer_max_size`` elements; parameters that do not fit keep an individual gradient."""
        need = {}
        for p in self._trainable:
            if _aligned_numel(p) <= self._buffer_max_size:
                key = (p.dtype, self._param2rank[p.name])
                need[key] = min(need.get(key, 0) + _aligned_numel(p), self._buffer_max_size)
        for (dtype, dst), size in need.items():
            st = GradStorage(size, dtype, self._trainable[0].device, dst)
            self._grad_storages.setdefault(dtype, {})[dst] = st
            self._grad_storage_list.append(st)
        for p in self._trainable:
            st = self._grad_storages.get(p.dtype, {}).get(self._param2rank[p.name])
            p._gs_grad_bucket = None
            if st is not None and st.can_add_grad_view(p):
                st.add_grad(p)

    def _rebuild_views(self):
        for st in self._grad_storage_list:
            st.rebuild()

    # ------------------------------------------------------------------ backward
    def _make_hook(self, p):
        owner_idx = self._param2rank[p.name]
        owner = self._group.ranks[owner_idx]

        @torch.no_grad()
        def hook(param):
            if not self._need_sync or param.grad is None:
                return
            if not self._callback_queued:
                self._callback_queued = True
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
            st = getattr(param, "_gs_grad_bucket", None)
            if st is not None and param.grad.data_ptr() >= st.buffer.data_ptr() and \
                    param.grad.data_ptr() < st.buffer.data_ptr() + st.buffer.numel() * st.buffer.element_size():
                st.params_checked_in += 1
                if st.all_checked_in and not st.sent:
                    st.sent = True
                    t = comm.reduce(st.buffer, owner, ReduceOp.AVG, self._group, sync_op=False)
                    self._tasks.append((st, owner_idx, t))
                return
            t = comm.reduce(param.grad, owner, ReduceOp.AVG, self._group, sync_op=False)
            self._tasks.append((param, owner_idx, t))
        return hook

    @torch.no_grad()
    def _finalize(self):
        # buckets whose parameters did not all produce a gradient (unused parameters) still have to leave
        for st in self._grad_storage_list:
            if self._need_sync and not st.sent and st.params_checked_in > 0:
                st.sent = True
                owner = self._group.ranks[st.destination]
                self._tasks.append((st, st.destination, comm.reduce(st.buffer, owner, ReduceOp.AVG, self._group, sync_op=False)))
        dp = self._dp_group is not None and self._dp_group.nranks > 1
        for obj, owner_idx, t in self._tasks:
            t.wait()
            mine = owner_idx == self._rank
            if isinstance(obj, GradStorage):
                if mine and dp:
                    comm.all_reduce(obj.buffer, ReduceOp.AVG, self._dp_group)
                if not mine:
                    obj.zero_()              # a send buffer only: ready for the next accumulation
            elif not mine:
                obj.grad = None
            elif dp:
                comm.all_reduce(obj.grad, ReduceOp.AVG, self._dp_group)
        for st in self._grad_storage_list:
            st.reset_checked_in()
        self._tasks = []
        self._callback_queued = False

    def forward(self, *inputs, **kwargs):
        if self._sync_buffers:
            with torch.no_grad():
                for b in self._layer.buffers():
                    comm.broadcast(b, self._group.ranks[0], self._group)
        self._rebuild_views()
        return self._layer(*inputs, **kwargs)

    def no_sync(self):
        @contextlib.contextmanager
        def ctx():
            prev = self._need_sync
            self._need_sync = False
            try:
                yield
            finally:
                self._need_sync = prev
        return ctx()

    def state_dict(self, *a, **k):
        return self._layer.state_dict(*a, **k)

    def set_state_dict(self, sd, use_structured_name=True):
        return self._layer.set_state_dict(sd, use_structured_name)

    def parameters(self, *a, **k):
        return self._layer.parameters(*a, **k)

    def named_parameters(self, *a, **k):
        return self._layer.named_parameters(*a, **k)

    def clear_gradients(self):
        self._clear_grads(set_to_zero=True)

    @torch.no_grad()
    def _clear_grads(self, set_to_zero=True):
        """One fill per bucket instead of one per parameter; unbucketed gradients are zeroed / dropped individually."""
        for st in self._grad_storage_list:
            st.zero_()
            st.rebuild()
        for p in self._trainable:
            if getattr(p, "_gs_grad_bucket", None) is None and p.grad is not None:
                if set_to_zero:
                    p.grad.zero_()
                else:
                    p.grad = None

    def to(self, device=None, dtype=None, blocking=True):
        self._layer.to(device=device, dtype=dtype)
        for st in self._grad_storage_list:
            st.to(next(self._layer.parameters()).device)
        return self
