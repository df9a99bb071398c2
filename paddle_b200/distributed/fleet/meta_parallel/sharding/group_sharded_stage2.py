"""GroupShardedStage2 (reference: .../sharding/group_sharded_stage2.py:46-700): ZeRO-2 model side - as backward
produces a gradient it is reduced (AVG) to the rank that owns the parameter; non-owners drop it.  Small gradients
travel in per-destination fused buckets up to ``buffer_max_size`` elements (GradStorage, :411-527)."""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.nn.layer.layers import Layer


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
        self._buffer_max_size = buffer_max_size
        self._param2rank = opt.param2rank
        self._trainable = [p for p in layer.parameters() if p.requires_grad]
        self._tasks = []
        self._callback_queued = False
        self._accumulate = 1
        # per-destination buckets of (param, view) for small grads
        self._buckets = {}
        for p in self._trainable:
            p.register_post_accumulate_grad_hook(self._make_hook(p))

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
            t = comm.reduce(param.grad, owner, ReduceOp.AVG, self._group, sync_op=False)
            self._tasks.append((param, owner_idx, t))
        return hook

    _need_sync = True

    @torch.no_grad()
    def _finalize(self):
        for p, owner_idx, t in self._tasks:
            t.wait()
            if owner_idx != self._rank:
                p.grad = None
            elif self._dp_group is not None and self._dp_group.nranks > 1:
                comm.all_reduce(p.grad, ReduceOp.AVG, self._dp_group)
        self._tasks = []
        self._callback_queued = False

    def forward(self, *inputs, **kwargs):
        if self._sync_buffers:
            with torch.no_grad():
                for b in self._layer.buffers():
                    comm.broadcast(b, self._group.ranks[0], self._group)
        return self._layer(*inputs, **kwargs)

    def no_sync(self):
        import contextlib

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
        for p in self._trainable:
            p.grad = None

    def to(self, device=None, dtype=None, blocking=True):
        self._layer.to(device=device, dtype=dtype)
        return self
