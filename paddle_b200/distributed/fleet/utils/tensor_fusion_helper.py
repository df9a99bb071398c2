"""Tensor fusion (reference: python/paddle/distributed/fleet/utils/tensor_fusion_helper.py:172-910:
assign_group_by_size, flatten_dense_tensors, FusedCommBuffer with ALL_REDUCE / REDUCE / REDUCE_SCATTER acts,
ShardingGradView, fused_parameters).

B200 design: a *flat* layout.  Parameters of one dtype are laid out contiguously (each padded to 256 B) in one
buffer whose length is padded to ``nranks x 256 B``; parameters and gradients become views of the flat param /
grad buffers, rank r owns the contiguous range ``[r*S, (r+1)*S)``.  Reduce-scatter, the sharded optimizer update and
the parameter all-gather therefore each touch ONE contiguous tensor (no per-parameter slices, no concat/split), which
is also the shape the fused NVLink kernels want (K2 rs_adamw_fused, K3 ag_params).
"""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.distributed.reducer import assign_group_by_size as _assign

ALIGN_BYTES = 256


class HOOK_ACTION:
    ALL_REDUCE = 0
    REDUCE = 1
    REDUCE_SCATTER = 2


def assign_group_by_size(parameters, group_size=128 * 1024 * 1024):
    is_sparse = [False] * len(parameters)
    idx_groups = _assign(parameters, is_sparse, [group_size, group_size])
    return {i: [parameters[j] for j in idxs] for i, idxs in enumerate(idx_groups)}


def _align(numel, elsize, mult=1):
    a = ALIGN_BYTES // elsize * mult
    return (numel + a - 1) // a * a


def flatten_dense_tensors(parameters, use_main_grad=False, fuse_param=True, warp_buffer=False, release_grad=False,
                          nranks=1, allocator=None):
    """-> (param_buffer or None, grad_buffer, offsets).  Parameters (and grads) are re-pointed at buffer views."""
    dtype = parameters[0].dtype
    dev = parameters[0].device
    gdtype = torch.float32 if use_main_grad else dtype
    offsets, off = [], 0
    for p in parameters:
        assert p.dtype == dtype, "all parameters in a fused buffer must share a dtype"
        offsets.append(off)
        off += _align(p.numel(), p.element_size())
    total = _align(off, parameters[0].element_size(), nranks)
    alloc = allocator or (lambda n, dt, d: torch.zeros(n, dtype=dt, device=d))
    pbuf = None
    if fuse_param:
        pbuf = alloc(total, dtype, dev)
        with torch.no_grad():
            for p, o in zip(parameters, offsets):
                pbuf[o:o + p.numel()].copy_(p.detach().reshape(-1))
                p.data = pbuf[o:o + p.numel()].view(p.shape)
    gbuf = alloc(total, gdtype, dev)
    for p, o in zip(parameters, offsets):
        view = gbuf[o:o + p.numel()].view(p.shape)
        if use_main_grad:
            p.main_grad = view
        else:
            p.grad = view
    return pbuf, gbuf, offsets


class FusedCommBuffer:
    """One fused gradient (and optionally parameter) buffer plus its collective action."""

    def __init__(self, id, params, comm_group, acc_steps=1, act=None, dst=-1, use_main_grad=None, fuse_param=False,
                 scale_after_comm=True, release_grads=False, use_reduce_avg=True, allocator=None):
        self._id = id
        self._params = list(params)
        self._acc_steps = acc_steps
        self._comm_group = comm_group
        self._scale_after_comm = scale_after_comm
        self._use_reduce_avg = use_reduce_avg
        self.use_main_grad = use_main_grad if use_main_grad is not None else hasattr(self._params[0], "main_grad")
        self._task = None
        self._params_step_dict = {}
        self._params_checked_in = 0
        self._act = act if act is not None else HOOK_ACTION.ALL_REDUCE
        self._dst = dst
        if self._act == HOOK_ACTION.REDUCE:
            assert dst != -1
        n = comm_group.nranks if comm_group is not None else 1
        self._nranks = n
        self.param_storage, self.grad_storage, self._offsets = flatten_dense_tensors(
            self._params, use_main_grad=self.use_main_grad, fuse_param=fuse_param,
            nranks=n if self._act == HOOK_ACTION.REDUCE_SCATTER else 1, allocator=allocator)
        self._shard_numel = self.grad_storage.numel() // n if self._act == HOOK_ACTION.REDUCE_SCATTER else None
        self._init_step_dict()

    @property
    def params(self):
        return self._params

    def _init_step_dict(self):
        self._params_step_dict = {p.name: 0 for p in self._params}

    def _reset_params_checked_in(self):
        self._task = None
        self._init_step_dict()
        self._params_checked_in = 0

    @property
    def _all_params_checked_in(self):
        return len(self._params) == self._params_checked_in and len(self._params_step_dict) == 0

    def grad_view(self, p):
        i = next(k for k, q in enumerate(self._params) if q is p)
        o = self._offsets[i]
        return self.grad_storage[o:o + p.numel()].view(p.shape)

    def add_grad(self, param, use_comm=True):
        assert param.name in self._params_step_dict
        g = param.main_grad if self.use_main_grad else param.grad
        view = self.grad_view(param)
        if g is not None and g.data_ptr() != view.data_ptr():
            # autograd produced a fresh tensor: fold it into the bucket view (reference asserts instead, :530-535)
            with torch.no_grad():
                view.add_(g.to(view.dtype)) if self._params_step_dict[param.name] > 0 else view.copy_(g)
            if self.use_main_grad:
                param.main_grad = view
            else:
                param.grad = view
        self._params_step_dict[param.name] += 1
        if self._params_step_dict[param.name] == self._acc_steps:
            self._params_checked_in += 1
            self._params_step_dict.pop(param.name)
        if self._all_params_checked_in and use_comm:
            self.comm_grads()

    def local_shard(self, buf=None):
        buf = self.grad_storage if buf is None else buf
        r = self._comm_group.rank if self._nranks > 1 else 0
        return buf[r * self._shard_numel:(r + 1) * self._shard_numel]

    @torch.no_grad()
    def comm_grads(self):
        assert self._all_params_checked_in, (
            f"Not all params checked in. Parameter number: {len(self._params)}, Check-in number: {self._params_checked_in}")
        self._comm_grads()

    def _comm_grads(self):
        if self._nranks == 1:
            return
        avg = self._use_reduce_avg
        if not self._scale_after_comm and not avg:
            self.grad_storage.mul_(1.0 / self._nranks)
        op = ReduceOp.AVG if avg else ReduceOp.SUM
        if self._act == HOOK_ACTION.ALL_REDUCE:
            self._task = comm.all_reduce(self.grad_storage, op, self._comm_group, sync_op=False)
        elif self._act == HOOK_ACTION.REDUCE:
            self._task = comm.reduce(self.grad_storage, self._dst, op, self._comm_group, sync_op=False)
        else:
            self._task = comm.reduce_scatter(self.local_shard(), self.grad_storage, op, self._comm_group, sync_op=False)

    @torch.no_grad()
    def scale_grads(self):
        if self._task is not None:
            self._task.wait()
        if self._nranks > 1 and self._scale_after_comm and not self._use_reduce_avg:
            tgt = self.local_shard() if self._act == HOOK_ACTION.REDUCE_SCATTER else self.grad_storage
            tgt.mul_(1.0 / self._nranks)
        self._reset_params_checked_in()

    @torch.no_grad()
    def sync_params(self, sync=True):
        """All-gather the parameter buffer from the per-rank shards (after the sharded optimizer step)."""
        assert self.param_storage is not None and self._act == HOOK_ACTION.REDUCE_SCATTER
        if self._nranks == 1:
            return None
        t = comm.all_gather(self.param_storage, self.local_shard(self.param_storage), self._comm_group, sync_op=sync)
        return t


def obtain_storage(parameters, use_main_grad=False, clip=True, dist=False, fuse_param=True, comm_overlap=False, act=None,
                   comm_group=None, dst=-1, acc_steps=1, scale_after_comm=False, group_size=256 * 1024 * 1024,
                   use_reduce_avg=True):
    if not parameters:
        return [], []
    groups = assign_group_by_size(parameters, group_size)
    buffers = []
    for gid, ps in groups.items():
        buffers.append(FusedCommBuffer(gid, ps, comm_group, acc_steps, act, dst, use_main_grad, fuse_param,
                                       scale_after_comm, use_reduce_avg=use_reduce_avg))
    return buffers


def fused_parameters(parameters, use_main_grad=False, fuse_param=True, comm_overlap=False, comm_group=None, act=None,
                     dst=-1, acc_step=1, scale_after_comm=False, group_params=False, apply_decay_param_fun=None,
                     use_reduce_avg=True, group_size=256 * 1024 * 1024):
    """Group parameters (by dtype) into FusedCommBuffers; returns (decay_fused, all_fused, comm_buffers) in the
    reference's shape where the first two are the parameter lists (unchanged objects, now buffer views)."""
    by_dtype = {}
    for p in parameters:
        if p.requires_grad:
            by_dtype.setdefault(p.dtype, []).append(p)
    buffers = []
    for ps in by_dtype.values():
        buffers += obtain_storage(ps, use_main_grad, fuse_param=fuse_param, comm_overlap=comm_overlap, act=act,
                                  comm_group=comm_group, dst=dst, acc_steps=acc_step, scale_after_comm=scale_after_comm,
                                  group_size=group_size, use_reduce_avg=use_reduce_avg)
    decay = [p for p in parameters if apply_decay_param_fun is None or apply_decay_param_fun(p.name)]
    return decay, list(parameters), buffers
