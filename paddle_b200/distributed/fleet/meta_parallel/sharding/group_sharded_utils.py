"""GroupSharded helpers (reference: .../sharding/group_sharded_utils.py): sharded global-norm clip and the scaler
wrapper whose found_inf is all-reduced (MAX) over the sharding group."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp


def _comm_device(dev):
    """Collectives need device tensors on NCCL: host-resident (offloaded) values take a round trip."""
    from paddle_b200.distributed import collective as C
    if dev.type == "cpu" and C._state.backend == "nccl" and torch.cuda.is_available():
        return torch.device("cuda", torch.cuda.current_device())
    return dev


class GroupShardedClipGrad:
    def __init__(self, clip, device, group):
        self._clip = clip
        self._device = device
        self._group = group

    @torch.no_grad()
    def _dygraph_clip(self, params_grads):
        """params_grads hold only this rank's slices/owned params except ``unslice`` ones (replicated)."""
        if not any(g is not None for _, g in params_grads):
            return params_grads
        coef = self.fold_coef(params_grads)
        for p, g in params_grads:
            if g is not None and getattr(p, "need_clip", True):
                g.mul_(coef.to(device=g.device, dtype=g.dtype))
        return params_grads

    @torch.no_grad()
    def fold_coef(self, params_grads):
        """The clip coefficient as a device scalar (never read by the host): local sum of squares of the sliced
        gradients (one multi-tensor launch), summed over the sharding group, plus the replicated ones.  The optimizer
        multiplies it into the gradients inside the fused update kernel instead of running a pass over them."""
        sliced, unsliced = [], []
        for p, g in params_grads:
            if g is None or getattr(p, "need_clip", True) is False:
                continue
            (unsliced if getattr(p, "_gs_unslice", False) else sliced).append(g)
        some = next(g for _, g in params_grads if g is not None)
        dev = some.device
        comm_dev = _comm_device(dev)
        sq = ops.squared_l2_norm(sliced).to(comm_dev).float().reshape(1) if sliced else torch.zeros(1, device=comm_dev)
        comm.all_reduce(sq, ReduceOp.SUM, self._group)
        if unsliced:
            sq = sq + ops.squared_l2_norm(unsliced).to(comm_dev).float().reshape(1)
        gn = torch.sqrt(sq)
        mx = torch.full_like(gn, self._clip.clip_norm)
        return (mx / torch.maximum(gn, mx)).to(dev)

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


def offload_optimizer_state(optimizer, params, wrap_step=True):
    """``offload=True`` of GroupSharded stage 2 / 3 (reference: group_sharded_stage3.py:98-127, 294;
    group_sharded_optimizer_stage2.py:120-160 ``_offload_*``).  Every parameter in ``params`` gets an fp32 shadow in
    pinned host memory that the inner optimizer owns instead (so master weights and moments are host tensors and the
    update runs on the CPU); each step copies the gradients down (``_offload_pre``), steps, and copies the updated
    values back up (``_offload_post``).  ``wrap_step=False`` leaves calling the two halves to the sharding wrapper."""
    from paddle_b200.framework.tensor_patch import Parameter
    ids = {id(p) for p in params}
    shadows = {}
    for p in params:
        host = torch.empty(p.shape, dtype=torch.float32, pin_memory=torch.cuda.is_available())
        host.copy_(p.detach().float())
        sh = Parameter(host, requires_grad=True, name=p.name + "@offload")
        sh.need_clip = getattr(p, "need_clip", True)
        sh.optimize_attr = getattr(p, "optimize_attr", {"learning_rate": 1.0})
        for k in ("_gs_unslice", "_origin_name"):
            if hasattr(p, k):
                setattr(sh, k, getattr(p, k))
        shadows[id(p)] = sh
    swap = lambda lst: [shadows[id(q)] if id(q) in ids else q for q in lst]   # noqa: E731
    optimizer._parameter_list = swap(optimizer._parameter_list)
    for g in optimizer._param_groups:
        g["params"] = swap(g["params"])
    optimizer._offload_pairs = [(p, shadows[id(p)]) for p in params]

    @torch.no_grad()
    def pre():
        for p, sh in optimizer._offload_pairs:
            g = getattr(p, "main_grad", None)
            g = g if g is not None else p.grad
            sh.grad = None if g is None else g.detach().to("cpu", torch.float32, non_blocking=False)

    @torch.no_grad()
    def post():
        for p, sh in optimizer._offload_pairs:
            p.data.copy_(sh.detach().to(p.dtype), non_blocking=True)
            sh.grad = None
    optimizer._offload_pre, optimizer._offload_post = pre, post
    if wrap_step:
        inner_step = optimizer.step

        def step():
            pre()
            inner_step()
            post()
        optimizer.step = step
    return optimizer
