"""GroupSharded stage 3 / ZeRO-3 (reference: python/paddle/distributed/fleet/meta_parallel/sharding/
group_sharded_stage3.py:101-1217).

Kept from the reference: parameters above ``segment_size`` elements are padded to 256 B and to a multiple of nranks
(in elements, :466-484) and only the rank's slice stays resident; smaller parameters stay replicated and their grads
are all-reduced; a forward pre-hook all-gathers a layer's parameters (and prefetches the next layer in the order
learned on the first step), a post-hook releases them; backward re-gathers; the optimizer steps on the slices with
fp32 master weights; ``get_all_parameters`` rebuilds full tensors for saving.

Changed by design (B200): the gradient path is a reduce-scatter (AVG) of the padded gradient straight into the
slice accumulator - half the wire volume of the reference's all-reduce-then-slice (:741-803); communication runs
async on the group's comm stream and is only awaited where its result is consumed, so gather / scatter overlap the
neighbouring layers' GEMMs.
"""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.framework.tensor_patch import Parameter
from paddle_b200.nn.clip import ClipGradByGlobalNorm
from paddle_b200.nn.layer.layers import Layer

from .group_sharded_utils import GroupShardedClipGrad

ALIGN_BYTES = 256


def _padded_numel(p, nranks):
    a = ALIGN_BYTES // p.element_size()
    n = (p.numel() + a - 1) // a * a
    return (n + nranks - 1) // nranks * nranks


class _Unit:
    """One layer's sliced parameters and their gather state."""

    def __init__(self, layer, params):
        self.layer = layer
        self.params = params
        self.gather_task = None      # list of (param, full_buffer, Task) while an async gather is in flight
        self.gathered = False
        self.pending_grads = 0


class _BackwardGather(torch.autograd.Function):
    """Identity on the layer outputs whose backward re-gathers the layer's parameters first."""

    @staticmethod
    def forward(ctx, stage3, unit, *outs):
        ctx.stage3, ctx.unit = stage3, unit
        return tuple(o.view_as(o) if isinstance(o, torch.Tensor) else o for o in outs)

    @staticmethod
    def backward(ctx, *grads):
        s3, unit = ctx.stage3, ctx.unit
        s3._ensure_gathered(unit)
        s3._prefetch_backward(unit)
        unit.pending_grads = sum(1 for p in unit.params if p.requires_grad)
        return (None, None, *grads)


class GroupShardedStage3(Layer):
    def __init__(self, layer, optimizer, group=None, sync_buffers=False, device="gpu", segment_size=2 ** 20,
                 pretrain_sync_models=True, offload=False, sync_comm=False, dp_group=None, exclude_layer=None):
        super().__init__()
        from paddle_b200.distributed import collective as C
        self._layer = layer
        self._optim = optimizer
        self._group = group if group is not None else C.new_group(list(range(C.get_world_size())))
        self._dp_group = dp_group
        self._world_size = self._group.nranks
        self._rank = self._group.rank
        self._segment_size = segment_size
        self._sync_buffers = sync_buffers
        self._sync_comm = sync_comm
        self._offload = offload
        if offload:
            raise NotImplementedError("offload=True: CPU offload of optimizer state is not implemented in this build")
        self._exclude = list(exclude_layer or [])
        self._units = []
        self._unit_of_layer = {}
        self._order_fw = []
        self._order_learned = False
        self._unslice_params = []
        self._slice_params = []
        self._grad_tasks = []
        self._use_main_grad = False
        if pretrain_sync_models:
            self._sync_params_and_buffers()
        self._segment_rank_params()
        self._register_hooks()
        self._redefine_opt()

    # ------------------------------------------------------------------ setup
    @torch.no_grad()
    def _sync_params_and_buffers(self):
        src = self._group.ranks[0]
        for p in self._layer.parameters():
            comm.broadcast(p, src, self._group)
        for b in self._layer.buffers():
            comm.broadcast(b, src, self._group)

    def _is_excluded(self, sub):
        return any((isinstance(e, type) and isinstance(sub, e)) or (isinstance(e, int) and id(sub) == e) or
                   (isinstance(e, str) and type(sub).__name__ == e) for e in self._exclude)

    @torch.no_grad()
    def _segment_rank_params(self):
        seen = set()
        for sub in self._layer.sublayers(include_self=True):
            own = [p for p in sub.parameters(include_sublayers=False) if id(p) not in seen]
            for p in own:
                seen.add(id(p))
            if not own:
                continue
            sliced = []
            for p in own:
                if not p.requires_grad or p.numel() < self._segment_size or self._is_excluded(sub):
                    p._gs_unslice = True
                    self._unslice_params.append(p)
                    continue
                self._slice_param(p)
                sliced.append(p)
            if sliced:
                u = _Unit(sub, sliced)
                self._units.append(u)
                self._unit_of_layer[id(sub)] = u

    def _slice_param(self, p):
        n = self._world_size
        padded = _padded_numel(p, n)
        per = padded // n
        flat = torch.zeros(padded, dtype=p.dtype, device=p.device)
        flat[: p.numel()].copy_(p.detach().reshape(-1))
        sl = flat[self._rank * per:(self._rank + 1) * per].clone()
        fw = Parameter(sl, requires_grad=True, name=p.name + "@slice")
        fw.data = sl
        fw.need_clip = getattr(p, "need_clip", True)
        fw.optimize_attr = getattr(p, "optimize_attr", {"learning_rate": 1.0})
        fw._origin_name = p.name
        p.fw_storage = fw
        p._gs_shape = tuple(p.shape)
        p._gs_numel = p.numel()
        p._gs_padded = padded
        p._gs_unslice = False
        p.bw_storage = None
        p.data = torch.empty(0, dtype=p.dtype, device=p.device)   # release the full tensor
        self._slice_params.append(p)

    def _register_hooks(self):
        for u in self._units:
            u.layer.register_forward_pre_hook(self._make_pre_hook(u))
            u.layer.register_forward_post_hook(self._make_post_hook(u))
            for p in u.params:
                p.register_post_accumulate_grad_hook(self._make_grad_hook(u, p))
        for p in self._unslice_params:
            if p.requires_grad:
                pass

    # ------------------------------------------------------------------ gather / release
    def _launch_gather(self, unit):
        if unit.gathered or unit.gather_task is not None:
            return
        tasks = []
        with torch.no_grad():
            for p in unit.params:
                full = torch.empty(p._gs_padded, dtype=p.dtype, device=p.fw_storage.device)
                t = comm.all_gather(full, p.fw_storage.detach(), self._group, sync_op=False)
                tasks.append((p, full, t))
        unit.gather_task = tasks

    def _ensure_gathered(self, unit):
        if unit.gathered:
            return
        self._launch_gather(unit)
        for p, full, t in unit.gather_task:
            t.wait()
            p.data = full[: p._gs_numel].view(p._gs_shape)
        unit.gather_task = None
        unit.gathered = True

    def _release(self, unit):
        if not unit.gathered:
            return
        for p in unit.params:
            p.data = torch.empty(0, dtype=p.dtype, device=p.fw_storage.device)
        unit.gathered = False

    def _make_pre_hook(self, unit):
        def hook(layer, inputs):
            self._ensure_gathered(unit)
            if not self._order_learned:
                self._order_fw.append(unit)
            else:
                try:
                    i = self._order_index[id(unit)]
                    if i + 1 < len(self._order_fw):
                        self._launch_gather(self._order_fw[i + 1])
                except KeyError:
                    pass
            return None
        return hook

    def _make_post_hook(self, unit):
        def hook(layer, inputs, outputs):
            self._release(unit)
            if not torch.is_grad_enabled():
                return outputs
            outs = outputs if isinstance(outputs, tuple) else (outputs,)
            if not any(isinstance(o, torch.Tensor) and o.requires_grad for o in outs):
                return outputs
            wrapped = _BackwardGather.apply(self, unit, *outs)
            return wrapped if isinstance(outputs, tuple) else wrapped[0]
        return hook

    def _prefetch_backward(self, unit):
        if not self._order_learned:
            return
        i = self._order_index.get(id(unit))
        if i is not None and i - 1 >= 0:
            self._launch_gather(self._order_fw[i - 1])

    def _make_grad_hook(self, unit, p):
        @torch.no_grad()
        def hook(param):
            g = param.grad
            if g is None:
                return
            n = self._world_size
            per = p._gs_padded // n
            flat = g.reshape(-1)
            if flat.numel() != p._gs_padded:
                tmp = torch.zeros(p._gs_padded, dtype=g.dtype, device=g.device)
                tmp[: flat.numel()].copy_(flat)
                flat = tmp
            out = torch.empty(per, dtype=g.dtype, device=g.device)
            t = comm.reduce_scatter(out, flat, ReduceOp.AVG, self._group, sync_op=False)
            self._grad_tasks.append((p, out, t, flat))
            param.grad = None
            unit.pending_grads -= 1
            if unit.pending_grads <= 0:
                self._release(unit)
        return hook

    @torch.no_grad()
    def _finish_grads(self):
        for p, out, t, _keep in self._grad_tasks:
            t.wait()
            if self._dp_group is not None and self._dp_group.nranks > 1:
                comm.all_reduce(out, ReduceOp.AVG, self._dp_group)
            if p.bw_storage is None:
                p.bw_storage = out.float() if self._use_main_grad else out
            else:
                p.bw_storage.add_(out.to(p.bw_storage.dtype))
        self._grad_tasks = []
        for p in self._unslice_params:
            if p.requires_grad and p.grad is not None:
                comm.all_reduce(p.grad, ReduceOp.AVG, self._group)
                if self._dp_group is not None and self._dp_group.nranks > 1:
                    comm.all_reduce(p.grad, ReduceOp.AVG, self._dp_group)

    # ------------------------------------------------------------------ optimizer redefinition (:819-838)
    def _redefine_opt(self):
        opt = self._optim
        slices = [p.fw_storage for p in self._slice_params]
        local = slices + [p for p in self._unslice_params if p.requires_grad]
        base = dict(opt._param_groups[0]) if getattr(opt, "_param_groups", None) else {}
        base["params"] = local
        opt._parameter_list = local
        opt._param_groups = [base]
        opt._group = self._group
        if isinstance(opt._grad_clip, ClipGradByGlobalNorm):
            opt._grad_clip = GroupShardedClipGrad(opt._grad_clip, None, self._group)
        stage3 = self
        inner_step = opt.step
        inner_clear = opt.clear_grad

        def step():
            stage3._finish_grads()
            for p in stage3._slice_params:
                fw = p.fw_storage
                fw.grad = p.bw_storage
            inner_step()

        def clear_grad(set_to_zero=True):
            for p in stage3._slice_params:
                p.bw_storage = None
                p.fw_storage.grad = None
                p.grad = None
            for p in stage3._unslice_params:
                if p.grad is not None:
                    if set_to_zero:
                        p.grad.zero_()
                    else:
                        p.grad = None
        opt.step = step
        opt.clear_grad = clear_grad
        opt.clear_gradients = clear_grad

    # ------------------------------------------------------------------ forward / utilities
    def forward(self, *inputs, **kwargs):
        if self._sync_buffers:
            with torch.no_grad():
                for b in self._layer.buffers():
                    comm.broadcast(b, self._group.ranks[0], self._group)
        out = self._layer(*inputs, **kwargs)
        if not self._order_learned and self._order_fw:
            self._order_learned = True
            self._order_index = {id(u): i for i, u in enumerate(self._order_fw)}
        return out

    @torch.no_grad()
    def get_all_parameters(self, convert2cpu=False):
        """Materialise every sliced parameter in full (for saving)."""
        for u in self._units:
            self._ensure_gathered(u)
        if convert2cpu:
            for p in self._slice_params:
                p.data = p.data.cpu()

    @torch.no_grad()
    def release_all_parameters(self):
        for u in self._units:
            self._release(u)

    def state_dict(self, *a, **k):
        return self._layer.state_dict(*a, **k)

    def set_state_dict(self, sd, use_structured_name=True):
        self.get_all_parameters()
        r = self._layer.set_state_dict(sd, use_structured_name)
        with torch.no_grad():
            n = self._world_size
            for p in self._slice_params:
                per = p._gs_padded // n
                flat = torch.zeros(p._gs_padded, dtype=p.dtype, device=p.fw_storage.device)
                flat[: p._gs_numel].copy_(p.detach().reshape(-1))
                p.fw_storage.data.copy_(flat[self._rank * per:(self._rank + 1) * per])
        self.release_all_parameters()
        return r

    def parameters(self, *a, **k):
        return self._layer.parameters(*a, **k)

    def named_parameters(self, *a, **k):
        return self._layer.named_parameters(*a, **k)

    def clear_gradients(self):
        self._optim.clear_grad()
