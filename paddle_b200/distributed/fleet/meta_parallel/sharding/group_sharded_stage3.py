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


class _NvlTransport:
    """Stage-3 traffic over NVLink symmetric memory instead of NCCL (K3 push all-gather, K2-pass-1 pull reduce-scatter).

    Every rank keeps its parameter slices inside one symmetric arena; gathered parameters live in a small ring of
    arena slots (a layer is gathered into the next slot while the previous layer computes), gradients are staged in
    two arena buffers from which every owner pulls and averages its slice through the switch.  All transport kernels
    run on one side stream in the same order on every rank; the tcgen05 GEMMs leave >90 % of the issue slots idle, so
    the copy kernels co-reside with them on the SMs."""
    SLOTS = 4

    def __init__(self, group, params, device):
        from paddle_b200.parallel import symm
        n = group.nranks
        al = lambda b: (b + 255) // 256 * 256   # noqa: E731
        slice_bytes = sum(al(_padded_numel(p, n) // n * p.element_size()) for p in params)
        self.max_param_bytes = max(al(_padded_numel(p, n) * p.element_size()) for p in params)
        self.slot_bytes = self.max_param_bytes   # grown by the owner once units are known (see reserve)
        self._pending = (group, slice_bytes, device)
        self.symm = symm
        self.arena = None

    def reserve(self, max_unit_bytes):
        group, slice_bytes, device = self._pending
        self.slot_bytes = max(self.slot_bytes, max_unit_bytes)
        total = slice_bytes + self.SLOTS * self.slot_bytes + 2 * self.max_param_bytes + (8 << 20)
        self.arena = self.symm.SymmArena(group, total)
        self.stream = torch.cuda.Stream(device=device)
        import os
        # grid caps of the transport kernels (0 = library default); measured: capping them exposes transfer time
        self.ag_ctas = int(os.environ.get("PDB200_STAGE3_AG_CTAS", "0"))
        self.symm.lib.kernels.pdb200_nvl_set_rs_max_ctas(int(os.environ.get("PDB200_STAGE3_RS_CTAS", "0")))
        self.slots = None
        self.slot_owner = [None] * self.SLOTS
        self.next_slot = 0
        self.gstage = None
        self.gstage_done = [None, None]
        self.next_gstage = 0
        self.rs_scratch = None

    def finish_layout(self):
        """Called after every slice has been placed: the ring and the gradient stages take the rest of the arena."""
        self.slots = [self.arena.empty((self.slot_bytes,), torch.uint8) for _ in range(self.SLOTS)]
        self.gstage = [self.arena.empty((self.max_param_bytes,), torch.uint8)]

    def new_slice(self, numel, dtype):
        return self.arena.empty((numel,), dtype)

    def _view(self, buf, byte_off, numel, dtype):
        arena, off = buf._symm
        t = arena.tensor(off + byte_off, (numel,), dtype)
        t._symm = (arena, off + byte_off)
        return t

    def gather(self, unit):
        slot = None
        for k in range(self.SLOTS):                 # next slot whose occupant has been released
            c = (self.next_slot + k) % self.SLOTS
            prev = self.slot_owner[c]
            if prev is None or prev is unit or not (prev.gathered or prev.gather_task is not None):
                slot = c
                break
        if slot is None:
            names = [type(o.layer).__name__ + ":" + (o.params[0].name or "?") for o in self.slot_owner]
            raise RuntimeError(f"stage-3 gather ring ({self.SLOTS} slots) is full: {names}")
        self.next_slot = (slot + 1) % self.SLOTS
        self.slot_owner[slot] = unit
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)           # the slot's previous contents and this rank's slices are final up to here
        fulls = []
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ev)
            off = 0
            for p in unit.params:
                full = self._view(self.slots[slot], off, p._gs_padded, p.dtype)
                self.symm.all_gather_push(p.fw_storage._symm_view, full, max_ctas=self.ag_ctas)
                fulls.append((p, full))
                off += (p._gs_padded * p.element_size() + 255) // 256 * 256
            done = torch.cuda.Event()
            done.record(self.stream)
        return fulls, done

    def reduce_scatter(self, p, grad, per, out_dtype):
        """Everything (staging copy, pull + average, cast) runs on the transport stream; the compute stream only
        records "gradient ready" and later waits for ``done``."""
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        grad.record_stream(self.stream)
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ev)
            stage = self._view(self.gstage[0], 0, p._gs_padded, grad.dtype)   # one buffer: the stream serialises its reuse
            flat = grad.reshape(-1)
            stage[: flat.numel()].copy_(flat)
            if flat.numel() != p._gs_padded:
                stage[flat.numel():].zero_()
            acc = torch.empty(per, dtype=torch.float32, device=grad.device)
            scratch = torch.empty(self.symm.lib.kernels.pdb200_nvl_rs_grid(int(per)), dtype=torch.float32, device=grad.device)
            self.symm.reduce_scatter_avg(stage, self.arena.rank * per, per, acc, scratch)
            out = acc if out_dtype == torch.float32 else acc.to(out_dtype)
            done = torch.cuda.Event()
            done.record(self.stream)
        out.record_stream(cur)
        return out, done

    def close(self):
        if self.arena is not None:
            self.arena.close()
            self.arena = None


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
        import os
        # how many layers ahead parameters are gathered (the reference prefetches one, :849-896); two hides the
        # gather of short layers (o_proj) behind the long ones
        self._prefetch_depth = max(1, int(os.environ.get("PDB200_STAGE3_PREFETCH", "2")))
        self._nvl = None
        if pretrain_sync_models:
            self._sync_params_and_buffers()
        self._init_transport()
        self._segment_rank_params()
        if self._nvl is not None:
            self._nvl.finish_layout()
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

    def _init_transport(self):
        """Use the NVLink symmetric-memory transport when the group is one NVSwitch node and every slice keeps the
        16-byte granularity of the copy kernels; otherwise NCCL (also the numerical oracle: FLAGS_use_fused_comm=0)."""
        from paddle_b200.framework.flags import flag
        from paddle_b200.parallel import symm
        dev = next(iter(self._layer.parameters())).device
        # opt-in: measured on 2 x B200 (Llama-2 7B, b=2): NCCL transport 397 ms/step, this transport 406-425 ms/step
        if not flag("stage3_nvl_transport") or dev.type != "cuda" or not symm.available(self._group):
            return
        if not hasattr(symm.lib.kernels.require(), "pdb200_nvl_allgather"):
            return
        n = self._world_size
        cands, unit_bytes, seen = [], 0, set()
        for sub in self._layer.sublayers(include_self=True):
            own = [p for p in sub.parameters(include_sublayers=False) if id(p) not in seen]
            ub = 0
            for p in own:
                seen.add(id(p))
                if p.requires_grad and p.numel() >= self._segment_size and not self._is_excluded(sub):
                    if p.dtype not in (torch.bfloat16, torch.float32) or (_padded_numel(p, n) // n) % 8 != 0:
                        return
                    cands.append(p)
                    ub += (_padded_numel(p, n) * p.element_size() + 255) // 256 * 256
            unit_bytes = max(unit_bytes, ub)
        if not cands:
            return
        try:
            t = _NvlTransport(self._group, cands, dev)
            t.reserve(unit_bytes)
            self._nvl = t
        except Exception as e:  # noqa: BLE001
            import warnings
            warnings.warn(f"stage 3: symmetric-memory transport unavailable ({e}); using NCCL")
            self._nvl = None

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
        if self._nvl is not None:
            sl = self._nvl.new_slice(per, p.dtype)
            sl.copy_(flat[self._rank * per:(self._rank + 1) * per])
        else:
            sl = flat[self._rank * per:(self._rank + 1) * per].clone()
        fw = Parameter(sl, requires_grad=True, name=p.name + "@slice")
        fw.data = sl
        fw._symm_view = sl if self._nvl is not None else None
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
        if self._nvl is not None:
            with torch.no_grad():
                unit.gather_task = self._nvl.gather(unit)
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
        if self._nvl is not None:
            fulls, done = unit.gather_task
            torch.cuda.current_stream().wait_event(done)
            for p, full in fulls:
                p.data = full[: p._gs_numel].view(p._gs_shape)
            unit.gather_task = None
            unit.gathered = True
            return
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
                    for k in range(1, self._prefetch_depth + 1):
                        if i + k < len(self._order_fw):
                            self._launch_gather(self._order_fw[i + k])
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
        if i is not None:
            for k in range(1, self._prefetch_depth + 1):
                if i - k >= 0:
                    self._launch_gather(self._order_fw[i - k])

    def _make_grad_hook(self, unit, p):
        @torch.no_grad()
        def hook(param):
            g = param.grad
            if g is None:
                return
            n = self._world_size
            per = p._gs_padded // n
            if self._nvl is not None:
                out, done = self._nvl.reduce_scatter(p, g, per, torch.float32 if self._use_main_grad else p.dtype)
                self._grad_tasks.append((p, out, done, None))
                param.grad = None
                unit.pending_grads -= 1
                if unit.pending_grads <= 0:
                    self._release(unit)
                return
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
            if self._nvl is not None:
                torch.cuda.current_stream().wait_event(t)
            else:
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
