"""ZeRO-1 sharding optimizers (reference: python/paddle/distributed/fleet/meta_optimizers/dygraph_optimizer/
dygraph_sharding_optimizer.py:44-960).

V1 (``DygraphShardingOptimizer``): whole parameters are assigned to ranks by a greedy size balance (:246-280);
gradients are reduced to the owner, the owner updates, updated parameters are broadcast.

V2 (``DygraphShardingOptimizerV2``, ``split_param``): every parameter is sliced across ranks inside flat fused
buffers; gradients are reduce-scattered (AVG), the optimizer runs on the local contiguous shard (fp32 master / m /
v of shard size only), parameters are all-gathered.  With FLAGS_use_fused_comm and symmetric memory the three
steps run as K2 (reduce-scatter + AdamW in one kernel pulling peer slices over NVLink) and K3 (multimem
all-gather); the composition below is the NCCL baseline and the numerical oracle.
"""
from __future__ import annotations

import os

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.distributed.fleet.utils import tensor_fusion_helper as tfh
from paddle_b200.framework.tensor_patch import Parameter


def _grad_of(p):
    mg = getattr(p, "main_grad", None)
    return mg if mg is not None else p.grad


def _flat_params(opt):
    pl = opt._parameter_list
    if pl and isinstance(pl[0], dict):
        out = []
        for g in pl:
            out.extend(g["params"])
        return out
    return list(pl)


class DygraphShardingOptimizer:
    def __init__(self, optimizer, hcg):
        self._inner_opt = optimizer
        self._hcg = hcg
        self._sharding_world_size = hcg.get_sharding_parallel_world_size()
        self._sharding_rank = hcg.get_sharding_parallel_rank()
        self._group = hcg.get_sharding_parallel_group()
        self._parameter_list = _flat_params(optimizer)
        self._origin_parameter_list = list(self._parameter_list)
        self._rank2params = self._partition_parameters()
        self._param2rank = {p.name: r for r, ps in self._rank2params.items() for p in ps}
        local = self._rank2params[self._sharding_rank]
        self._set_inner_opt_attr("_parameter_list", local)
        self._set_inner_opt_attr("_param_groups", [{"params": local}])
        from paddle_b200.distributed import fleet
        st = fleet.fleet._user_defined_strategy
        sc = st.hybrid_configs["sharding_configs"] if st is not None else {}
        ppc = st.hybrid_configs["pp_configs"] if st is not None else {}
        self.tensor_fusion = bool(sc.get("tensor_fusion", False))
        self.comm_overlap = bool(sc.get("comm_overlap", False))
        self.accumulate_steps = int(sc.get("accumulate_steps", 1))
        self._use_reduce_avg = bool(sc.get("use_reduce_avg", True))
        self._comm_buffers = []
        # the pipeline engine owns the overlap when pp_configs.sharding_comm_overlap is set (it knows the micro-batch count)
        if (self.tensor_fusion or self.comm_overlap) and not ppc.get("sharding_comm_overlap", False):
            self._build_fused_buffers(int(sc.get("comm_buffer_size_MB", 256)) << 20)

    def _build_fused_buffers(self, group_bytes):
        """``tensor_fusion``: the gradients of every owner rank's parameters live in fused buffers, so a step needs one
        reduce per buffer instead of one per parameter.  ``comm_overlap``: a buffer is reduced from the backward hook of
        the last of its parameters (after ``accumulate_steps`` backward passes), overlapping the rest of the backward
        (reference: dygraph_sharding_optimizer.py:120-230, tensor_fusion_helper.fused_parameters)."""
        for r, ps in sorted(self._rank2params.items()):
            ps = [q for q in ps if q.requires_grad]
            if not ps:
                continue
            for group in tfh.assign_group_by_size(ps, group_bytes).values():
                buf = tfh.FusedCommBuffer(len(self._comm_buffers), group, self._group, self.accumulate_steps,
                                          tfh.HOOK_ACTION.REDUCE, dst=self._group.ranks[r], use_reduce_avg=self._use_reduce_avg)
                self._comm_buffers.append(buf)
                if self.comm_overlap:
                    for q in group:
                        q.register_post_accumulate_grad_hook(lambda *_, b=buf, prm=q: b.add_grad(prm))

    def _set_inner_opt_attr(self, name, value):
        opt = self._inner_opt
        while opt is not None:
            if hasattr(opt, name) or not hasattr(opt, "_inner_opt"):
                if name == "_param_groups":
                    # keep per-group hyper-parameters of the first group
                    base = dict(opt._param_groups[0]) if getattr(opt, "_param_groups", None) else {}
                    base.update(value[0])
                    value = [base]
                setattr(opt, name, value)
            opt = getattr(opt, "_inner_opt", None)

    def _partition_parameters(self):
        mapping = {r: [] for r in range(self._sharding_world_size)}
        sizes = [0.0] * self._sharding_world_size
        params = list(self._parameter_list)
        if os.environ.get("FLAGS_sharding_sort_parameters", "1") in ("1", "true", "True"):
            params.sort(key=lambda p: p.numel(), reverse=True)
        for p in params:
            r = sizes.index(min(sizes))
            mapping[r].append(p)
            sizes[r] += max(p.numel(), 1)
        return mapping

    def filter_parameters(self, parameter_list, hcg):
        local = {id(p) for p in self._rank2params[self._sharding_rank]}
        return [p for p in parameter_list if id(p) in local]

    @torch.no_grad()
    def reduce_gradients(self, parameter_list, hcg):
        if self._comm_buffers:
            for buf in self._comm_buffers:
                if not self.comm_overlap:                      # fused, not overlapped: check everything in, then one reduce
                    buf._params_step_dict = {}
                    buf._params_checked_in = len(buf.params)
                    for q in buf.params:
                        g = _grad_of(q)
                        view = buf.grad_view(q)
                        if g is None:
                            view.zero_()
                        elif g.data_ptr() != view.data_ptr():
                            view.copy_(g)
                        if buf.use_main_grad:
                            q.main_grad = view
                        else:
                            q.grad = view
                    buf.comm_grads()
                buf.scale_grads()
            return
        tasks = []
        for p in parameter_list:
            g = _grad_of(p)
            if g is None or not p.requires_grad:
                continue
            owner = self._group.ranks[self._param2rank[p.name]]
            tasks.append(comm.reduce(g, owner, ReduceOp.AVG, self._group, sync_op=False))
        for t in tasks:
            t.wait()

    @torch.no_grad()
    def _sharding_sync_parameters(self):
        tasks = []
        for r, ps in self._rank2params.items():
            src = self._group.ranks[r]
            for p in ps:
                tasks.append(comm.broadcast(p, src, self._group, sync_op=False))
        for t in tasks:
            t.wait()

    def clear_grad(self, set_to_zero=True):
        for p in self._origin_parameter_list:
            if p.grad is not None:
                if set_to_zero:
                    p.grad.zero_()
                else:
                    p.grad = None
            mg = getattr(p, "main_grad", None)
            if mg is not None:
                mg.zero_()

    clear_gradients = clear_grad

    def _amp_defer_unscale(self, scaler):
        """GradScaler.unscale_ hook: gradients are reduced to their owner in ``reduce_gradients`` after
        ``scaler.step`` called ``unscale_``; unscaling only the owned ones before that would mix scaled and unscaled
        contributions.  ``step`` unscales + checks the owned, reduced gradients instead."""
        self.__dict__["_amp_scaler"] = scaler
        return True

    @torch.no_grad()
    def step(self):
        scaler = self.__dict__.pop("_amp_scaler", None)
        if scaler is not None:
            grads = [g for g in (_grad_of(p) for p in self._rank2params[self._sharding_rank]) if g is not None]
            scaler._unscale_grads(grads)
            if scaler._found_inf is not None:
                comm.all_reduce(scaler._found_inf, ReduceOp.MAX, self._group)
            scaler._sync_found_inf(self)
        # gradients of parameters owned elsewhere are not needed for the update
        self._inner_opt.step()
        self._sharding_sync_parameters()

    def minimize(self, loss, startup_program=None, parameters=None, no_grad_set=None):
        self.step()
        return None, None

    def set_state_dict(self, state_dict):
        local_names = {p.name for p in self._rank2params[self._sharding_rank]}
        keep = {}
        for k, v in state_dict.items():
            if k in ("master_weights", "LR_Scheduler"):
                keep[k] = {kk: vv for kk, vv in v.items() if kk in local_names} if k == "master_weights" else v
                continue
            if any(k.startswith(n + "_") for n in local_names):
                keep[k] = v
        return self._inner_opt.set_state_dict(keep)

    def __getattr__(self, item):
        return getattr(self.__dict__["_inner_opt"], item)


class DygraphShardingOptimizerV2:
    """split_param sharding on flat buffers."""

    def __init__(self, optimizer, hcg):
        self._inner_opt = optimizer
        self._hcg = hcg
        self._group = hcg.get_sharding_parallel_group()
        self._sharding_world_size = hcg.get_sharding_parallel_world_size()
        self._sharding_rank = hcg.get_sharding_parallel_rank()
        self._parameter_list = _flat_params(optimizer)
        self._origin_parameter_list = list(self._parameter_list)
        from paddle_b200.distributed import fleet
        st = fleet.fleet._user_defined_strategy
        sc = st.hybrid_configs["sharding_configs"] if st is not None else {}
        self._use_reduce_avg = sc.get("use_reduce_avg", True)
        self._acc_steps = sc.get("accumulate_steps", 1)
        self.comm_overlap = sc.get("comm_overlap", False)
        self._buffer_bytes = int(sc.get("comm_buffer_size_MB", 256)) << 20
        self._use_main_grad = any(hasattr(p, "main_grad") for p in self._parameter_list)
        self._build_buffers()
        self._all_gather_overlap_forward = False

    def _build_buffers(self):
        by_dtype = {}
        for p in self._parameter_list:
            if p.requires_grad:
                by_dtype.setdefault((p.dtype, bool(getattr(p, "is_distributed", False))), []).append(p)
        self._comm_buffers = []
        self._slice_params = []
        self._slice_of = {}
        idx = 0
        allocator = self._symm_allocator(by_dtype)
        for (_, is_dist), ps in by_dtype.items():
            groups = tfh.assign_group_by_size(ps, self._buffer_bytes)
            for _, gps in groups.items():
                buf = tfh.FusedCommBuffer(idx, gps, self._group, self._acc_steps, tfh.HOOK_ACTION.REDUCE_SCATTER,
                                          use_main_grad=self._use_main_grad, fuse_param=True,
                                          use_reduce_avg=self._use_reduce_avg, scale_after_comm=True,
                                          allocator=allocator)
                idx += 1
                self._comm_buffers.append(buf)
                # the local contiguous shard of the parameter buffer becomes ONE optimizer parameter
                shard = buf.local_shard(buf.param_storage)
                sp = Parameter(shard, requires_grad=True, name=f"sharding_v2_slice_{self._sharding_rank}_{buf._id}")
                sp.data = shard            # share storage with the flat parameter buffer
                sp._segments = self._segments(buf)
                sp.need_clip = True
                sp.is_distributed = is_dist   # keeps the mp dist / replicated split for the global-norm clip
                self._slice_params.append(sp)
                self._slice_of[buf._id] = sp
        opt = self._inner_opt
        while opt is not None:
            if hasattr(opt, "_parameter_list"):
                opt._parameter_list = self._slice_params
                base = dict(opt._param_groups[0]) if getattr(opt, "_param_groups", None) else {}
                base["params"] = self._slice_params
                opt._param_groups = [base]
            opt = getattr(opt, "_inner_opt", None)

    def _symm_allocator(self, by_dtype):
        """Flat parameter/gradient buffers in NVLink symmetric memory so that K2+K3 (reduce-scatter + AdamW +
        all-gather in one kernel) can run on them; None -> ordinary allocations and the NCCL composition."""
        self._arena = None
        from paddle_b200.parallel import symm
        some = next((p for ps in by_dtype.values() for p in ps), None)
        if some is None or not some.is_cuda or not symm.available(self._group):
            return None
        need = 0
        for ps in by_dtype.values():
            elems = sum((p.numel() + 127) // 128 * 128 for p in ps) + 4096 * self._sharding_world_size * 4
            need += elems * ps[0].element_size() + elems * (4 if self._use_main_grad else ps[0].element_size())
        try:
            self._arena = symm.get_arena(self._group, need + (16 << 20), key=f"shardv2_{id(self)}")
        except Exception:  # noqa: BLE001
            self._arena = None
            return None
        arena = self._arena

        def alloc(numel, dtype, device):
            t = arena.empty((numel,), dtype)
            t.zero_()
            return t
        return alloc

    def _fused_ready(self):
        """Fused K2+K3 preconditions: symmetric bf16 buffers, native multi-precision AdamW with uniform decay."""
        if self._arena is None:
            return False
        base = self._inner_opt
        while hasattr(base, "_inner_opt"):
            base = base._inner_opt
        ok = (getattr(base, "_decoupled", False) and base._multi_precision and
              getattr(base, "_apply_decay_param_fun", None) is None and getattr(base, "_lr_ratio_fn", None) is None and
              all(b.param_storage.dtype == torch.bfloat16 and b._shard_numel % 8 == 0 for b in self._comm_buffers))
        self._base_opt = base
        return ok

    def _segments(self, buf):
        """[(param, start_in_shard, start_in_param, length)] overlap of each parameter with this rank's shard."""
        S = buf._shard_numel
        lo, hi = self._sharding_rank * S, (self._sharding_rank + 1) * S
        segs = []
        for p, off in zip(buf._params, buf._offsets):
            a, b = max(off, lo), min(off + p.numel(), hi)
            if a < b:
                segs.append((p, a - lo, a - off, b - a))
        return segs

    def filter_parameters(self, parameter_list, hcg):
        return []   # the dp all-reduce runs on the reduced shards inside reduce_gradients

    @torch.no_grad()
    def reduce_gradients(self, parameter_list, hcg):
        self._fused_pending = False
        for buf in self._comm_buffers:
            for p in buf._params:
                g = _grad_of(p)
                view = buf.grad_view(p)
                if g is None:
                    view.zero_()
                elif g.data_ptr() != view.data_ptr():
                    view.copy_(g)
                    if buf.use_main_grad:
                        p.main_grad = view
                    else:
                        p.grad = view
        if self._fused_ready():
            self._fused_pending = True       # the reduction happens inside the fused update kernel(s) in step()
            self._hcg_for_step = hcg
            return
        for buf in self._comm_buffers:
            buf._params_checked_in = len(buf._params)
            buf._params_step_dict = {}
            buf._comm_grads()
        for buf in self._comm_buffers:
            buf.scale_grads()
        # hand the reduced shard to the slice parameter as its gradient
        for buf in self._comm_buffers:
            sp = self._slice_of[buf._id]
            gshard = buf.local_shard()
            if buf.use_main_grad:
                sp.main_grad = gshard
            else:
                sp.grad = gshard
        dp = hcg.get_data_parallel_group() if hcg.get_data_parallel_world_size() > 1 else None
        if dp is not None:
            for buf in self._comm_buffers:
                comm.all_reduce(buf.local_shard(), ReduceOp.AVG, dp)

    @torch.no_grad()
    def _sharding_sync_parameters(self):
        tasks = [buf.sync_params(sync=False) for buf in self._comm_buffers]
        for t in tasks:
            if t is not None:
                t.wait()

    def _amp_defer_unscale(self, scaler):
        """GradScaler.unscale_ hook: when ``scaler.step`` runs, the flat gradient buffers still hold this rank's
        unreduced gradients (the reduce-scatter happens in ``reduce_gradients`` / the fused kernel).  Remember the
        scaler; ``step`` unscales and checks the reduced shard instead."""
        self.__dict__["_amp_scaler"] = scaler
        return True

    @torch.no_grad()
    def step(self):
        scaler = self.__dict__.pop("_amp_scaler", None)
        if getattr(self, "_fused_pending", False):
            self._fused_step(scaler)
            self._fused_pending = False
            return
        if scaler is not None:
            grads = [g for g in (_grad_of(sp) for sp in self._slice_params) if g is not None]
            scaler._unscale_grads(grads)
            if scaler._found_inf is not None:
                comm.all_reduce(scaler._found_inf, ReduceOp.MAX, self._group)
            scaler._sync_found_inf(self)
        self._inner_opt.step()
        self._sharding_sync_parameters()

    @torch.no_grad()
    def _fused_step(self, scaler=None):
        """K2+K3 over NVLink: no clip -> ONE kernel per buffer (reduce-scatter + AdamW + all-gather); with a
        global-norm clip, a dp axis or a loss scaler -> reduce-scatter+norm kernel, scalar exchange, AdamW+all-gather
        kernel.  With a scaler the squared norm of the reduced (still scaled) shard doubles as the inf/nan check, and
        1/scale is folded into the kernel's gradient multiplier."""
        from paddle_b200.parallel import symm
        opt = self._base_opt
        hcg = self._hcg_for_step
        opt._create_accumulators(self._slice_params)
        lr = opt.get_lr()
        grp = opt._param_groups[0]
        glr = lr * grp.get("learning_rate", 1.0)
        coeff = opt._coeff(grp)
        pows = opt._pows(self._slice_params[0])
        clip = opt._grad_clip
        dp = hcg.get_data_parallel_group() if hcg.get_data_parallel_world_size() > 1 else None
        two_pass = clip is not None or dp is not None or scaler is not None
        S_rank = self._sharding_rank
        local = {}
        if two_pass:
            dev = self._slice_params[0].device
            dist_sq = torch.zeros(1, device=dev)
            not_dist_sq = torch.zeros(1, device=dev)
            for buf in self._comm_buffers:
                S = buf._shard_numel
                g32 = torch.empty(S, dtype=torch.float32, device=dev)
                sq = symm.reduce_scatter_norm(buf.grad_storage, S_rank * S, S, g32)
                if dp is not None:
                    comm.all_reduce(g32, ReduceOp.AVG, dp)
                    sq = (g32 * g32).sum().reshape(1)
                local[buf._id] = g32
                if self._slice_of[buf._id].is_distributed:
                    dist_sq += sq
                else:
                    not_dist_sq += sq
            coef = None
            inv = None
            if scaler is not None:
                scaler._init_state(dev)
                inv = (1.0 / scaler._scale).float()
                bad = (~torch.isfinite(dist_sq + not_dist_sq)).float()
                scaler._found_inf.copy_(torch.maximum(scaler._found_inf, bad))
                comm.all_reduce(scaler._found_inf, ReduceOp.MAX, self._group)
                scaler._sync_found_inf(self)
                dist_sq = dist_sq * inv * inv          # norms of the unscaled gradients for the clip
                not_dist_sq = not_dist_sq * inv * inv
            if clip is not None:
                if hasattr(clip, "_global_norm"):
                    clip._global_norm(dist_sq, not_dist_sq)
                else:
                    comm.all_reduce(not_dist_sq, group=self._group)
                    comm.all_reduce(dist_sq, group=self._group)
                gnorm = torch.sqrt(dist_sq + not_dist_sq)
                mx = torch.full_like(gnorm, float(clip.clip_norm))
                coef = mx / (torch.maximum(gnorm, mx) + 1.0e-6)
            if inv is not None:
                coef = inv if coef is None else coef * inv
        else:
            coef = None
        found_inf = scaler._found_inf if scaler is not None else opt._found_inf
        prev_found = opt._found_inf
        opt._found_inf = found_inf                # _advance_pows must skip the step too
        scale = coef if opt._grad_scale_dev is None else (opt._grad_scale_dev if coef is None else coef * opt._grad_scale_dev)
        for buf in self._comm_buffers:
            sp = self._slice_of[buf._id]
            S = buf._shard_numel
            symm.rs_adamw_ag(buf.grad_storage, buf.param_storage, S_rank * S, S, opt._master_weights[sp.name],
                             opt._get_accumulator("moment1", sp), opt._get_accumulator("moment2", sp), glr, opt._beta1,
                             opt._beta2, opt._epsilon, coeff, pows, grad_scale=scale, found_inf=found_inf,
                             local_grad=local.get(buf._id))
        for _, pw in opt._beta_pows.items():
            opt._advance_pows(pw)
        opt._found_inf = prev_found
        opt._step_count += 1

    def minimize(self, loss, startup_program=None, parameters=None, no_grad_set=None):
        self.step()
        return None, None

    def clear_grad(self, set_to_zero=True):
        for buf in self._comm_buffers:
            buf.grad_storage.zero_()

    clear_gradients = clear_grad

    def _set_all_gather_overlap_forward(self, flag, layers=None):
        self._all_gather_overlap_forward = flag

    def __getattr__(self, item):
        return getattr(self.__dict__["_inner_opt"], item)
