"""HybridParallelOptimizer / HybridParallelClipGrad (reference: python/paddle/distributed/fleet/meta_optimizers/
dygraph_optimizer/hybrid_parallel_optimizer.py:41-520).

Gradient-sync orchestration for hybrid parallelism: sharding reduce -> dp (or dp x sep) fused all-reduce ->
global-norm clip across sharding/mp/pp groups -> inner step -> mp parameter/moment sync.
The squared norms of the "distributed" and "replicated" parameter classes are each ONE multi-tensor kernel (K15);
the clip coefficient stays on the device and, for the native AdamW, is folded into the update kernel instead of a
separate pass over every gradient."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.distributed.fleet.base.topology import ParallelMode
from paddle_b200.distributed.fleet.utils.hybrid_parallel_util import (fused_allreduce_gradients,
                                                                      sharding_reduce_gradients)
from paddle_b200.distributed.fleet.utils.log_util import logger
from paddle_b200.nn.clip import ClipGradByGlobalNorm


def _grad_of(p):
    mg = getattr(p, "main_grad", None)
    return mg if mg is not None else p.grad


class HybridParallelClipGrad:
    def __init__(self, clip, hcg):
        self._clip = clip
        self._hcg = hcg
        self.not_sharding_stage1 = True
        self._fold_target = None      # optimizer whose kernel applies the coefficient (set by HybridParallelOptimizer)

    def _global_norm(self, dist_sq, not_dist_sq):
        hcg = self._hcg
        sharding_flag = hcg.get_sharding_parallel_world_size() > 1
        dp_flag = hcg.get_data_parallel_world_size() > 1
        mp_flag = hcg.get_model_parallel_world_size() > 1
        pp_flag = hcg.get_pipe_parallel_world_size() > 1
        if sharding_flag and self.not_sharding_stage1 is not False:
            if mp_flag:
                comm.all_reduce(dist_sq, group=hcg.get_sharding_parallel_group())
            comm.all_reduce(not_dist_sq, group=hcg.get_sharding_parallel_group())
        if mp_flag:
            if not (dp_flag and sharding_flag):
                comm.all_reduce(dist_sq, group=hcg.get_check_parallel_group(sharding_flag))
            else:
                comm.all_reduce(dist_sq, group=hcg.get_model_parallel_group())
                if pp_flag:
                    comm.all_reduce(dist_sq, group=hcg.get_pipe_parallel_group())
        if pp_flag:
            comm.all_reduce(not_dist_sq, group=hcg.get_pipe_parallel_group())

    @torch.no_grad()
    def _dygraph_clip(self, params_grads):
        dist, not_dist = [], []
        for p, g in params_grads:
            if g is None or getattr(p, "need_clip", True) is False:
                continue
            if hasattr(p, "is_firstly_shared") and not getattr(p, "is_firstly_shared", True):
                continue
            (dist if getattr(p, "is_distributed", False) else not_dist).append(g)
        some = next((g for _, g in params_grads if g is not None), None)
        if some is None:
            return params_grads
        dev = some.device
        dist_sq = ops.squared_l2_norm(dist).to(dev) if dist else torch.zeros(1, device=dev)
        not_dist_sq = ops.squared_l2_norm(not_dist).to(dev) if not_dist else torch.zeros(1, device=dev)
        return self._comm_and_clip(params_grads, dist_sq.float(), not_dist_sq.float())

    def _comm_and_clip(self, params_grads, dist_sq, not_dist_sq):
        self._global_norm(dist_sq, not_dist_sq)
        gnorm = torch.sqrt(dist_sq + not_dist_sq)
        max_norm = torch.full_like(gnorm, self.clip_norm)
        clip_var = max_norm / (torch.maximum(gnorm, max_norm) + 1.0e-6)
        self._last_global_norm = gnorm
        tgt = self._fold_target
        # folding multiplies EVERY gradient inside the update kernel, so it is only exact when every parameter is clipped
        if tgt is not None and all(tgt._native(p) and getattr(p, "need_clip", True) for p, _ in params_grads) \
                and hasattr(tgt, "_beta1"):
            prev = tgt._grad_scale_dev
            tgt._grad_scale_dev = clip_var if prev is None else prev * clip_var
            tgt._fold_reset = True
            return params_grads
        for p, g in params_grads:
            if g is None or getattr(p, "need_clip", True) is False:
                continue
            g.mul_(clip_var.to(g.dtype))
        return params_grads

    def __getattr__(self, item):
        return getattr(self.__dict__["_clip"], item)

    def __call__(self, params_grads):
        return self._dygraph_clip(params_grads)


def unwrap_optimizer(optimizer, optimizer_instances=()):
    opt = optimizer
    while isinstance(opt, optimizer_instances):
        opt = opt._inner_opt
    return opt


class HybridParallelOptimizer:
    def __init__(self, optimizer, hcg, strategy):
        from .dygraph_sharding_optimizer import DygraphShardingOptimizer, DygraphShardingOptimizerV2
        from paddle_b200.distributed.fleet.utils.mix_precision_utils import MixPrecisionOptimizer
        self._all_params = list(_flat_params(optimizer))     # before sharding narrows the inner parameter list
        if hcg.get_sharding_parallel_world_size() > 1:
            split_param = strategy.hybrid_configs["sharding_configs"]["split_param"]
            cls = DygraphShardingOptimizerV2 if split_param else DygraphShardingOptimizer
            optimizer = cls(optimizer, hcg)
        self._inner_opt = optimizer
        self._strategy = strategy
        self._hcg = hcg
        self._use_dp_mode = hcg.get_parallel_mode() == ParallelMode.DATA_PARALLEL
        self._need_dp = hcg.get_data_parallel_world_size() > 1
        self._dp_enable = not self._use_dp_mode and self._need_dp
        self._sharding_enable = hcg.get_sharding_parallel_world_size() > 1
        self._sep_enable = hcg.get_sep_parallel_world_size() > 1
        inner = unwrap_optimizer(self._inner_opt, (MixPrecisionOptimizer, DygraphShardingOptimizer,
                                                   DygraphShardingOptimizerV2))
        self._base_opt = inner
        if isinstance(inner._grad_clip, ClipGradByGlobalNorm) and not self._use_dp_mode:
            logger.warning("While using ClipGradByGlobalNorm in TensorParallel, PipelineParallel or Sharding, the grad "
                           "clip of original optimizer will be changed.")
            inner._grad_clip = HybridParallelClipGrad(inner._grad_clip, hcg)
            inner._grad_clip._fold_target = inner
            if isinstance(self._inner_opt, (DygraphShardingOptimizer,)) and not isinstance(self._inner_opt, DygraphShardingOptimizerV2):
                inner._grad_clip.not_sharding_stage1 = True
            for item in getattr(inner, "_param_groups", []):
                if isinstance(item, dict) and item.get("grad_clip") is not None:
                    item["grad_clip"] = inner._grad_clip
        mpc = strategy.hybrid_configs["mp_configs"]
        self._mp_sync = dict(param=mpc["sync_param"], grad=mpc["sync_grad"], moment=mpc["sync_moment"],
                             mode=mpc["sync_mode"], names=list(mpc.get("sync_param_name", [])))

    # ---- mp sync of replicated params / grads / moments (reference :352-479)
    def _filter_fn(self, p):
        if getattr(p, "is_distributed", False):
            return False
        names = self._mp_sync["names"]
        return not names or any(n in p.name for n in names)

    def _mp_sync_tensors(self, tensors):
        g = self._hcg.get_model_parallel_group()
        src = self._hcg.get_model_parallel_group_src_rank()
        for t in tensors:
            if self._mp_sync["mode"] == "average":
                comm.all_reduce(t, ReduceOp.AVG, g)
            else:
                comm.broadcast(t, src, g)

    @torch.no_grad()
    def _hybrid_sync_grad(self, parameter_list):
        dp_parameter_list = parameter_list
        if self._sharding_enable:
            self._inner_opt.reduce_gradients(parameter_list, self._hcg)
            dp_parameter_list = self._inner_opt.filter_parameters(parameter_list, self._hcg)
        if self._dp_enable or self._sep_enable:
            fused_allreduce_gradients(dp_parameter_list, self._hcg)

    @torch.no_grad()
    def step(self):
        parameter_list = self._all_params
        self._hybrid_sync_grad(parameter_list)
        mp = self._hcg.get_model_parallel_world_size() > 1
        if mp and self._mp_sync["grad"]:
            self._mp_sync_tensors([_grad_of(p) for p in parameter_list if self._filter_fn(p) and _grad_of(p) is not None])
        try:
            self._inner_opt.step()
        finally:
            if getattr(self._base_opt, "_fold_reset", False):
                self._base_opt._grad_scale_dev = None
                self._base_opt._fold_reset = False
        if mp and self._mp_sync["param"]:
            self._mp_sync_tensors([p for p in parameter_list if self._filter_fn(p)])
        if mp and self._mp_sync["moment"]:
            ms = []
            for name in ("moment1", "moment2"):
                for p in parameter_list:
                    acc = self._base_opt._accumulators.get(name, {}).get(p.name)
                    if acc is not None and self._filter_fn(p):
                        ms.append(acc)
            self._mp_sync_tensors(ms)

    @torch.no_grad()
    def minimize(self, loss, startup_program=None, parameters=None, no_grad_set=None):
        self.step()
        return None, None

    def clear_grad(self, set_to_zero=True):
        self._inner_opt.clear_grad(set_to_zero)

    clear_gradients = clear_grad

    def state_dict(self):
        return self._inner_opt.state_dict()

    def set_state_dict(self, sd):
        return self._inner_opt.set_state_dict(sd)

    def __getattr__(self, item):
        return getattr(self.__dict__["_inner_opt"], item)


def _flat_params(opt):
    pl = opt._parameter_list
    if pl and isinstance(pl[0], dict):
        for g in pl:
            yield from g["params"]
    else:
        yield from pl
