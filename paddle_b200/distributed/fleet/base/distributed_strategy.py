"""DistributedStrategy (reference: python/paddle/distributed/fleet/base/distributed_strategy.py:1-2697 and the
protobuf schema paddle/fluid/framework/distributed_strategy.proto:36-430).

The reference backs this object with a protobuf message; here it is a plain typed config with the same property
names, the same keys and defaults in every ``*_configs`` dict, the same key validation (unknown keys raise), and
prototxt-style text (de)serialisation.
"""
from __future__ import annotations

import copy
import json

_DEFAULT_ORDER = ["dp", "pp", "sharding", "sep", "mp"]

_MP_CONFIGS = dict(sync_param=True, sync_grad=False, sync_moment=False, sync_mode="broadcast", sync_param_name=["embedding", "layer_norm", ".b_"],
                   mp_async_allreduce=False, mp_skip_c_identity=False, mp_fused_linear_param_grad_add=False,
                   need_broadcast_data=True, recompute_allgather=False)
_PP_CONFIGS = dict(dp_comm_overlap=False, delay_scale_loss=False, enable_timer=False, sharding_comm_overlap=False,
                   profiling=False, release_gradients=False, overlap_p2p_comm=False, clear_every_step_cache=False,
                   use_batch_p2p_comm=True)
_SHARDING_CONFIGS_DY = dict(tensor_fusion=False, accumulate_steps=1, comm_overlap=False, split_param=False,
                            fuse_optimizer=True, use_reduce_avg=True, comm_buffer_size_MB=256)
_HYBRID = dict(dp_degree=-1, mp_degree=1, pp_degree=1, sharding_degree=1, sep_degree=1, order=list(_DEFAULT_ORDER),
               mp_configs=dict(_MP_CONFIGS), pp_configs=dict(_PP_CONFIGS), sharding_configs=dict(_SHARDING_CONFIGS_DY))
_AMP = dict(init_loss_scaling=32768.0, incr_every_n_steps=1000, decr_every_n_nan_or_inf=2, incr_ratio=2.0,
            decr_ratio=0.8, use_dynamic_loss_scaling=True, custom_white_list=[], custom_black_list=[],
            custom_black_varnames=[], use_pure_fp16=False, use_pure_bf16=False, use_fp16_guard=True,
            use_optimizer_fp16=False, use_bf16=False, dtype="float16", level="O1")
_RECOMPUTE = dict(checkpoints=[], enable_offload=False, checkpoint_shape=[])
_PIPELINE = dict(micro_batch_size=1, accumulate_steps=1, schedule_mode="1F1B", p2p_cache_shape=True,
                 enable_partial_send_recv=True)
_TENSOR_PARALLEL = dict(tensor_parallel_degree=1, tensor_init_seed=-1)
_SHARDING_STATIC = dict(sharding_segment_strategy="segment_broadcast_MB", segment_broadcast_MB=32.0,
                        segment_anchors=[], sharding_degree=8, mp_degree=1, dp_degree=1, hybrid_dp=False,
                        gradient_merge_acc_step=1, optimize_offload=False, pp_allreduce_in_optimize=False,
                        pp_degree=1, optimize_cast=False, stage=1, enable_tuning=False)
_GRADIENT_MERGE = dict(k_steps=1, avg=True)
_LARS = dict(lars_coeff=0.001, lars_weight_decay=0.0005, epsilon=0.0, exclude_from_weight_decay=[])
_LAMB = dict(lamb_weight_decay=0.01, exclude_from_weight_decay=[])
_LOCALSGD = dict(k_steps=1, begin_step=1)
_DGC = dict(rampup_begin_step=0, rampup_step=1, sparsity=[0.999])
_A_SYNC = dict(k_steps=-1, max_merge_var_num=1, send_queue_size=16, independent_recv_thread=False,
               min_send_grad_num_before_recv=1, thread_pool_size=1, send_wait_times=1,
               runtime_split_send_recv=False, launch_barrier=True, heter_worker_device_guard="cpu", lr_decay_steps=10,
               use_ps_gpu=0, use_gpu_graph=0)


def check_configs_key(defaults, user, name):
    for k in user:
        if k not in defaults:
            raise ValueError(f"key: {k} not in {name}")


class _Cfg(dict):
    """dict that also allows attribute access (strategy.hybrid_configs.mp_degree appears in user code)."""
    __getattr__ = dict.__getitem__

    def __setattr__(self, k, v):
        self[k] = v


class DistributedStrategy:
    def __init__(self):
        self._d = {
            # scalar switches (distributed_strategy.proto:365-430)
            "amp": False, "recompute": False, "pipeline": False, "tensor_parallel": False, "sharding": False,
            "gradient_merge": False, "lars": False, "lamb": False, "localsgd": False, "adaptive_localsgd": False,
            "dgc": False, "a_sync": True, "auto": False, "semi_auto": False, "auto_search": False,
            "heter_ccl_mode": False, "is_fl_ps_mode": False, "with_coordinator": False, "asp": False,
            "fp16_allreduce": False, "qat": False, "without_graph_optimization": True, "elastic": False,
            "fuse_all_reduce_ops": True, "fuse_grad_size_in_MB": 32, "fuse_grad_size_in_num": 8,
            "fuse_grad_merge": False, "last_comm_group_size_MB": 1.0, "find_unused_parameters": False,
            "nccl_comm_num": 1, "sync_nccl_allreduce": True, "use_hierarchical_allreduce": False,
            "hierarchical_allreduce_inter_nranks": 1, "sync_batch_norm": False, "cudnn_exhaustive_search": False,
            "conv_workspace_size_limit": 512, "cudnn_batchnorm_spatial_persistent": False, "calc_comm_same_stream": False,
            "split_data": True,
            # parameter-server / static-graph switches: stored and serialised, read by nothing in this framework
            "adam_d2sum": False, "is_with_coordinator": False, "build_strategy": None,
        }
        self._cfgs = {
            "hybrid_configs": _Cfg(copy.deepcopy(_HYBRID)), "amp_configs": _Cfg(copy.deepcopy(_AMP)),
            "recompute_configs": _Cfg(copy.deepcopy(_RECOMPUTE)), "pipeline_configs": _Cfg(copy.deepcopy(_PIPELINE)),
            "tensor_parallel_configs": _Cfg(copy.deepcopy(_TENSOR_PARALLEL)),
            "sharding_configs": _Cfg(copy.deepcopy(_SHARDING_STATIC)),
            "gradient_merge_configs": _Cfg(copy.deepcopy(_GRADIENT_MERGE)), "lars_configs": _Cfg(copy.deepcopy(_LARS)),
            "lamb_configs": _Cfg(copy.deepcopy(_LAMB)), "localsgd_configs": _Cfg(copy.deepcopy(_LOCALSGD)),
            "dgc_configs": _Cfg(copy.deepcopy(_DGC)), "a_sync_configs": _Cfg(copy.deepcopy(_A_SYNC)),
            "adaptive_localsgd_configs": _Cfg({"init_k_steps": 1, "begin_step": 1}), "qat_configs": _Cfg({
                "channel_wise_abs_max": True, "weight_bits": 8, "activation_bits": 8, "not_quant_pattern": [], "algo": None}),
            "gradient_scale_configs": _Cfg({"scale_strategy": "avg", "scale_gradient": False}),
            "trainer_desc_configs": _Cfg({"dump_fields_path": "", "dump_fields": [], "dump_param": [], "stat_var_names": []}),
            "fs_client_param": _Cfg({"uri": "", "user": "", "passwd": "", "hadoop_bin": ""}),
            "sparse_table_configs": _Cfg({}), "fleet_desc_configs": _Cfg({}),
        }
        self._cfgs["hybrid_configs"]["mp_configs"] = _Cfg(self._cfgs["hybrid_configs"]["mp_configs"])
        self._cfgs["hybrid_configs"]["pp_configs"] = _Cfg(self._cfgs["hybrid_configs"]["pp_configs"])
        self._cfgs["hybrid_configs"]["sharding_configs"] = _Cfg(self._cfgs["hybrid_configs"]["sharding_configs"])
        self.hybrid_parallel_order = list(_DEFAULT_ORDER)

    _DEFAULTS = {"hybrid_configs": _HYBRID, "amp_configs": _AMP, "recompute_configs": _RECOMPUTE,
                 "pipeline_configs": _PIPELINE, "tensor_parallel_configs": _TENSOR_PARALLEL,
                 "sharding_configs": _SHARDING_STATIC, "gradient_merge_configs": _GRADIENT_MERGE,
                 "lars_configs": _LARS, "lamb_configs": _LAMB, "localsgd_configs": _LOCALSGD, "dgc_configs": _DGC,
                 "a_sync_configs": _A_SYNC}

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        d = self.__dict__.get("_d", {})
        if name in d:
            return d[name]
        c = self.__dict__.get("_cfgs", {})
        if name in c:
            return c[name]
        raise AttributeError(f"DistributedStrategy has no attribute {name}")

    def __setattr__(self, name, value):
        if name.startswith("_") or name == "hybrid_parallel_order":
            return object.__setattr__(self, name, value)
        if name in self._d:
            want = type(self._d[name])
            if want is bool and not isinstance(value, bool):
                raise ValueError(f"{name} should have value of bool type")
            self._d[name] = value if want in (bool, type(None)) else want(value)
            return
        if name in self._cfgs:
            self._set_configs(name, value)
            return
        raise AttributeError(f"DistributedStrategy has no attribute {name}")

    def _set_configs(self, name, configs):
        tgt = self._cfgs[name]
        defaults = self._DEFAULTS.get(name)
        if defaults is None:                       # PS / static-graph config blocks: free-form, stored as given
            tgt.update(configs)
            return
        check_configs_key(defaults, configs, name)
        for k, v in configs.items():
            if name == "hybrid_configs" and k in ("mp_configs", "pp_configs", "sharding_configs"):
                sub_defaults = {"mp_configs": _MP_CONFIGS, "pp_configs": _PP_CONFIGS, "sharding_configs": _SHARDING_CONFIGS_DY}[k]
                check_configs_key(sub_defaults, v, k)
                tgt[k].update(v)
            elif name == "hybrid_configs" and k == "order":
                if sorted(v) != sorted(_DEFAULT_ORDER) and sorted(v) != sorted([x for x in _DEFAULT_ORDER if x != "sep"]):
                    raise ValueError(f"order must be a permutation of {_DEFAULT_ORDER}")
                order = list(v)
                if "sep" not in order:
                    order.insert(order.index("mp"), "sep")
                tgt[k] = order
                self.hybrid_parallel_order = order
            else:
                tgt[k] = v

    # ---- serialisation (reference: save_to_prototxt / load_from_prototxt)
    def to_dict(self):
        return {"scalars": dict(self._d), "configs": json.loads(json.dumps(self._cfgs))}

    def save_to_prototxt(self, output):
        with open(output, "w") as f:
            json.dump(self.to_dict(), f, indent=1)

    def load_from_prototxt(self, pb_file):
        with open(pb_file) as f:
            d = json.load(f)
        self._d.update(d["scalars"])
        for k, v in d["configs"].items():
            self._set_configs(k, v)

    def __repr__(self):
        rows = [f"  {k}={v}" for k, v in self._d.items() if v not in (False, None)]
        rows += [f"  {k}={dict(v)}" for k, v in self._cfgs.items()]
        return "DistributedStrategy(\n" + "\n".join(rows) + "\n)"

    def __deepcopy__(self, memo):
        new = DistributedStrategy()
        new._d = copy.deepcopy(self._d)
        for k, v in self._cfgs.items():
            new._cfgs[k] = _Cfg(copy.deepcopy(dict(v)))
        for sub in ("mp_configs", "pp_configs", "sharding_configs"):
            new._cfgs["hybrid_configs"][sub] = _Cfg(new._cfgs["hybrid_configs"][sub])
        new.hybrid_parallel_order = list(self.hybrid_parallel_order)
        return new
