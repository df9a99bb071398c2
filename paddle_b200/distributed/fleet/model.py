"""fleet.distributed_model (reference: python/paddle/distributed/fleet/model.py:30-178)."""
from __future__ import annotations

from paddle_b200.distributed.parallel import DataParallel

from .base.topology import ParallelMode


def distributed_model(model):
    from .fleet import fleet as fleet_env
    hcg = fleet_env._hcg
    strategy = fleet_env._user_defined_strategy
    if hcg is None:
        raise RuntimeError("call fleet.init(is_collective=True) before fleet.distributed_model")
    if strategy.amp:
        from paddle_b200 import amp
        cfg = strategy.amp_configs
        level = "O2" if cfg["use_pure_fp16"] or cfg["use_pure_bf16"] else "O1"
        dtype = "bfloat16" if cfg["use_pure_bf16"] or cfg.get("use_bf16") else cfg.get("dtype", "float16")
        if level == "O2":
            model = amp.decorate(models=model, level="O2", dtype=dtype)
    mode = hcg.get_parallel_mode()
    from .meta_parallel import (PipelineLayer, PipelineParallel, PipelineParallelWithInterleave,
                                PipelineParallelWithInterleaveFthenB, SegmentParallel, ShardingParallel, TensorParallel)
    if mode == ParallelMode.SHARDING_PARALLEL:
        return ShardingParallel(model, hcg, strategy=strategy)
    if mode == ParallelMode.DATA_PARALLEL:
        if hcg.get_data_parallel_world_size() == 1:
            return model
        return DataParallel(model, comm_buffer_size=strategy.fuse_grad_size_in_MB,
                            last_comm_buffer_size=strategy.last_comm_group_size_MB,
                            find_unused_parameters=strategy.find_unused_parameters,
                            group=hcg.get_data_parallel_group())
    if mode == ParallelMode.SEGMENT_PARALLEL:
        return SegmentParallel(model, hcg, strategy=strategy)
    if mode == ParallelMode.TENSOR_PARALLEL:
        return TensorParallel(model, hcg, strategy=strategy)
    if mode == ParallelMode.PIPELINE_PARALLEL:
        assert isinstance(model, PipelineLayer), "For pipeline parallel, the model should an instance of PipelineLayer"
        if model.get_num_virtual_stages() == 1:
            return PipelineParallel(model, hcg, strategy=strategy)
        acc = strategy.pipeline_configs["accumulate_steps"]
        pp = hcg.get_pipe_parallel_world_size()
        if pp <= acc < 2 * pp:
            return PipelineParallelWithInterleaveFthenB(model, hcg, strategy=strategy)
        return PipelineParallelWithInterleave(model, hcg, strategy=strategy)
    return model
