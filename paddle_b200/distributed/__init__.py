"""paddle.distributed public namespace (reference: python/paddle/distributed/__init__.py:17-176)."""
from . import collective, communication, watchdog  # noqa: F401
from .collective import (Group, ReduceOp, barrier, destroy_process_group, get_backend, get_group, is_available,  # noqa: F401
                         is_initialized, new_group, wait)
from .communication import (P2POp, all_gather, all_gather_object, all_reduce, all_to_all, alltoall, alltoall_single,  # noqa: F401
                            batch_isend_irecv, broadcast, broadcast_object_list, gather, gather_object, irecv, isend,
                            recv, reduce, reduce_scatter, scatter, scatter_object_list, send, stream)
from .parallel import DataParallel, ParallelEnv, get_rank, get_world_size, init_parallel_env  # noqa: F401
from .spawn import spawn  # noqa: F401
from . import fleet  # noqa: F401
from . import checkpoint, launch, sharding, utils  # noqa: F401
from .checkpoint import load_state_dict, save_state_dict  # noqa: F401
from .fleet.layers.mpu.mp_ops import split  # noqa: F401
from .auto_parallel import (DistAttr, DistModel, Partial, Placement, ProcessMesh, Replicate, Shard, Strategy,  # noqa: F401
                            dtensor_from_fn, reshard, shard_dataloader, shard_layer, shard_optimizer, shard_tensor,
                            to_static, unshard_dtensor, ShardingStage1, ShardingStage2, ShardingStage3)
from .sharding import group_sharded_parallel, save_group_sharded_model  # noqa: F401
from .utils.moe_utils import global_gather, global_scatter  # noqa: F401


class ParallelMode:
    DATA_PARALLEL = 0
    TENSOR_PARALLEL = 1
    PIPELINE_PARALLEL = 2
    SHARDING_PARALLEL = 3
