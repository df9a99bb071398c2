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
from . import checkpoint, launch, rpc, sharding, utils  # noqa: F401
from .checkpoint import load_state_dict, save_state_dict  # noqa: F401
from .fleet.layers.mpu.mp_ops import split  # noqa: F401
from .auto_parallel import (DistAttr, DistModel, Partial, Placement, ProcessMesh, Replicate, Shard, Strategy,  # noqa: F401
                            dtensor_from_fn, reshard, shard_dataloader, shard_layer, shard_optimizer, shard_tensor,
                            to_static, unshard_dtensor, ShardingStage1, ShardingStage2, ShardingStage3, ReduceType,
                            create_mesh, dtensor_from_local, dtensor_to_local, get_mesh, set_mesh, shard_scaler)
from .sharding import group_sharded_parallel, save_group_sharded_model  # noqa: F401
from .utils.moe_utils import global_gather, global_scatter  # noqa: F401


class ParallelMode:
    DATA_PARALLEL = 0
    TENSOR_PARALLEL = 1
    PIPELINE_PARALLEL = 2
    SHARDING_PARALLEL = 3


# ---- CPU-side gloo helpers (reference: python/paddle/distributed/parallel_with_gloo.py) ---------------------------
_gloo_pg = [None]


def gloo_init_parallel_env(rank_id, rank_num, server_endpoint):
    """Stand-alone gloo group over TCP for CPU barriers outside of init_parallel_env."""
    import datetime

    import torch.distributed as tdist
    host, port = server_endpoint.split(":")
    store = tdist.TCPStore(host, int(port), rank_num, rank_id == 0, timeout=datetime.timedelta(seconds=300))
    _gloo_pg[0] = tdist.ProcessGroupGloo(store, rank_id, rank_num, datetime.timedelta(seconds=300))


def gloo_barrier():
    assert _gloo_pg[0] is not None, "call gloo_init_parallel_env first"
    _gloo_pg[0].barrier().wait()


def gloo_release():
    _gloo_pg[0] = None


class _PSOnly:
    """Parameter-server data feeding (InMemoryDataset / QueueDataset and the sparse-table entry policies) belongs to the
    PS runtime, which is outside this collective-training framework."""

    def __init__(self, *a, **k):
        raise NotImplementedError(f"{type(self).__name__}: {_PSOnly.__doc__}")


class InMemoryDataset(_PSOnly):
    pass


class QueueDataset(_PSOnly):
    pass


class CountFilterEntry(_PSOnly):
    pass


class ShowClickEntry(_PSOnly):
    pass


class ProbabilityEntry(_PSOnly):
    pass


from . import io  # noqa: E402,F401
