"""paddle.distributed.fleet (reference: python/paddle/distributed/fleet/__init__.py)."""
from . import base, layers, meta_optimizers, meta_parallel, recompute as _recompute_pkg, utils  # noqa: F401
from .base.distributed_strategy import DistributedStrategy  # noqa: F401
from .base.role_maker import PaddleCloudRoleMaker, Role, UserDefinedRoleMaker  # noqa: F401
from .base.topology import CommunicateTopology, HybridCommunicateGroup, ParallelMode  # noqa: F401
from .fleet import Fleet, fleet  # noqa: F401
from .layers.mpu.random import get_rng_state_tracker  # noqa: F401
from .recompute import recompute, recompute_hybrid, recompute_sequential  # noqa: F401

init = fleet.init
distributed_model = fleet.distributed_model
distributed_optimizer = fleet.distributed_optimizer
distributed_scaler = fleet.distributed_scaler
get_hybrid_communicate_group = fleet.get_hybrid_communicate_group
worker_index = fleet.worker_index
worker_num = fleet.worker_num
node_num = fleet.node_num
local_rank = fleet.local_rank
is_first_worker = fleet.is_first_worker
is_worker = fleet.is_worker
is_server = fleet.is_server
worker_endpoints = fleet.worker_endpoints
barrier_worker = fleet.barrier_worker
init_worker = fleet.init_worker
stop_worker = fleet.stop_worker
collective_perf = fleet.collective_perf
