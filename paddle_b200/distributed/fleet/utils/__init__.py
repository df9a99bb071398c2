"""paddle.distributed.fleet.utils (reference: python/paddle/distributed/fleet/utils/__init__.py)."""
from . import hybrid_parallel_util, log_util, mix_precision_utils, sequence_parallel_utils, tensor_fusion_helper, timer_helper  # noqa: F401
from ..recompute import recompute  # noqa: F401
from .fs import DistributedInfer, HDFSClient, LocalFS  # noqa: F401
