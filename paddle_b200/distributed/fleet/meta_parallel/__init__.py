"""paddle.distributed.fleet.meta_parallel (reference: python/paddle/distributed/fleet/meta_parallel/__init__.py)."""
from ..layers.mpu import (ColumnParallelLinear, ParallelCrossEntropy, RNGStatesTracker, RowParallelLinear,  # noqa: F401
                          VocabParallelEmbedding, get_rng_state_tracker, model_parallel_random_seed)
from .meta_parallel_base import MetaParallelBase, SegmentParallel, ShardingParallel, TensorParallel  # noqa: F401
from .parallel_layers import LayerDesc, PipelineLayer, SharedLayerDesc  # noqa: F401
from .pipeline_parallel import (PipelineParallel, PipelineParallelWithInterleave,  # noqa: F401
                                PipelineParallelWithInterleaveFthenB)
from . import sharding  # noqa: F401
