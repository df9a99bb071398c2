from .dygraph_sharding_optimizer import DygraphShardingOptimizer, DygraphShardingOptimizerV2  # noqa: F401
from .hybrid_parallel_gradscaler import HybridParallelGradScaler  # noqa: F401
from .hybrid_parallel_optimizer import HybridParallelClipGrad, HybridParallelOptimizer  # noqa: F401
