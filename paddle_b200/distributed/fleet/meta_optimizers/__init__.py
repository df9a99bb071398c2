from . import dygraph_optimizer  # noqa: F401
from .dygraph_optimizer import (DygraphShardingOptimizer, DygraphShardingOptimizerV2, HybridParallelGradScaler,  # noqa: F401
                                HybridParallelOptimizer)
