"""fleet.distributed_optimizer (reference: python/paddle/distributed/fleet/optimizer.py:65-92)."""
from __future__ import annotations


def distributed_optimizer(optimizer, strategy=None):
    from .fleet import fleet as fleet_env
    hcg = fleet_env._hcg
    st = strategy or fleet_env._user_defined_strategy
    if hcg is None:
        raise RuntimeError("call fleet.init(is_collective=True) before fleet.distributed_optimizer")
    if hcg.nranks == 1:
        return optimizer
    from .meta_optimizers.dygraph_optimizer import HybridParallelOptimizer
    return HybridParallelOptimizer(optimizer, hcg, st)
