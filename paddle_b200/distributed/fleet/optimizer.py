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
    hp_optim = HybridParallelOptimizer(optimizer, hcg, st)
    ppc = st.hybrid_configs["pp_configs"] if st is not None else {}
    # the pipeline engine reduces these gradients itself, overlapped with the backward pass (reference :80-92)
    if ppc.get("dp_comm_overlap") if isinstance(ppc, dict) else getattr(ppc, "dp_comm_overlap", False):
        hp_optim._dp_enable = False
    if ppc.get("sharding_comm_overlap") if isinstance(ppc, dict) else getattr(ppc, "sharding_comm_overlap", False):
        hp_optim._sharding_enable = False
    return hp_optim
