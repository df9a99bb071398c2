"""fleet.distributed_scaler (reference: python/paddle/distributed/fleet/scaler.py + meta_optimizers/dygraph_optimizer/
hybrid_parallel_gradscaler.py): found_inf is all-reduced (MAX) over the check group so every rank skips together."""
from __future__ import annotations

import types

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp


def distributed_scaler(scaler):
    from .fleet import fleet as fleet_env
    hcg = fleet_env._hcg

    def _sync_found_inf(self, optimizer):
        if hcg is None or hcg.nranks == 1:
            return
        comm.all_reduce(self._found_inf, ReduceOp.MAX, group=None)
    scaler._sync_found_inf = types.MethodType(_sync_found_inf, scaler)
    return scaler
