"""HybridParallelGradScaler (reference: .../dygraph_optimizer/hybrid_parallel_gradscaler.py): found_inf MAX
all-reduce over the whole job so every rank takes the same skip decision."""
from __future__ import annotations

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp


class HybridParallelGradScaler:
    def __init__(self, scaler, hcg):
        self._scaler = scaler
        self._hcg = hcg
        self._use_dp_mode = False
        outer = self

        def _sync(optimizer):
            if hcg is not None and hcg.nranks > 1 and scaler._found_inf is not None:
                comm.all_reduce(scaler._found_inf, ReduceOp.MAX, group=None)
        scaler._sync_found_inf = _sync

    def scale(self, var):
        return self._scaler.scale(var)

    def minimize(self, optimizer, *args, **kwargs):
        self._scaler.step(optimizer)
        self._scaler.update()

    def step(self, optimizer):
        return self._scaler.step(optimizer)

    def update(self):
        return self._scaler.update()

    def __getattr__(self, item):
        return getattr(self.__dict__["_scaler"], item)
