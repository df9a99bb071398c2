"""Model wrappers selected by ParallelMode (reference: python/paddle/distributed/fleet/meta_parallel/
{meta_parallel_base,tensor_parallel,sharding_parallel,segment_parallel}.py)."""
from __future__ import annotations

import torch

from paddle_b200.distributed.fleet.utils.hybrid_parallel_util import (broadcast_dp_parameters, broadcast_input_data,
                                                                      broadcast_mp_parameters,
                                                                      broadcast_sep_parameters,
                                                                      broadcast_sharding_parameters)
from paddle_b200.nn.layer.layers import Layer


class MetaParallelBase(Layer):
    def __init__(self, layers, hcg, strategy):
        super().__init__()
        self._layers = layers
        self._hcg = hcg
        self._strategy = strategy
        self._prepare_for_model()

    def _prepare_for_model(self):
        pass

    def _pre_forward(self, *inputs, **kwargs):
        pass

    def _post_forward(self, output):
        pass

    def forward(self, *inputs, **kwargs):
        self._pre_forward(*inputs, **kwargs)
        out = self._layers(*inputs, **kwargs)
        self._post_forward(out)
        return out

    def state_dict(self, *a, **k):
        return self._layers.state_dict(*a, **k)

    def set_state_dict(self, sd, use_structured_name=True):
        return self._layers.set_state_dict(sd, use_structured_name)

    def parameters(self, *a, **k):
        return self._layers.parameters(*a, **k)

    def named_parameters(self, *a, **k):
        return self._layers.named_parameters(*a, **k)


class TensorParallel(MetaParallelBase):
    def _prepare_for_model(self):
        broadcast_mp_parameters(self._layers, self._hcg)
        if self._hcg.get_sep_parallel_world_size() > 1:
            broadcast_sep_parameters(self._layers, self._hcg)
        if self._hcg.get_sharding_parallel_world_size() > 1:
            broadcast_sharding_parameters(self._layers, self._hcg)
        broadcast_dp_parameters(self._layers, self._hcg)

    def _pre_forward(self, *inputs, **kwargs):
        need = self._strategy.hybrid_configs["mp_configs"].get("need_broadcast_data", True) if self._strategy else True
        if need:
            broadcast_input_data(self._hcg, *inputs, **kwargs)


class ShardingParallel(MetaParallelBase):
    def _prepare_for_model(self):
        broadcast_sharding_parameters(self._layers, self._hcg)
        broadcast_dp_parameters(self._layers, self._hcg)


class SegmentParallel(MetaParallelBase):
    def _prepare_for_model(self):
        broadcast_sep_parameters(self._layers, self._hcg)
        if self._hcg.get_sharding_parallel_world_size() > 1:
            broadcast_sharding_parameters(self._layers, self._hcg)
        broadcast_dp_parameters(self._layers, self._hcg)
