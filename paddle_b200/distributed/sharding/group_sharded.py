"""group_sharded_parallel / save_group_sharded_model (reference: python/paddle/distributed/sharding/group_sharded.py:39-278)."""
from __future__ import annotations

import os

import torch

from paddle_b200.distributed.fleet.meta_parallel.sharding import (GroupShardedOptimizerStage2, GroupShardedScaler,
                                                                  GroupShardedStage2, GroupShardedStage3)
from paddle_b200.framework import io as _io


def group_sharded_parallel(model, optimizer, level, scaler=None, group=None, offload=False, sync_buffers=False,
                           buffer_max_size=2 ** 23, segment_size=2 ** 20, sync_comm=False, dp_group=None,
                           exclude_layer=None):
    assert isinstance(model, torch.nn.Module), "The model must be the instance of paddle.nn.Layer."
    assert level in ("os", "os_g", "p_g_os"), "The level must be os, os_g or p_g_os."
    if level in ("os", "os_g"):
        optimizer = GroupShardedOptimizerStage2(params=optimizer._parameter_list, optim=optimizer, group=group,
                                                offload=offload, dp_group=dp_group)
        model = GroupShardedStage2(model, optimizer, group=group, sync_buffers=sync_buffers,
                                   buffer_max_size=buffer_max_size, dp_group=dp_group)
    else:
        model = GroupShardedStage3(model, optimizer=optimizer, group=group, sync_buffers=sync_buffers,
                                   segment_size=segment_size, offload=offload, sync_comm=sync_comm, dp_group=dp_group,
                                   exclude_layer=exclude_layer)
    if scaler is not None:
        scaler = GroupShardedScaler(scaler)
    return model, optimizer, scaler


def save_group_sharded_model(model, output, optimizer=None):
    assert not os.path.isfile(output), f"Saving directory ({output}) should be a directory, not a file"
    os.makedirs(output, exist_ok=True)
    if isinstance(model, GroupShardedStage3):
        model.get_all_parameters(convert2cpu=False)
        _io.save(model._layer.state_dict(), os.path.join(output, "model.pdmodel"))
        model.release_all_parameters()
    elif isinstance(model, GroupShardedStage2):
        _io.save(model._layer.state_dict(), os.path.join(output, "model.pdmodel"))
    else:
        raise ValueError("Please use the layer which is wrapped with group_sharded_parallel.")
    if optimizer is not None:
        assert hasattr(optimizer, "_optim") or hasattr(optimizer, "state_dict")
        inner = getattr(optimizer, "_optim", optimizer)
        _io.save(inner.state_dict(), os.path.join(output, "model.pdopt"))
