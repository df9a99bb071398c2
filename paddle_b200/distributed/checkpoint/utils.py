"""flatten/unflatten of nested state dicts (reference: python/paddle/distributed/checkpoint/utils.py:65-109)."""
from __future__ import annotations

import torch


def flatten_state_dict(state_dict):
    flat, mapping = {}, {}

    def rec(obj, path):
        if isinstance(obj, dict):
            for k, v in obj.items():
                rec(v, path + (k,))
        elif isinstance(obj, torch.Tensor):
            key = ".".join(str(p) for p in path)
            if key in flat:
                raise ValueError(f"duplicate flattened key {key}")
            flat[key] = obj
            mapping[key] = path
        else:
            key = ".".join(str(p) for p in path)
            flat[key] = obj
            mapping[key] = path
    rec(state_dict, ())
    return flat, mapping


def unflatten_state_dict(flat, mapping):
    out = {}
    for key, val in flat.items():
        path = mapping[key]
        d = out
        for p in path[:-1]:
            d = d.setdefault(p, {})
        d[path[-1]] = val
    return out


def compute_local_shape_and_global_offset(global_shape, placements_desc):
    """placements_desc: list of (mesh_dim_size, coord, shard_dim or None)."""
    shape, off = list(global_shape), [0] * len(global_shape)
    for size, coord, dim in placements_desc:
        if dim is None:
            continue
        assert shape[dim] % size == 0
        shape[dim] //= size
        off[dim] += coord * shape[dim]
    return tuple(shape), tuple(off)
