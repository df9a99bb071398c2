"""dist.save_state_dict (reference: python/paddle/distributed/checkpoint/save_state_dict.py:76-241).

Every rank writes ``{rank}_{uid}.distcp`` = paddle.save({key: local_tensor}); the coordinator writes
``{uid}.metadata`` with each tensor's shards (global offset + local shape) and which file holds them; replicated
tensors are de-duplicated to the lowest rank.  A tensor is "sharded" when it carries ``_dist_meta`` =
(global_shape, global_offset) - set by the TP layers' ``is_distributed``/``split_axis`` via ``shard_info`` or by
``paddle.distributed.shard_tensor``."""
from __future__ import annotations

import os

import torch

from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm
from paddle_b200.framework import io as _io

from .metadata import LocalTensorIndex, LocalTensorMetadata, Metadata
from .utils import flatten_state_dict


def shard_info(t):
    """-> (global_shape, global_offset) of a local tensor."""
    meta = getattr(t, "_dist_meta", None)
    if meta is not None:
        return tuple(meta[0]), tuple(meta[1])
    if getattr(t, "is_distributed", False) and hasattr(t, "split_axis"):
        from paddle_b200.distributed import fleet
        hcg = fleet.get_hybrid_communicate_group()
        if hcg is not None and hcg.get_model_parallel_world_size() > 1:
            n, r, ax = hcg.get_model_parallel_world_size(), hcg.get_model_parallel_rank(), t.split_axis
            gshape = list(t.shape)
            gshape[ax] *= n
            off = [0] * t.dim()
            off[ax] = r * t.shape[ax]
            return tuple(gshape), tuple(off)
    return tuple(t.shape), tuple([0] * t.dim())


def check_file_name(file_name, process_group):
    names = []
    comm.all_gather_object(names, file_name, process_group)
    assert all(n == names[0] for n in names), f"id:{file_name} in different rank have different ids"


def _unique_id(path, process_group):
    uid = -1
    if os.path.isdir(path):
        for f in os.listdir(path):
            if f.endswith(".metadata"):
                try:
                    uid = max(uid, int(f.split(".")[0]))
                except ValueError:
                    pass
    ids = []
    comm.all_gather_object(ids, uid + 1, process_group)
    return max(ids)


def save_state_dict(state_dict, path, process_group=None, coordinator_rank=0):
    assert isinstance(state_dict, dict), "The state_dict should be a dictionary."
    flat, mapping = flatten_state_dict(state_dict)
    if len(flat) > 0:
        for v in flat.values():
            assert isinstance(v, torch.Tensor), "The value of state_dict should be a paddle.Tensor."
    os.makedirs(path, exist_ok=True)
    rank = C.get_rank()
    multi = C.is_initialized() and C.get_world_size() > 1
    uid = _unique_id(path, process_group) if multi else _unique_id_local(path)
    file_name = f"{rank}_{uid}.distcp"
    local_meta, local_storage, local_tensors = {}, {}, {}
    for key, val in flat.items():
        gshape, goff = shard_info(val)
        replicated = tuple(gshape) == tuple(val.shape)
        local_meta[key] = (LocalTensorMetadata(goff, tuple(val.shape), str(val.dtype).replace("torch.", "")), replicated)
        local_storage[LocalTensorIndex(key, goff)] = file_name
        if hasattr(val, "placements"):
            from paddle_b200.distributed.auto_parallel.dtensor import local_of
            val = local_of(val).detach()
        local_tensors[key] = val
    gathered = []
    if multi:
        comm.all_gather_object(gathered, (rank, local_meta, local_storage), process_group)
    else:
        gathered = [(rank, local_meta, local_storage)]
    # de-duplicate: a (key, offset) shard is written by the lowest rank that holds it
    owner = {}
    for r, meta, storage in sorted(gathered, key=lambda x: x[0]):
        for idx, fname in storage.items():
            owner.setdefault(idx, (r, fname))
    to_write = {k: v for k, v in local_tensors.items()
                if owner[LocalTensorIndex(k, local_meta[k][0].global_offset)][0] == rank}
    _io.save(to_write, os.path.join(path, file_name))
    if rank == coordinator_rank:
        md = Metadata()
        for r, meta, storage in sorted(gathered, key=lambda x: x[0]):
            for key, (ltm, _) in meta.items():
                lst = md.state_dict_metadata.setdefault(key, [])
                if not any(x.global_offset == ltm.global_offset for x in lst):
                    lst.append(ltm)
        md.storage_metadata = {idx: fname for idx, (r, fname) in owner.items()}
        md.flat_mapping = mapping
        _io.save(md, os.path.join(path, f"{uid}.metadata"))
    if multi:
        C.barrier(process_group)


def _unique_id_local(path):
    uid = -1
    for f in os.listdir(path):
        if f.endswith(".metadata"):
            try:
                uid = max(uid, int(f.split(".")[0]))
            except ValueError:
                pass
    return uid + 1
