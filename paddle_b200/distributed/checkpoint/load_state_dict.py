"""dist.load_state_dict (reference: python/paddle/distributed/checkpoint/load_state_dict.py:32-557): reads all
``.metadata``, computes for every target shard its overlaps with the stored shards and fills the target in place -
i.e. it reshards across different parallel layouts (save mp=2, load mp=4).  All checkpoint files are visible to
every rank on this single-node target, so each rank reads what it needs directly."""
from __future__ import annotations

import os
import pickle

import torch

from paddle_b200.distributed import collective as C
from paddle_b200.framework import io as _io

from .save_state_dict import shard_info
from .utils import flatten_state_dict


def _load_metadata(path):
    def uid(f):                                  # "<unique id>.metadata": numeric order, so save 10 comes after save 2
        stem = f[: -len(".metadata")]
        return (0, int(stem)) if stem.isdigit() else (1, stem)
    files = sorted((f for f in os.listdir(path) if f.endswith(".metadata")), key=uid)
    assert files, f"no .metadata file found under {path}"
    mds = []
    for f in files:
        with open(os.path.join(path, f), "rb") as fh:
            mds.append(pickle.load(fh))
    return mds


def compute_overlap(cur_offset, cur_shape, st_offset, st_shape):
    """-> (cur_slices, stored_slices) or None when the boxes do not intersect."""
    cur_sl, st_sl = [], []
    for co, cs, so, ss in zip(cur_offset, cur_shape, st_offset, st_shape):
        lo, hi = max(co, so), min(co + cs, so + ss)
        if lo >= hi:
            return None
        cur_sl.append(slice(lo - co, hi - co))
        st_sl.append(slice(lo - so, hi - so))
    return tuple(cur_sl), tuple(st_sl)


def load_state_dict(state_dict, path, process_group=None, coordinator_rank=0):
    assert isinstance(state_dict, dict), "The state_dict should be a dictionary."
    flat, _ = flatten_state_dict(state_dict)
    mds = _load_metadata(path)
    cache = {}

    def read(fname):
        if fname not in cache:
            cache[fname] = _io.load(os.path.join(path, fname))
        return cache[fname]

    with torch.no_grad():
        for key, target in flat.items():
            if not isinstance(target, torch.Tensor):
                continue
            gshape, goff = shard_info(target)
            if hasattr(target, "placements"):
                # an auto_parallel DistTensor: fill its local shard (ops on the tensor itself would be global ops)
                from paddle_b200.distributed.auto_parallel.dtensor import local_of
                target = local_of(target)
            filled = False
            # coverage of the target shard: every element must come from some stored shard (the newest save of a key
            # wins, older ones only fill what it does not cover)
            covered = torch.zeros(tuple(target.shape), dtype=torch.bool) if target.dim() > 0 else None
            for md in reversed(mds):
                if key not in md.state_dict_metadata:
                    continue
                for ltm in md.state_dict_metadata[key]:
                    ov = compute_overlap(goff, tuple(target.shape), tuple(ltm.global_offset), tuple(ltm.local_shape))
                    if ov is None and target.dim() > 0:
                        continue
                    fname = None
                    for idx, fn in md.storage_metadata.items():
                        if idx.tensor_key == key and tuple(idx.global_offset) == tuple(ltm.global_offset):
                            fname = fn
                            break
                    assert fname is not None, f"storage for {key}@{ltm.global_offset} not found"
                    stored = read(fname)[key]
                    if target.dim() == 0:
                        target.copy_(stored.to(target.device, target.dtype))
                    else:
                        cur_sl, st_sl = ov
                        fresh = ~covered[cur_sl]
                        if bool(fresh.all()):
                            target[cur_sl].copy_(stored[st_sl].to(target.device, target.dtype))
                        elif bool(fresh.any()):
                            piece = target[cur_sl]
                            m = fresh.to(target.device)
                            piece[m] = stored[st_sl].to(target.device, target.dtype)[m]
                        covered[cur_sl] = True
                    filled = True
            assert filled, f"key {key} not found in checkpoint {path}"
            if covered is not None:
                assert bool(covered.all()), (f"checkpoint {path} covers only {int(covered.sum())} of {covered.numel()} "
                                             f"elements of {key} (target offset {goff}, shape {tuple(target.shape)})")
    if C.is_initialized() and C.get_world_size() > 1:
        C.barrier(process_group)
