"""Hybrid-parallel helpers (reference: python/paddle/distributed/fleet/utils/hybrid_parallel_util.py:86-287):
parameter broadcast over mp/sep/sharding/dp groups, input broadcast inside the mp group, fused gradient
all-reduce for the dp (or dp x sep) axis."""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.distributed.parallel import sync_params_buffers

_FUSE_BYTES = 128 << 20


def _apply_collective_grads(parameters, comm_group, bucket_size=_FUSE_BYTES, scale=None):
    """concat-free fused all-reduce: grads are coalesced per dtype into <=128 MB flat buffers."""
    grads = []
    for p in parameters:
        if not p.requires_grad:
            continue
        g = getattr(p, "main_grad", None)
        g = g if g is not None else p.grad
        if g is not None:
            grads.append(g)
    if not grads or comm_group is None or comm_group.nranks == 1:
        return
    nranks = comm_group.nranks
    scale = (1.0 / nranks) if scale is None else scale
    by_dtype = {}
    for g in grads:
        by_dtype.setdefault(g.dtype, []).append(g)
    with torch.no_grad():
        for lst in by_dtype.values():
            chunk, size = [], 0
            for g in lst + [None]:
                if g is not None:
                    chunk.append(g)
                    size += g.numel() * g.element_size()
                if chunk and (g is None or size >= bucket_size):
                    if len(chunk) == 1:
                        flat = chunk[0].reshape(-1)
                        if scale != 1.0:
                            flat.mul_(scale)
                        comm.all_reduce(flat, ReduceOp.SUM, comm_group)
                    else:
                        flat = torch.cat([c.reshape(-1) for c in chunk])
                        if scale != 1.0:
                            flat.mul_(scale)
                        comm.all_reduce(flat, ReduceOp.SUM, comm_group)
                        off = 0
                        for c in chunk:
                            c.copy_(flat[off:off + c.numel()].view_as(c))
                            off += c.numel()
                    chunk, size = [], 0


def fused_allreduce_gradients_with_group(parameter_list, group, bucket_size=_FUSE_BYTES, scale=None):
    _apply_collective_grads(parameter_list, group, bucket_size, scale)


def fused_allreduce_gradients(parameter_list, hcg):
    """dp gradient sync of the hybrid path (scaled by 1/dp); sep axis is summed unscaled (Appendix B.4)."""
    if hcg is None:
        from paddle_b200.distributed import collective as C
        return _apply_collective_grads(parameter_list, C._get_default_group())
    group, scale = hcg.get_data_parallel_group(), None
    if hcg.get_sep_parallel_world_size() > 1:
        group = hcg.get_dp_sep_parallel_group()
        scale = 1.0 / hcg.get_data_parallel_world_size()
    _apply_collective_grads(parameter_list, group, scale=scale)


def broadcast_input_data(hcg, *inputs, **kwargs):
    g = hcg.get_model_parallel_group()
    src = hcg.get_model_parallel_group_src_rank()
    dev = "cuda" if torch.cuda.is_available() and g.process_group is not None and \
        str(getattr(g, "backend", "")).lower() == "nccl" else None
    for v in list(inputs) + list(kwargs.values()):
        if isinstance(v, torch.Tensor):
            with torch.no_grad():
                comm.broadcast(v, src, g)
    return inputs, kwargs


def broadcast_mp_parameters(model, hcg):
    sync_params_buffers(model, hcg.get_model_parallel_group(), hcg.get_model_parallel_group_src_rank(),
                        is_model_parallel=True, fuse_params=False)


def broadcast_dp_parameters(model, hcg):
    sync_params_buffers(model, hcg.get_data_parallel_group(), hcg.get_data_parallel_group_src_rank(),
                        is_model_parallel=False, fuse_params=False)


def broadcast_sharding_parameters(model, hcg):
    sync_params_buffers(model, hcg.get_sharding_parallel_group(), hcg.get_sharding_parallel_group_src_rank(),
                        is_model_parallel=False, fuse_params=False)


def broadcast_sep_parameters(model, hcg):
    sync_params_buffers(model, hcg.get_sep_parallel_group(), hcg.get_sep_parallel_group().ranks[0],
                        is_model_parallel=False, fuse_params=False)


def sharding_reduce_gradients(parameter_list, hcg):
    """ZeRO-1 without tensor fusion: average every grad over the sharding group."""
    g = hcg.get_sharding_parallel_group()
    with torch.no_grad():
        for p in parameter_list:
            gr = getattr(p, "main_grad", None)
            gr = gr if gr is not None else p.grad
            if p.requires_grad and gr is not None:
                comm.all_reduce(gr, ReduceOp.AVG, g)
