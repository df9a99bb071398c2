"""paddle.geometric (reference: python/paddle/geometric/{math,message_passing/send_recv}.py): segment reductions and
graph message passing, as index_reduce / scatter operations."""
from __future__ import annotations

import torch


def _segment(data, segment_ids, reduce):
    ids = segment_ids.long()
    n = int(ids.max().item()) + 1 if ids.numel() else 0
    out = torch.zeros((n,) + tuple(data.shape[1:]), dtype=data.dtype, device=data.device)
    if n == 0:
        return out
    idx = ids.reshape((-1,) + (1,) * (data.dim() - 1)).expand_as(data)
    if reduce == "sum":
        return out.scatter_add_(0, idx, data)
    if reduce == "mean":
        return out.scatter_reduce_(0, idx, data, "mean", include_self=False)
    return out.scatter_reduce_(0, idx, data, "amax" if reduce == "max" else "amin", include_self=False)


def segment_sum(data, segment_ids, name=None):
    return _segment(data, segment_ids, "sum")


def segment_mean(data, segment_ids, name=None):
    return _segment(data, segment_ids, "mean")


def segment_max(data, segment_ids, name=None):
    return _segment(data, segment_ids, "max")


def segment_min(data, segment_ids, name=None):
    return _segment(data, segment_ids, "min")


def _reduce_to(msg, dst_index, reduce_op, out_size, like):
    n = int(out_size) if out_size is not None else like.shape[0]
    out = torch.zeros((n,) + tuple(msg.shape[1:]), dtype=msg.dtype, device=msg.device)
    idx = dst_index.long().reshape((-1,) + (1,) * (msg.dim() - 1)).expand_as(msg)
    if reduce_op == "sum":
        return out.scatter_add_(0, idx, msg)
    return out.scatter_reduce_(0, idx, msg, {"mean": "mean", "max": "amax", "min": "amin"}[reduce_op], include_self=False)


def send_u_recv(x, src_index, dst_index, reduce_op="sum", out_size=None, name=None):
    """Gather node features along edges and reduce them at the destination nodes."""
    return _reduce_to(x[src_index.long()], dst_index, reduce_op, out_size, x)


def send_ue_recv(x, y, src_index, dst_index, message_op="add", reduce_op="sum", out_size=None, name=None):
    m = x[src_index.long()]
    m = {"add": m + y, "sub": m - y, "mul": m * y, "div": m / y}[message_op]
    return _reduce_to(m, dst_index, reduce_op, out_size, x)


def send_uv(x, y, src_index, dst_index, message_op="add", name=None):
    a, b = x[src_index.long()], y[dst_index.long()]
    return {"add": a + b, "sub": a - b, "mul": a * b, "div": a / b}[message_op]
