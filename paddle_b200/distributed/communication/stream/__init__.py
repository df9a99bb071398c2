"""paddle.distributed.stream.* (reference: python/paddle/distributed/communication/stream/*.py).
Same collectives with the extra ``use_calc_stream`` switch, legal only together with ``sync_op=True``
(stream/all_reduce.py:121-124)."""
from __future__ import annotations

from ... import communication as _c
from ...collective import ReduceOp


def _check(sync_op, use_calc_stream):
    if not sync_op and use_calc_stream:
        raise RuntimeError("use_calc_stream can only be true in sync op behavior.")


def all_reduce(tensor, op=ReduceOp.SUM, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.all_reduce(tensor, op, group, sync_op, use_calc_stream)


def all_gather(tensor_or_tensor_list, tensor, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.all_gather(tensor_or_tensor_list, tensor, group, sync_op, use_calc_stream)


def reduce_scatter(tensor, tensor_or_tensor_list, op=ReduceOp.SUM, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.reduce_scatter(tensor, tensor_or_tensor_list, op, group, sync_op, use_calc_stream)


def alltoall(out_tensor_or_tensor_list, in_tensor_or_tensor_list, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    import torch
    if isinstance(out_tensor_or_tensor_list, torch.Tensor):
        return _c.alltoall_single(out_tensor_or_tensor_list, in_tensor_or_tensor_list, None, None, group, sync_op)
    return _c.alltoall(out_tensor_or_tensor_list, in_tensor_or_tensor_list, group, sync_op)


def alltoall_single(out_tensor, in_tensor, out_split_sizes=None, in_split_sizes=None, group=None, sync_op=True,
                    use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.alltoall_single(out_tensor, in_tensor, in_split_sizes, out_split_sizes, group, sync_op)


def broadcast(tensor, src, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.broadcast(tensor, src, group, sync_op)


def reduce(tensor, dst=0, op=ReduceOp.SUM, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.reduce(tensor, dst, op, group, sync_op)


def scatter(tensor, tensor_or_tensor_list=None, src=0, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.scatter(tensor, tensor_or_tensor_list, src, group, sync_op)


def gather(tensor, gather_list=None, dst=0, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.gather(tensor, gather_list, dst, group, sync_op)


def send(tensor, dst=0, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.send(tensor, dst, group, sync_op)


def recv(tensor, src=0, group=None, sync_op=True, use_calc_stream=False):
    _check(sync_op, use_calc_stream)
    return _c.recv(tensor, src, group, sync_op)
