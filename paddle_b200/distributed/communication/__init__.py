"""Collective communication API (reference: python/paddle/distributed/communication/*.py and stream/*.py; native
side paddle/fluid/distributed/collective/process_group_nccl.cc:203-682).

Every function has the reference signature ``(..., group=None, sync_op=True)`` and returns a ``Task``.  ``src`` /
``dst`` are GLOBAL ranks.  NCCL runs on torch's per-group communication stream and ``task.wait()`` orders the
current stream after it; gloo (CPU) blocks.  Gaps of the gloo backend (reduce_scatter, all_to_all, AVG, 16-bit
floats) are filled by composition so that every hybrid-parallel code path can be exercised without a GPU.
"""
from __future__ import annotations

import pickle

import numpy as np
import torch
import torch.distributed as tdist

from . import check as _chk

from .. import collective as C
from ..collective import ReduceOp, Task, _finish, _in_group, _resolve, _state
from .. import watchdog as _wd


def _trivial(g):
    return g.nranks == 1 or not tdist.is_initialized()


def _native(g, *tensors):
    """The native process group of ``g`` when FLAGS_use_native_process_group routes these tensors through it."""
    from .. import native_pg
    if _state.backend != "nccl" or not all(native_pg.enabled_for(t) and t.is_contiguous() for t in tensors):
        return None
    return native_pg.NativeProcessGroup.for_group(g)


class _NativeWork:
    """Adapts a NativeTask to the torch Work interface ``Task`` expects."""

    def __init__(self, task):
        self._t = task

    def wait(self):
        return self._t.wait()

    def is_completed(self):
        return self._t.is_completed()


def _is_gloo():
    return _state.backend == "gloo"


def _check_tensor(t, name="tensor"):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a Tensor, got {type(t)}")


class _F32Proxy:
    """gloo has no 16-bit float reductions: stage through fp32 and copy back."""

    def __init__(self, t):
        self.orig = t
        self.need = _is_gloo() and t.dtype in (torch.bfloat16, torch.float16)
        self.buf = t.float() if self.need else t

    def back(self):
        if self.need:
            with torch.no_grad():
                self.orig.copy_(self.buf.to(self.orig.dtype))


# ------------------------------------------------------------------------------------------- all_reduce
def all_reduce(tensor, op=ReduceOp.SUM, group=None, sync_op=True, use_calc_stream=False):
    _check_tensor(tensor)
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    _chk.same_shape(tensor, tensor, C._state.backend, "all_reduce")
    _chk.dynamic_check(tensor, 0, g, "all_reduce")
    tag = _wd.task_started(g, "all_reduce", tensor.numel(), sync_op)
    if _trivial(g):
        return _finish(None, sync_op, tag=tag)
    npg = _native(g, tensor)
    if npg is not None:
        with torch.no_grad():
            return _finish(_NativeWork(npg.all_reduce(tensor, op, sync_op=False, use_calc_stream=use_calc_stream)), sync_op, tag=tag)
    px = _F32Proxy(tensor)
    top = op
    post_scale = None
    if op == ReduceOp.AVG and _is_gloo():
        top, post_scale = ReduceOp.SUM, 1.0 / g.nranks
    with torch.no_grad():
        work = tdist.all_reduce(px.buf, op=C._TORCH_OP[top], group=g.process_group, async_op=True)

    def post():
        if post_scale is not None:
            with torch.no_grad():
                px.buf.mul_(post_scale)
        px.back()
    return _finish(work, sync_op, post, tag)


# ------------------------------------------------------------------------------------------- broadcast / reduce
def broadcast(tensor, src, group=None, sync_op=True, use_calc_stream=False):
    _check_tensor(tensor)
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    tag = _wd.task_started(g, "broadcast", tensor.numel(), sync_op)
    if _trivial(g):
        return _finish(None, sync_op, tag=tag)
    if src not in g.ranks:
        raise ValueError(f"src rank {src} is not in group {g.ranks}")
    npg = _native(g, tensor)
    if npg is not None:
        with torch.no_grad():
            return _finish(_NativeWork(npg.broadcast(tensor, src, sync_op=False, use_calc_stream=use_calc_stream)), sync_op, tag=tag)
    with torch.no_grad():
        work = tdist.broadcast(tensor, src=src, group=g.process_group, async_op=True)
    return _finish(work, sync_op, tag=tag)


def reduce(tensor, dst, op=ReduceOp.SUM, group=None, sync_op=True, use_calc_stream=False):
    _check_tensor(tensor)
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    tag = _wd.task_started(g, "reduce", tensor.numel(), sync_op)
    if _trivial(g):
        return _finish(None, sync_op, tag=tag)
    px = _F32Proxy(tensor)
    top, post_scale = op, None
    if op == ReduceOp.AVG and _is_gloo():
        top, post_scale = ReduceOp.SUM, 1.0 / g.nranks
    with torch.no_grad():
        work = tdist.reduce(px.buf, dst=dst, op=C._TORCH_OP[top], group=g.process_group, async_op=True)

    def post():
        if post_scale is not None and _state.global_rank == dst:
            with torch.no_grad():
                px.buf.mul_(post_scale)
        if _state.global_rank == dst:
            px.back()
    return _finish(work, sync_op, post, tag)


# ------------------------------------------------------------------------------------------- all_gather
def all_gather(tensor_or_tensor_list, tensor, group=None, sync_op=True, use_calc_stream=False):
    """``all_gather(tensor_list, tensor)``: list is filled (created when empty); the stream variant also takes a
    single pre-allocated output tensor of nranks x the input size."""
    _check_tensor(tensor)
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    n = g.nranks
    out = tensor_or_tensor_list
    tag = _wd.task_started(g, "all_gather", tensor.numel(), sync_op)
    if isinstance(out, torch.Tensor):
        _chk.gather_like(out, tensor, n, C._state.backend, "all_gather")
        if _trivial(g):
            with torch.no_grad():
                out.reshape(-1).copy_(tensor.reshape(-1))
            return _finish(None, sync_op, tag=tag)
        npg = _native(g, out, tensor)
        if npg is not None:
            with torch.no_grad():
                return _finish(_NativeWork(npg.all_gather(out, tensor, sync_op=False, use_calc_stream=use_calc_stream)), sync_op, tag=tag)
        with torch.no_grad():
            work = tdist.all_gather_into_tensor(out.view(-1) if out.is_contiguous() else out, tensor.contiguous().view(-1),
                                                group=g.process_group, async_op=True) \
                if not _is_gloo() else None
            if work is None:
                parts = list(out.view(-1).chunk(n))
                work = tdist.all_gather(parts, tensor.contiguous().view(-1), group=g.process_group, async_op=True)
        return _finish(work, sync_op, tag=tag)
    if not isinstance(out, list):
        raise TypeError("tensor_list must be a list (or a Tensor for the stream variant)")
    if len(out) == 0:
        out.extend(torch.empty_like(tensor) for _ in range(n))
    elif len(out) != n:
        raise ValueError(f"tensor_list must hold {n} tensors")
    if _trivial(g):
        with torch.no_grad():
            out[0].copy_(tensor)
        return _finish(None, sync_op, tag=tag)
    with torch.no_grad():
        work = tdist.all_gather(out, tensor.contiguous(), group=g.process_group, async_op=True)
    return _finish(work, sync_op, tag=tag)


def all_gather_object(object_list, obj, group=None):
    g = _resolve(group)
    if not _in_group(g):
        return
    if _trivial(g):
        object_list.clear()
        object_list.append(obj)
        return
    tmp = [None] * g.nranks
    tdist.all_gather_object(tmp, obj, group=g.process_group)
    object_list.clear()
    object_list.extend(tmp)


# ------------------------------------------------------------------------------------------- reduce_scatter
def reduce_scatter(tensor, tensor_or_tensor_list, op=ReduceOp.SUM, group=None, sync_op=True, use_calc_stream=False):
    _check_tensor(tensor)
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    n = g.nranks
    src = tensor_or_tensor_list
    tag = _wd.task_started(g, "reduce_scatter", tensor.numel() * n, sync_op)
    flat_in = torch.cat([t.reshape(-1) for t in src]) if isinstance(src, (list, tuple)) else src.contiguous().view(-1)
    _chk.scatter_like(tensor, flat_in, n, C._state.backend, "reduce_scatter")
    if _trivial(g):
        with torch.no_grad():
            tensor.copy_(flat_in.view_as(tensor))
        return _finish(None, sync_op, tag=tag)
    if _is_gloo():
        buf = flat_in.float() if flat_in.dtype in (torch.bfloat16, torch.float16) else flat_in.clone()
        top = ReduceOp.SUM if op == ReduceOp.AVG else op
        work = tdist.all_reduce(buf, op=C._TORCH_OP[top], group=g.process_group, async_op=True)

        def post():
            with torch.no_grad():
                part = buf.view(n, -1)[g.rank]
                if op == ReduceOp.AVG:
                    part = part / n
                tensor.copy_(part.view_as(tensor).to(tensor.dtype))
        return _finish(work, sync_op, post, tag)
    npg = _native(g, tensor, flat_in)
    if npg is not None:
        with torch.no_grad():
            return _finish(_NativeWork(npg.reduce_scatter(tensor, flat_in, op, sync_op=False, use_calc_stream=use_calc_stream)),
                           sync_op, tag=tag)
    with torch.no_grad():
        out = tensor if tensor.is_contiguous() else tensor.contiguous()
        work = tdist.reduce_scatter_tensor(out.view(-1), flat_in, op=C._TORCH_OP[op], group=g.process_group, async_op=True)

    def post2():
        if out is not tensor:
            with torch.no_grad():
                tensor.copy_(out)
    return _finish(work, sync_op, post2, tag)


# ------------------------------------------------------------------------------------------- all_to_all
def alltoall(out_tensor_list, in_tensor_list, group=None, sync_op=True, use_calc_stream=False):
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    n = g.nranks
    if len(in_tensor_list) != n:
        raise ValueError(f"in_tensor_list must hold {n} tensors")
    if len(out_tensor_list) == 0:
        out_tensor_list.extend(torch.empty_like(t) for t in in_tensor_list)
    tag = _wd.task_started(g, "alltoall", sum(t.numel() for t in in_tensor_list), sync_op)
    if _trivial(g):
        with torch.no_grad():
            out_tensor_list[0].copy_(in_tensor_list[0])
        return _finish(None, sync_op, tag=tag)
    if _is_gloo():
        return _finish(_P2PAllToAll(out_tensor_list, in_tensor_list, g), sync_op, tag=tag)
    with torch.no_grad():
        work = tdist.all_to_all(out_tensor_list, [t.contiguous() for t in in_tensor_list], group=g.process_group, async_op=True)
    return _finish(work, sync_op, tag=tag)


all_to_all = alltoall


class _P2PAllToAll:
    """gloo fallback: pairwise isend/irecv."""

    def __init__(self, outs, ins, g):
        self.reqs = []
        me = g.rank
        with torch.no_grad():
            outs[me].copy_(ins[me])
        for peer in range(g.nranks):
            if peer == me:
                continue
            self.reqs.append(tdist.isend(ins[peer].contiguous(), g.ranks[peer], group=g.process_group))
            self.reqs.append(tdist.irecv(outs[peer], g.ranks[peer], group=g.process_group))

    def wait(self):
        for r in self.reqs:
            r.wait()

    def is_completed(self):
        return all(r.is_completed() for r in self.reqs)


def alltoall_single(out_tensor, in_tensor, in_split_sizes=None, out_split_sizes=None, group=None, sync_op=True,
                    use_calc_stream=False):
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    tag = _wd.task_started(g, "alltoall_single", in_tensor.numel(), sync_op)
    if _trivial(g):
        with torch.no_grad():
            out_tensor.copy_(in_tensor)
        return _finish(None, sync_op, tag=tag)
    if _is_gloo():
        n = g.nranks
        ins = list(in_tensor.split(in_split_sizes, 0)) if in_split_sizes else list(in_tensor.chunk(n, 0))
        outs = list(out_tensor.split(out_split_sizes, 0)) if out_split_sizes else list(out_tensor.chunk(n, 0))
        return _finish(_P2PAllToAll(outs, ins, g), sync_op, tag=tag)
    with torch.no_grad():
        work = tdist.all_to_all_single(out_tensor, in_tensor.contiguous(), out_split_sizes, in_split_sizes,
                                       group=g.process_group, async_op=True)
    return _finish(work, sync_op, tag=tag)


# ------------------------------------------------------------------------------------------- scatter / gather
def scatter(tensor, tensor_or_tensor_list=None, src=0, group=None, sync_op=True, use_calc_stream=False):
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    tag = _wd.task_started(g, "scatter", tensor.numel(), sync_op)
    lst = tensor_or_tensor_list
    if isinstance(lst, torch.Tensor):
        lst = list(lst.chunk(g.nranks, 0))
    if _trivial(g):
        with torch.no_grad():
            tensor.copy_(lst[0].view_as(tensor))
        return _finish(None, sync_op, tag=tag)
    is_src = _state.global_rank == src
    if _is_gloo() or True:
        # uniform implementation on both backends: root sends slices (NCCL has no native scatter either,
        # process_group_nccl.cc:468-601 does grouped send/recv)
        if is_src:
            reqs = []
            for i, r in enumerate(g.ranks):
                if r == src:
                    with torch.no_grad():
                        tensor.copy_(lst[i].view_as(tensor))
                else:
                    reqs.append(tdist.isend(lst[i].contiguous(), r, group=g.process_group))
            work = _Reqs(reqs)
        else:
            work = _Reqs([tdist.irecv(tensor, src, group=g.process_group)])
    return _finish(work, sync_op, tag=tag)


class _Reqs:
    def __init__(self, reqs):
        self.reqs = reqs

    def wait(self):
        for r in self.reqs:
            r.wait()

    def is_completed(self):
        return all(r.is_completed() for r in self.reqs)


def gather(tensor, gather_list=None, dst=0, group=None, sync_op=True, use_calc_stream=False):
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    tag = _wd.task_started(g, "gather", tensor.numel(), sync_op)
    is_dst = _state.global_rank == dst
    if is_dst:
        if gather_list is None:
            raise ValueError("gather_list must be given on the dst rank")
        if len(gather_list) == 0:
            gather_list.extend(torch.empty_like(tensor) for _ in range(g.nranks))
    if _trivial(g):
        with torch.no_grad():
            gather_list[0].copy_(tensor)
        return _finish(None, sync_op, tag=tag)
    if is_dst:
        reqs = []
        for i, r in enumerate(g.ranks):
            if r == dst:
                with torch.no_grad():
                    gather_list[i].copy_(tensor)
            else:
                reqs.append(tdist.irecv(gather_list[i], r, group=g.process_group))
        work = _Reqs(reqs)
    else:
        work = _Reqs([tdist.isend(tensor.contiguous(), dst, group=g.process_group)])
    return _finish(work, sync_op, tag=tag)


# ------------------------------------------------------------------------------------------- p2p
def send(tensor, dst=0, group=None, sync_op=True, use_calc_stream=False):
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    tag = _wd.task_started(g, "send", tensor.numel(), sync_op)
    if _trivial(g):
        return _finish(None, sync_op, tag=tag)
    work = tdist.isend(tensor.contiguous(), dst, group=g.process_group)
    return _finish(work, sync_op, tag=tag)


def recv(tensor, src=0, group=None, sync_op=True, use_calc_stream=False):
    g = _resolve(group)
    if not _in_group(g):
        return Task()
    tag = _wd.task_started(g, "recv", tensor.numel(), sync_op)
    if _trivial(g):
        return _finish(None, sync_op, tag=tag)
    work = tdist.irecv(tensor, src, group=g.process_group)
    return _finish(work, sync_op, tag=tag)


def isend(tensor, dst, group=None):
    return send(tensor, dst, group, sync_op=False)


def irecv(tensor, src=None, group=None):
    return recv(tensor, src, group, sync_op=False)


class P2POp:
    def __init__(self, op, tensor, peer, group=None):
        if op not in (isend, irecv):
            raise RuntimeError("Invalid ``op`` function. Expected ``op`` to be of type ``paddle.distributed.isend`` or "
                               "``paddle.distributed.irecv``.")
        self.op, self.tensor, self.peer, self.group = op, tensor, peer, group


def batch_isend_irecv(p2p_op_list):
    """Reference: communication/batch_isend_irecv.py:23-181 (ncclGroupStart/End around the list)."""
    if not p2p_op_list:
        return []
    g = _resolve(p2p_op_list[0].group)
    if _trivial(g):
        return [Task()]
    tag = _wd.task_started(g, "batch_isend_irecv", sum(o.tensor.numel() for o in p2p_op_list), False)
    ops = [tdist.P2POp(tdist.isend if o.op is isend else tdist.irecv, o.tensor, o.peer, g.process_group)
           for o in p2p_op_list]
    works = tdist.batch_isend_irecv(ops)
    return [Task(w, tag=tag if i == 0 else "") for i, w in enumerate(works)]


# ------------------------------------------------------------------------------------------- object collectives
def _obj_to_tensor(obj, device):
    data = np.frombuffer(pickle.dumps(obj), dtype=np.uint8).copy()
    return torch.from_numpy(data).to(device), data.size


def _comm_device():
    return torch.device("cuda", torch.cuda.current_device()) if _state.backend == "nccl" else torch.device("cpu")


def broadcast_object_list(object_list, src, group=None):
    g = _resolve(group)
    if not _in_group(g) or _trivial(g):
        return
    tdist.broadcast_object_list(object_list, src=src, group=g.process_group, device=_comm_device())


def scatter_object_list(out_object_list, in_object_list=None, src=0, group=None):
    g = _resolve(group)
    if not _in_group(g):
        return
    if _trivial(g):
        out_object_list.clear()
        out_object_list.append(in_object_list[0])
        return
    tmp = [None]
    tdist.scatter_object_list(tmp, in_object_list if _state.global_rank == src else None, src=src, group=g.process_group)
    out_object_list.clear()
    out_object_list.append(tmp[0])


def gather_object(obj, object_gather_list=None, dst=0, group=None):
    g = _resolve(group)
    if not _in_group(g):
        return
    if _trivial(g):
        object_gather_list.clear()
        object_gather_list.append(obj)
        return
    tmp = [None] * g.nranks if _state.global_rank == dst else None
    tdist.gather_object(obj, tmp, dst=dst, group=g.process_group)
    if tmp is not None:
        object_gather_list.clear()
        object_gather_list.extend(tmp)


from . import stream  # noqa: E402,F401
