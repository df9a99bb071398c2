"""Python face of the native process group (csrc/runtime/process_group.cc; reference binding:
paddle/fluid/pybind/distributed_py.cc - ProcessGroupNCCL methods returning Task objects).

``NativeProcessGroup.for_group(group)`` lazily creates the NCCL communicator of a paddle group (collective over the
group; the unique id travels through this library's TCP store, whose address rank 0 announces through the
torch.distributed store that ``init_parallel_env`` already set up).  The methods mirror the reference's:
``all_reduce(tensor, op, sync_op, use_calc_stream)`` ... and return a ``NativeTask`` whose ``wait()`` orders the current
stream after the collective on the device.  ``FLAGS_use_native_process_group`` routes ``paddle.distributed``
collectives on CUDA tensors through it.
"""
from __future__ import annotations

import ctypes
import os
import socket

import torch

from .. import lib
from ..framework.flags import flag

_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2, torch.float64: 3, torch.int32: 4, torch.int64: 5,
       torch.uint8: 6, torch.int8: 7, torch.bool: 6}
_OP = {"SUM": 0, "MAX": 1, "MIN": 2, "PROD": 3, "AVG": 4}
_store = [None]
_groups = {}


def available():
    return (lib.runtime.available() and hasattr(lib.runtime.require(), "pdb200_pg_create") and
            lib.runtime.pdb200_pg_available() == 1 and torch.cuda.is_available())


def enabled_for(tensor):
    return bool(flag("use_native_process_group")) and isinstance(tensor, torch.Tensor) and tensor.is_cuda and \
        tensor.dtype in _DT and available()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _global_store():
    """One native TCP store per job: global rank 0 hosts it and announces host:port through torch's store."""
    if _store[0] is not None:
        return _store[0]
    import torch.distributed as tdist
    from .store import TCPStore
    from . import collective as C
    world, rank = C.get_world_size(), C.get_rank()
    ts = tdist.distributed_c10d._get_default_store()
    host = os.environ.get("MASTER_ADDR", "127.0.0.1")
    if rank == 0:
        port = _free_port()
        st = TCPStore(host, port, is_master=True, world_size=world)
        ts.set("pdb200_native_store", f"{host}:{port}")
    else:
        addr = ts.get("pdb200_native_store").decode()
        h, p = addr.rsplit(":", 1)
        st = TCPStore(h, int(p), is_master=False, world_size=world)
    _store[0] = st
    return st


class NativeTask:
    """Handle of one native collective (reference: ProcessGroup::Task)."""

    def __init__(self, pg, handle, keep=()):
        self._pg, self._h, self._keep = pg, handle, keep      # keep: tensors alive until the collective is done

    def wait(self, timeout=None):
        if self._h:
            lib.runtime.pdb200_pg_task_wait(self._h, torch.cuda.current_stream().cuda_stream)
            self._release()
        return True

    def is_completed(self):
        return (not self._h) or lib.runtime.pdb200_pg_task_is_completed(self._h) == 1

    def synchronize(self):
        if self._h:
            lib.runtime.pdb200_pg_task_synchronize(self._h)
            self._release()

    def _release(self):
        if self._h:
            # the current stream now depends on the collective: tensors may be freed in stream order
            for t in self._keep:
                t.record_stream(torch.cuda.current_stream())
            lib.runtime.pdb200_pg_task_free(self._pg._h, self._h)
            self._h, self._keep = None, ()

    def __del__(self):
        try:
            if self._h:
                lib.runtime.pdb200_pg_task_synchronize(self._h)
                lib.runtime.pdb200_pg_task_free(self._pg._h, self._h)
                self._h = None
        except Exception:  # noqa: BLE001
            pass


class NativeProcessGroup:
    def __init__(self, group):
        rt = lib.runtime.require()
        self.group = group
        self.rank, self.nranks = group.rank, group.nranks
        store = _global_store()
        self._h = rt.pdb200_pg_create(store._h, b"pdb200", int(group.id), int(group.rank), int(group.nranks), 120000)
        if not self._h:
            raise RuntimeError(f"native process group: {rt.pdb200_rt_last_error().decode()}")
        self.comm_stream = torch.cuda.ExternalStream(rt.pdb200_pg_comm_stream(self._h))

    @classmethod
    def for_group(cls, group):
        pg = _groups.get(group.id)
        if pg is None:
            pg = _groups[group.id] = cls(group)
        return pg

    # ---- helpers
    def _launch(self, fn, args, tensors, sync_op, use_calc_stream):
        cur = torch.cuda.current_stream().cuda_stream
        for t in tensors:
            if not use_calc_stream:
                t.record_stream(self.comm_stream)          # allocator safety: the comm stream uses this memory
        h = fn(self._h, *args, cur, int(bool(use_calc_stream)), int(bool(sync_op)))
        if not h:
            raise RuntimeError(f"native process group: {lib.runtime.pdb200_rt_last_error().decode()}")
        task = NativeTask(self, h, tuple(tensors))
        if sync_op:
            task.wait()
        return task

    @staticmethod
    def _op(op):
        if isinstance(op, int):                 # paddle ReduceOp constants: SUM 0, MAX 1, MIN 2, PROD 3, AVG 4
            return int(op)
        name = getattr(op, "name", None) or str(op).split(".")[-1]
        return _OP[name.upper()]

    def _local(self, global_rank):
        return self.group.ranks.index(global_rank)

    # ---- collectives (tensors must be contiguous CUDA tensors)
    def all_reduce(self, tensor, op, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_all_reduce, (tensor.data_ptr(), tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype],
                                                      self._op(op)), [tensor], sync_op, use_calc_stream)

    def broadcast(self, tensor, src, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_broadcast, (tensor.data_ptr(), tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype],
                                                     self._local(src)), [tensor], sync_op, use_calc_stream)

    def reduce(self, tensor, dst, op, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_reduce, (tensor.data_ptr(), tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype],
                                                  self._op(op), self._local(dst)), [tensor], sync_op, use_calc_stream)

    def all_gather(self, out, tensor, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_all_gather, (tensor.data_ptr(), out.data_ptr(), tensor.numel(), _DT[tensor.dtype]),
                            [tensor, out], sync_op, use_calc_stream)

    def reduce_scatter(self, out, tensor, op, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_reduce_scatter, (tensor.data_ptr(), out.data_ptr(), out.numel(), _DT[tensor.dtype],
                                                          self._op(op)), [tensor, out], sync_op, use_calc_stream)

    def all_to_all(self, out, tensor, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_all_to_all, (tensor.data_ptr(), out.data_ptr(), tensor.numel() // self.nranks,
                                                      _DT[tensor.dtype]), [tensor, out], sync_op, use_calc_stream)

    def send(self, tensor, dst, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_send, (tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype], self._local(dst)),
                            [tensor], sync_op, use_calc_stream)

    def recv(self, tensor, src, sync_op=True, use_calc_stream=False):
        rt = lib.runtime
        return self._launch(rt.pdb200_pg_recv, (tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype], self._local(src)),
                            [tensor], sync_op, use_calc_stream)

    def abort(self):
        lib.runtime.pdb200_pg_abort(self._h)

    def close(self):
        if self._h:
            lib.runtime.pdb200_pg_destroy(self._h)
            self._h = None
            _groups.pop(self.group.id, None)


def shutdown():
    for pg in list(_groups.values()):
        pg.close()
    _store[0] = None
