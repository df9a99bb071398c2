"""Process-group management (reference: python/paddle/distributed/collective.py:132-265,
communication/group.py:22-374).

One process per GPU.  ``torch.distributed`` (NCCL over NVLink/NVSwitch on GPUs, gloo on CPU) is the plumbing under
Paddle's group semantics: ``new_group`` must be called by every rank in the same order, ids come from a local
counter (``max(nrings, 9) + n``), non-members get ``rank == -1`` and no process group.
"""
from __future__ import annotations

import datetime
import os

import torch
import torch.distributed as tdist


class ReduceOp:
    SUM = 0
    MAX = 1
    MIN = 2
    PROD = 3
    AVG = 4


_TORCH_OP = {ReduceOp.SUM: tdist.ReduceOp.SUM, ReduceOp.MAX: tdist.ReduceOp.MAX, ReduceOp.MIN: tdist.ReduceOp.MIN,
             ReduceOp.PROD: tdist.ReduceOp.PRODUCT, ReduceOp.AVG: tdist.ReduceOp.AVG}


class Group:
    """A set of global ranks with its communicator (reference: communication/group.py:22-108)."""

    def __init__(self, rank_in_group, id, ranks, pg=None, name=None):
        self._rank_in_group = rank_in_group
        self._world_size = len(ranks) if rank_in_group >= 0 else -1
        self._id = id
        self._ranks = list(ranks)
        self._pg = pg
        self._name = name

    @property
    def rank(self):
        return self._rank_in_group

    @property
    def ranks(self):
        return self._ranks

    @property
    def nranks(self):
        return len(self._ranks)

    @property
    def world_size(self):
        return len(self._ranks)

    @property
    def name(self):
        return self._name

    @property
    def process_group(self):
        return self._pg

    @property
    def id(self):
        return self._id

    @property
    def backend(self):
        return _state.backend

    def is_member(self):
        return self._rank_in_group >= 0 and self.nranks > 1 or (self._rank_in_group >= 0)

    def get_group_rank(self, rank):
        return self._ranks.index(rank) if rank in self._ranks else -1

    def get_global_rank(self, group_rank):
        return self._ranks[group_rank]

    def __repr__(self):
        return f"rank: {self.rank}, nranks: {self.nranks}, id: {self.id}, ranks: {', '.join(map(str, self.ranks))}; name: {self.name}"


class _State:
    def __init__(self):
        self.backend = None
        self.default_group: Group | None = None
        self.groups: dict[int, Group] = {}
        self.ring_counter = 0
        self.global_rank = 0
        self.world_size = 1
        self.store = None          # torch store used for rendezvous
        self.native_store = None   # libpdb200_runtime TCP store (handle exchange, watchdog)


_state = _State()


def _new_ring_id():
    nrings = int(os.environ.get("FLAGS_nccl_nrings", "1"))
    _state.ring_counter += 1
    return max(nrings, 9) + _state.ring_counter


def is_initialized():
    return _state.default_group is not None


def is_available():
    return tdist.is_available()


def get_backend(group=None):
    return (_state.backend or "").upper() if _state.backend else None


def _get_default_group():
    if _state.default_group is None:
        raise RuntimeError("Call paddle.distributed.init_parallel_env first to initialize the distributed environment.")
    return _state.default_group


def _get_global_group():
    return _get_default_group()


def get_group(id=0):
    if id in _state.groups:
        return _state.groups[id]
    return None


def _resolve(group):
    """-> Group (default when None)."""
    return group if group is not None else _get_default_group()


def _in_group(group) -> bool:
    return group is None or group.rank >= 0


def new_group(ranks=None, backend=None, timeout=datetime.timedelta(seconds=1800), nccl_comm_init_option=0):
    default = _get_default_group()
    gid = _new_ring_id()
    if ranks is None:
        ranks = list(range(_state.world_size))
    ranks = sorted(int(r) for r in ranks)
    if backend is not None and backend.lower() not in ("nccl", "gloo"):
        raise ValueError(f"backend {backend!r} is not supported by this build (nccl, gloo)")
    me = _state.global_rank
    if _state.world_size == 1 or not tdist.is_initialized():
        pg = None
    else:
        # every rank must call in the same order, members or not (Appendix B.1)
        pg = tdist.new_group(ranks=ranks, backend=(backend or _state.backend).lower(), timeout=timeout)
    rank_in_group = ranks.index(me) if me in ranks else -1
    g = Group(rank_in_group, gid, ranks, pg if rank_in_group >= 0 else None, name=f"_default_pg{gid}")
    _state.groups[gid] = g
    from . import watchdog
    watchdog.note_group(g)
    return g


def destroy_process_group(group=None):
    if group is None or group is _state.default_group:
        from . import watchdog
        watchdog.shutdown()
        if tdist.is_initialized():
            try:
                tdist.destroy_process_group()
            except Exception:  # noqa: BLE001
                pass
        _state.groups.clear()
        _state.default_group = None
        _state.backend = None
        _state.ring_counter = 0
        _state.world_size = 1
        _state.global_rank = 0
    else:
        if group.process_group is not None:
            tdist.destroy_process_group(group.process_group)
        _state.groups.pop(group.id, None)


def get_rank(group=None):
    if group is not None:
        return group.rank
    if _state.default_group is not None:
        return _state.global_rank
    from .parallel import ParallelEnv
    return ParallelEnv().rank


def get_world_size(group=None):
    if group is not None:
        return group.nranks if group.rank >= 0 else -1
    if _state.default_group is not None:
        return _state.world_size
    from .parallel import ParallelEnv
    return ParallelEnv().world_size


class Task:
    """Handle of an asynchronous collective (reference: ProcessGroup::Task, process_group.h:47-120).
    ``wait()`` makes the *current stream* wait (NCCL) or blocks the host (gloo); it never spins the host on GPU."""

    def __init__(self, work=None, post=None, tag=""):
        self._work = work
        self._post = post
        self._done = work is None
        self._tag = tag

    def wait(self, timeout=None):
        if self._work is not None:
            self._work.wait()
            self._work = None
        if self._post is not None:
            self._post()
            self._post = None
        self._done = True
        from . import watchdog
        watchdog.task_finished(self._tag)
        return True

    def is_completed(self):
        if self._work is None:
            return True
        return self._work.is_completed()

    def is_sync(self):
        return self._done

    def synchronize(self):
        self.wait()
        if torch.cuda.is_available():
            torch.cuda.current_stream().synchronize()


def _finish(work, sync_op, post=None, tag=""):
    if tag:
        from . import watchdog
        watchdog.task_enqueued(tag, work)
    t = Task(work, post, tag)
    if sync_op:
        t.wait()
        from ..framework.flags import flag
        if torch.cuda.is_available() and (flag("benchmark_nccl") or flag("nccl_blocking_wait")):
            torch.cuda.synchronize()
    return t


def barrier(group=None):
    g = _resolve(group)
    if not _in_group(g) or g.nranks == 1 or not tdist.is_initialized():
        return
    if _state.backend == "nccl":
        tdist.barrier(group=g.process_group, device_ids=[torch.cuda.current_device()])
    else:
        tdist.barrier(group=g.process_group)


def wait(tensor, group=None, use_calc_stream=True):
    """Reference: communication/group.py wait().  Collectives issued with sync_op already order the calc stream;
    for the async case make the current stream catch up with everything enqueued so far."""
    if tensor.is_cuda and not use_calc_stream:
        torch.cuda.current_stream().synchronize()


def _get_group_map_by_name():
    return {g.name: g for g in _state.groups.values()}
