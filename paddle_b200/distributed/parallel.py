"""init_parallel_env / ParallelEnv / DataParallel (reference: python/paddle/distributed/parallel.py:147-1289).

Process model: one process per GPU, rank/world/endpoints from the environment (PADDLE_* set by
``paddle.distributed.launch``/``spawn``; torchrun's RANK/WORLD_SIZE/MASTER_* are accepted too).  Rendezvous
uses a TCP store on the master endpoint; the default group has id 0 and is named ``_default_pg``.
"""
from __future__ import annotations

import contextlib
import datetime
import os

import torch
import torch.distributed as tdist

from ..framework import flags as _flags
from ..framework.place import CPUPlace, CUDAPlace, set_device
from . import collective as C
from .collective import Group, _state
from .reducer import EagerReducer, assign_group_by_size


class ParallelEnv:
    """Reference: parallel.py:644-878."""

    def __init__(self):
        e = os.environ
        self._rank = int(e.get("PADDLE_TRAINER_ID", e.get("RANK", "0")))
        self._world_size = int(e.get("PADDLE_TRAINERS_NUM", e.get("WORLD_SIZE", "1")))
        sel = e.get("FLAGS_selected_gpus", e.get("LOCAL_RANK", "0"))
        self._device_id = int(str(sel).split(",")[0] or 0)
        self._local_rank = int(e.get("PADDLE_LOCAL_RANK", e.get("PADDLE_RANK_IN_NODE", e.get("LOCAL_RANK", "0"))))
        eps = e.get("PADDLE_TRAINER_ENDPOINTS", "")
        self._trainer_endpoints = eps.split(",") if eps else []
        self._current_endpoint = e.get("PADDLE_CURRENT_ENDPOINT", "")
        self._nrings = int(e.get("FLAGS_nccl_nrings", "1"))
        self._pg_timeout = int(e.get("PADDLE_PG_TIMEOUT", "1800000"))

    @property
    def rank(self):
        return self._rank

    local_rank = rank

    @property
    def world_size(self):
        return self._world_size

    nranks = world_size

    @property
    def device_id(self):
        return self._device_id

    dev_id = device_id

    @property
    def device_type(self):
        return "gpu" if torch.cuda.is_available() else "cpu"

    @property
    def current_endpoint(self):
        return self._current_endpoint

    @property
    def trainer_endpoints(self):
        return self._trainer_endpoints

    @property
    def nrings(self):
        return self._nrings

    @property
    def pg_timeout(self):
        return self._pg_timeout


def _master_endpoint(env: ParallelEnv):
    e = os.environ
    if e.get("PADDLE_MASTER"):
        host, port = e["PADDLE_MASTER"].rsplit(":", 1)
        return host, int(port)
    if e.get("MASTER_ADDR") and e.get("MASTER_PORT"):
        return e["MASTER_ADDR"], int(e["MASTER_PORT"])
    if env.trainer_endpoints:
        host, port = env.trainer_endpoints[0].rsplit(":", 1)
        return host, int(port)
    return "127.0.0.1", 6170


def _print_modified_flags():
    mod = _flags.modified_flags()
    if mod and os.environ.get("PADDLE_B200_VERBOSE"):
        print("modified FLAGS:", mod, flush=True)


def init_parallel_env():
    """Reference: parallel.py:945-1131.  Returns the default Group (id 0)."""
    if C.is_initialized():
        return _state.default_group
    env = ParallelEnv()
    backend = os.environ.get("PADDLE_DISTRI_BACKEND", "auto").lower()
    if backend == "auto":
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend not in ("nccl", "gloo"):
        raise ValueError(f"paddle.distributed initialize error, backend argument can only be one of 'nccl', 'gloo' "
                         f"in this build, but got {backend}")
    if backend == "nccl":
        set_device(CUDAPlace(env.device_id % max(torch.cuda.device_count(), 1)))
    else:
        set_device(CPUPlace())
    _state.backend = backend
    _state.global_rank, _state.world_size = env.rank, env.world_size
    if env.world_size > 1:
        host, port = _master_endpoint(env)
        timeout = datetime.timedelta(milliseconds=env.pg_timeout)
        store = tdist.TCPStore(host, port, env.world_size, is_master=(env.rank == 0), timeout=timeout,
                               wait_for_workers=False)
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
        tdist.init_process_group(backend, store=store, rank=env.rank, world_size=env.world_size, timeout=timeout, **kw)
        _state.store = store
        pg = tdist.group.WORLD
    else:
        pg = None
    g = Group(env.rank, 0, list(range(env.world_size)), pg, name="_default_pg")
    _state.default_group = g
    _state.groups[0] = g
    from . import watchdog
    watchdog.note_group(g)
    watchdog.start(env.rank)
    _print_modified_flags()
    if env.world_size > 1 and _flags.flag("eager_communication_connection"):
        t = torch.zeros(1, device="cuda" if backend == "nccl" else "cpu")
        tdist.all_reduce(t)
    return g


def get_rank(group=None):
    return C.get_rank(group)


def get_world_size(group=None):
    return C.get_world_size(group)


@torch.no_grad()
def sync_params_buffers(model, comm_group=None, src_rank=0, is_model_parallel=False, fuse_params=True):
    """Broadcast parameters and buffers from src_rank (reference: parallel.py:147-199); skips ``no_sync`` params
    and, for model-parallel groups, ``is_distributed`` ones."""
    from . import communication as comm
    ts = []
    for _, p in model.state_dict().items():
        if not isinstance(p, torch.Tensor):
            continue
        if getattr(p, "no_sync", False):
            continue
        if is_model_parallel and getattr(p, "is_distributed", False):
            continue
        ts.append(p)
    if not ts:
        return
    g = comm_group if comm_group is not None else C._get_default_group()
    if g.nranks == 1 or g.rank < 0:
        return
    if not fuse_params:
        for t in ts:
            comm.broadcast(t, src_rank, g)
        return
    by_dtype = {}
    for t in ts:
        by_dtype.setdefault((t.dtype, t.device), []).append(t)
    for (_, _), lst in by_dtype.items():
        # 128 MB fused broadcasts
        chunk, size = [], 0
        for t in lst + [None]:
            if t is not None:
                chunk.append(t)
                size += t.numel() * t.element_size()
            if chunk and (t is None or size >= 128 << 20):
                flat = torch.cat([c.detach().reshape(-1) for c in chunk])
                comm.broadcast(flat, src_rank, g)
                off = 0
                for c in chunk:
                    c.detach().copy_(flat[off:off + c.numel()].view_as(c))
                    off += c.numel()
                chunk, size = [], 0


class DataParallel(torch.nn.Module):
    """Reference: parallel.py:202-642.  Wraps a Layer; gradients are all-reduced in buckets as backward produces
    them (``EagerReducer``), overlapped with the rest of backward."""

    def __init__(self, layers, strategy=None, comm_buffer_size=25, last_comm_buffer_size=1,
                 find_unused_parameters=False, group=None):
        super().__init__()
        self._layers = layers
        self.find_unused_parameters = find_unused_parameters
        self.grad_need_sync = True
        self.group = group if group is not None else (C._get_default_group() if C.is_initialized() else None)
        self.comm_buffer_size = int(comm_buffer_size * 1024 * 1024)
        self.last_comm_buffer_size = int(last_comm_buffer_size * 1024 * 1024)
        self._reducer = None
        if self.group is not None and self.group.nranks > 1:
            sync_params_buffers(self._layers, self.group, src_rank=self.group.ranks[0], fuse_params=False)
            self.init_reducer()

    def init_reducer(self):
        params = [p for p in self._layers.parameters() if p.requires_grad]
        if not params:
            raise ValueError("This model does not have any parameters to train, and does not need to use DataParallel")
        is_sparse = [bool(getattr(p, "_is_sparse_grad", False)) for p in params]
        group_indices = assign_group_by_size(params, is_sparse, [self.last_comm_buffer_size, self.comm_buffer_size])
        allocator, fused = self._fused_comm_hooks(params)
        self._reducer = EagerReducer(params, list(reversed(group_indices)), is_sparse, self.group,
                                     [self.last_comm_buffer_size, self.comm_buffer_size], self.find_unused_parameters,
                                     bucket_allocator=allocator, fused_allreduce=fused)

    def _fused_comm_hooks(self, params):
        """K1: buckets live in NVLink symmetric memory and are reduced by ONE kernel (scale + NVLS all-reduce) on a
        side stream with a handful of CTAs, so the reduction overlaps the rest of backward.  Falls back to NCCL when
        the fused path is unavailable (CPU/gloo, multi-node, FLAGS_use_fused_comm=0)."""
        from ..parallel import symm
        if not (params and params[0].is_cuda and symm.available(self.group)):
            return None, None
        total = sum((p.numel() + 127) // 128 * 128 * p.element_size() for p in params)
        try:
            arena = symm.get_arena(self.group, total + (8 << 20), key=f"dp{id(self)}")
        except Exception:  # noqa: BLE001 - e.g. driver without VMM export; NCCL path still works
            return None, None
        side = torch.cuda.Stream()

        def allocator(numel, dtype, device):
            t = arena.empty((numel,), dtype)
            t.zero_()
            return t

        class _EvTask:
            def __init__(self, ev):
                self.ev = ev

            def wait(self):
                torch.cuda.current_stream().wait_event(self.ev)

            def is_completed(self):
                return self.ev.query()

        def fused(buf, nranks):
            ready = torch.cuda.Event()
            ready.record()
            with torch.cuda.stream(side):
                side.wait_event(ready)
                symm.allreduce_(buf, scale=1.0 / nranks, max_ctas=16)
                done = torch.cuda.Event()
                done.record(side)
            return _EvTask(done)
        return allocator, fused

    def forward(self, *inputs, **kwargs):
        outputs = self._layers(*inputs, **kwargs)
        if self._reducer is not None and self.grad_need_sync and torch.is_grad_enabled():
            self._reducer.prepare_for_backward(outputs)
        return outputs

    @contextlib.contextmanager
    def no_sync(self):
        prev = self.grad_need_sync
        self.grad_need_sync = False
        if self._reducer is not None:
            self._reducer.enabled = False
        try:
            yield
        finally:
            self.grad_need_sync = prev
            if self._reducer is not None:
                self._reducer.enabled = prev

    def scale_loss(self, loss):
        return loss

    def apply_collective_grads(self):
        return

    def state_dict(self, *args, **kwargs):
        return self._layers.state_dict(*args, **kwargs)

    def set_state_dict(self, state_dict, use_structured_name=True):
        return self._layers.set_state_dict(state_dict, use_structured_name)

    set_dict = set_state_dict
    load_dict = set_state_dict

    def parameters(self, *a, **k):
        return self._layers.parameters(*a, **k)

    def named_parameters(self, *a, **k):
        return self._layers.named_parameters(*a, **k)

    def sublayers(self, *a, **k):
        return self._layers.sublayers(*a, **k)

    def train(self, mode=True):
        self._layers.train(mode)
        return super().train(mode)

    def eval(self):
        self._layers.eval()
        return super().eval()

    def clear_gradients(self, *a, **k):
        return self._layers.clear_gradients(*a, **k)
