"""EagerReducer: bucketed data-parallel gradient all-reduce overlapped with backward
(reference: paddle/fluid/distributed/collective/reducer.{h,cc}: Eager_AssignGroupBySize :47-158, hook registration
:483-538, MarkVarReady/MarkGroupReady :798-985, FusedAllReduceSchedule :1093-1127, FinalizeBackward :1069-1091).

Differences by design (B200-first):
* no concat/split kernels - each bucket is one persistent flat buffer and every ``param.grad`` is a view into it, so
  autograd accumulates straight into the bucket;
* the 1/nranks scale is folded into the reduction (NCCL AVG on the baseline path; the fused NVLS kernel of
  ``paddle_b200.parallel.symm`` does scale + multimem all-reduce in one launch when FLAGS_use_fused_comm is on and
  the bucket lives in symmetric memory);
* buckets launch strictly in order (Appendix B.3) from post-accumulate hooks; the tail is awaited in a callback
  queued on the autograd engine, which only makes the compute stream wait - the host never blocks.
Bucket assignment and the per-iteration bookkeeping (which parameter was marked, which bucket filled, which buckets may
be reduced now - strictly in bucket order - and which parameters stayed unused) are native
(libpdb200_runtime.so ``pdb200_assign_group_by_size``, ``pdb200_reducer_*`` in csrc/runtime/reducer.cc); this module does
the tensor work and issues the collectives.  A Python mirror of the bookkeeping serves builds without the library.

Sparse gradients (``Embedding(sparse=True)``; reference AllReduceSparse :1129-1350) live in singleton buckets and are
reduced by gathering every rank's (row ids, values) and coalescing them - the dense bucket machinery never sees them.
With ``find_unused_parameters`` every rank contributes a "used in this backward" bitmap that is MAX-reduced over the
group (reference local_used_vars_ :996-1067): a parameter no rank used keeps ``grad = None`` (so weight decay or
momentum do not move it), one that any rank used takes part in the all-reduce with zeros where it was unused.
"""
from __future__ import annotations

import ctypes

import torch

from .. import lib
from ..framework.flags import flag
from . import communication as comm
from .collective import ReduceOp

_DT_ID = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2, torch.float64: 3}


def _assign_py(sizes_bytes, dtype_ids, is_sparse, limits):
    res, next_group, limit_idx = [], {}, {}
    for i, (sz, dt) in enumerate(zip(sizes_bytes, dtype_ids)):
        if is_sparse[i]:
            res.append([i])
            continue
        idxs, tot = next_group.get(dt, ([], 0))
        idxs = idxs + [i]
        tot += sz
        li = limit_idx.setdefault(dt, 0)
        if tot >= limits[li]:
            res.append(idxs)
            next_group[dt] = ([], 0)
            limit_idx[dt] = min(li + 1, len(limits) - 1)
        else:
            next_group[dt] = (idxs, tot)
    for dt in sorted(next_group):
        if next_group[dt][0]:
            res.append(next_group[dt][0])
    res.sort(key=lambda g: g[0])
    return res


def assign_group_by_size(tensors, is_sparse_gradient, group_size_limits, tensor_indices=None):
    """core.eager_assign_group_by_size: per-dtype buckets, limits [first, rest...] in bytes."""
    sizes = [t.numel() * t.element_size() for t in tensors]
    dts = [_DT_ID.get(t.dtype, 9) for t in tensors]
    rt = lib.runtime
    if rt.available() and hasattr(rt.require(), "pdb200_assign_group_by_size"):
        n = len(tensors)
        a_sizes = (ctypes.c_int64 * n)(*sizes)
        a_dts = (ctypes.c_int * n)(*dts)
        a_sp = (ctypes.c_int * n)(*[int(b) for b in is_sparse_gradient])
        a_lim = (ctypes.c_int64 * len(group_size_limits))(*[int(x) for x in group_size_limits])
        out_group = (ctypes.c_int * n)()
        ng = rt.pdb200_assign_group_by_size(ctypes.cast(a_sizes, ctypes.c_void_p), ctypes.cast(a_dts, ctypes.c_void_p),
                                            ctypes.cast(a_sp, ctypes.c_void_p), n,
                                            ctypes.cast(a_lim, ctypes.c_void_p), len(group_size_limits),
                                            ctypes.cast(out_group, ctypes.c_void_p))
        if ng < 0:
            raise RuntimeError("pdb200_assign_group_by_size failed")
        groups = [[] for _ in range(ng)]
        for i in range(n):
            groups[out_group[i]].append(i)
        return groups
    return _assign_py(sizes, dts, is_sparse_gradient, [int(x) for x in group_size_limits])


class _Bucket:
    def __init__(self, index, params, device):
        self.index = index
        self.params = params
        self.dtype = params[0].dtype
        self.offsets = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + 127) // 128 * 128   # keep every view 256-B aligned (vector loads, NVLS atoms)
        self.numel = off
        self.buffer = None
        self.device = device
        self.pending = len(params)
        self.ready = False
        self.task = None
        self.marked = set()

    def allocate(self, allocator=None):
        if self.buffer is None:
            self.buffer = (allocator(self.numel, self.dtype, self.device) if allocator is not None
                           else torch.zeros(self.numel, dtype=self.dtype, device=self.device))
        return self.buffer

    def view(self, i):
        p = self.params[i]
        return self.buffer[self.offsets[i]:self.offsets[i] + p.numel()].view(p.shape)


class EagerReducer:
    def __init__(self, params, group_indices, is_sparse_gradient, group, group_size_limits=None,
                 find_unused_parameters=False, bucket_allocator=None, fused_allreduce=None):
        self.params = list(params)
        self.group = group
        self.nranks = group.nranks
        self.find_unused = find_unused_parameters
        self.enabled = True
        self._allocator = bucket_allocator
        self._fused = fused_allreduce          # callable(bucket_buffer, nranks) -> Task, or None for NCCL
        self.buckets = [_Bucket(i, [self.params[j] for j in idxs], self.params[idxs[0]].device)
                        for i, idxs in enumerate(group_indices)]
        sparse = list(is_sparse_gradient) if is_sparse_gradient is not None else [False] * len(self.params)
        self._sparse_ids = {id(p) for p, sp in zip(self.params, sparse) if sp}
        self._where = {}
        for b in self.buckets:
            b.sparse = len(b.params) == 1 and id(b.params[0]) in self._sparse_ids
            if not b.sparse:
                b.allocate(self._allocator)
            for i, p in enumerate(b.params):
                self._where[id(p)] = (b, i)
        self._index_of = {id(p): i for i, p in enumerate(self.params)}
        self._core = None
        if lib.runtime.available() and hasattr(lib.runtime.require(), "pdb200_reducer_create"):
            n = len(self.params)
            bucket_of = (ctypes.c_int32 * max(n, 1))(*[self._where[id(p)][0].index for p in self.params])
            sparse_b = (ctypes.c_uint8 * max(len(self.buckets), 1))(*[int(b.sparse) for b in self.buckets])
            self._core = lib.runtime.pdb200_reducer_create(ctypes.cast(bucket_of, ctypes.c_void_p), n,
                                                           ctypes.cast(sparse_b, ctypes.c_void_p), len(self.buckets))
            self._launch_buf = (ctypes.c_int32 * max(len(self.buckets), 1))()
            self._missing_buf = (ctypes.c_int32 * max(n, 1))()
            self._n_launch = ctypes.c_int32(0)
        self._used_task = None
        self._used_bitmap = None
        self.next_bucket = 0
        self._armed = False
        self._callback_queued = False
        self._hooks = [p.register_post_accumulate_grad_hook(self._make_hook(p)) for p in self.params]

    # --------------------------------------------------------------- per-iteration state
    def prepare_for_backward(self, outputs=None):
        self._armed = True
        self._callback_queued = False
        self.next_bucket = 0
        if self._core:
            lib.runtime.pdb200_reducer_prepare(self._core)
        for b in self.buckets:
            b.pending = len(b.params)
            b.ready = False
            b.task = None
            b.marked = set()

    def _make_hook(self, p):
        def hook(param):
            if not (self.enabled and self._armed):
                return
            self._mark_ready(param)
        return hook

    @torch.no_grad()
    def _mark_ready(self, p):
        if self._core:
            return self._mark_ready_native(p)
        b, i = self._where[id(p)]
        if id(p) in b.marked:
            return
        b.marked.add(id(p))
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self.finalize_backward)
        if b.sparse:
            b.pending = 0
            b.ready = True
            self._launch_ready()
            return
        view = b.view(i)
        g = p.grad
        if g is not None and g.data_ptr() != view.data_ptr():
            view.copy_(g)
            p.grad = view          # from now on autograd accumulates into the bucket directly
        elif g is None:
            view.zero_()
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self.finalize_backward)
        b.pending -= 1
        if b.pending == 0:
            b.ready = True
            self._launch_ready()

    def _adopt(self, b, i, p):
        """The gradient of p lives in (or moves into) its slot of the bucket buffer."""
        view = b.view(i)
        g = p.grad
        if g is not None and g.data_ptr() != view.data_ptr():
            view.copy_(g)
            p.grad = view          # from now on autograd accumulates into the bucket directly
        elif g is None:
            view.zero_()

    def _mark_ready_native(self, p):
        b, i = self._where[id(p)]
        if id(p) not in b.marked and not b.sparse:
            self._adopt(b, i, p)                          # before the core may release the bucket
        n = lib.runtime.pdb200_reducer_mark_ready(self._core, self._index_of[id(p)],
                                                  ctypes.cast(self._launch_buf, ctypes.c_void_p), len(self.buckets))
        if n < 0:
            return                                        # not armed, or marked before in this iteration
        b.marked.add(id(p))
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self.finalize_backward)
        for k in range(n):
            bk = self.buckets[self._launch_buf[k]]
            bk.ready = True
            bk.task = self._all_reduce(bk)
            self.next_bucket = bk.index + 1

    def _launch_ready(self):
        while self.next_bucket < len(self.buckets) and self.buckets[self.next_bucket].ready:
            b = self.buckets[self.next_bucket]
            b.task = self._all_reduce(b)
            self.next_bucket += 1

    def _all_reduce(self, b):
        if b.sparse:
            return self._all_reduce_sparse(b.params[0])
        if self._fused is not None and flag("use_fused_comm"):
            return self._fused(b.buffer, self.nranks)
        return comm.all_reduce(b.buffer, op=ReduceOp.AVG, group=self.group, sync_op=False)

    @torch.no_grad()
    def _all_reduce_sparse(self, p):
        """AVG all-reduce of a sparse (COO) gradient: every rank's (row ids, values) are gathered - padded to the
        largest count - concatenated and coalesced (reference: reducer.cc AllReduceSparse gathers rows and values of
        the SelectedRows).  A rank without a gradient contributes zero rows."""
        g = p.grad
        dev = p.device
        if g is None:
            idx = torch.zeros(1, 0, dtype=torch.int64, device=dev)
            val = torch.zeros((0,) + tuple(p.shape[1:]), dtype=p.dtype, device=dev)
        else:
            g = g.coalesce() if g.is_sparse else g.to_sparse(1)
            idx, val = g.indices(), g.values()
        n = self.nranks
        cnt = torch.tensor([idx.shape[1]], dtype=torch.int64, device=dev)
        cnts = [torch.zeros_like(cnt) for _ in range(n)]
        comm.all_gather(cnts, cnt, self.group)
        cnts = [int(c) for c in cnts]               # row counts are data dependent: one host read per sparse gradient
        mx = max(max(cnts), 1)
        pad_i = torch.zeros(mx, dtype=torch.int64, device=dev)
        pad_i[: idx.shape[1]] = idx[0]
        pad_v = torch.zeros((mx,) + tuple(p.shape[1:]), dtype=p.dtype, device=dev)
        pad_v[: val.shape[0]] = val
        all_i = [torch.empty_like(pad_i) for _ in range(n)]
        all_v = [torch.empty_like(pad_v) for _ in range(n)]
        comm.all_gather(all_i, pad_i, self.group)
        comm.all_gather(all_v, pad_v, self.group)
        rows = torch.cat([all_i[r][: cnts[r]] for r in range(n)])
        vals = torch.cat([all_v[r][: cnts[r]] for r in range(n)]) / n
        p.grad = torch.sparse_coo_tensor(rows[None], vals, size=tuple(p.shape), device=dev).coalesce()
        return None

    @torch.no_grad()
    def _reduce_used_bitmap(self):
        """MAX all-reduce of "received a gradient in this backward" over the group."""
        dev = self.params[0].device
        used = torch.zeros(len(self.params), dtype=torch.int32, device=dev)
        marked = [self._index_of[id(p)] for b in self.buckets for p in b.params if id(p) in b.marked]
        if marked:
            used[torch.tensor(marked, dtype=torch.int64, device=dev)] = 1
        comm.all_reduce(used, op=ReduceOp.MAX, group=self.group)
        return used

    @torch.no_grad()
    def _finalize_native(self):
        n_params = len(self.params)
        missing = lib.runtime.pdb200_reducer_finalize(self._core, int(self.find_unused),
                                                      ctypes.cast(self._missing_buf, ctypes.c_void_p), n_params,
                                                      ctypes.cast(self._launch_buf, ctypes.c_void_p), len(self.buckets),
                                                      ctypes.byref(self._n_launch))
        if missing and not self.find_unused:
            names = [self.params[self._missing_buf[k]].name for k in range(min(missing, 4))]
            self._armed = False
            raise RuntimeError(
                "DataParallel: some parameters did not receive a gradient in this backward "
                f"({names}...). Construct DataParallel with find_unused_parameters=True.")
        for k in range(min(missing, n_params)):           # unused here: zeros take part in the reduction
            p = self.params[self._missing_buf[k]]
            b, i = self._where[id(p)]
            if not b.sparse:
                b.view(i).zero_()
                p.grad = b.view(i)
        for k in range(self._n_launch.value):
            bk = self.buckets[self._launch_buf[k]]
            bk.ready = True
            bk.task = self._all_reduce(bk)
            self.next_bucket = bk.index + 1
        globally_unused = []
        if self.find_unused:
            used = (ctypes.c_int32 * max(n_params, 1))()
            lib.runtime.pdb200_reducer_used_bitmap(self._core, ctypes.cast(used, ctypes.c_void_p))
            bitmap = torch.tensor(list(used)[:n_params], dtype=torch.int32, device=self.params[0].device)
            comm.all_reduce(bitmap, op=ReduceOp.MAX, group=self.group)
            self._used_bitmap = bitmap
            globally_unused = [self.params[i] for i, u in enumerate(bitmap.tolist()) if not u]
        for b in self.buckets:
            if b.task is not None:
                b.task.wait()
                b.task = None
        for p in globally_unused:            # nobody produced a gradient for it: keep "no gradient", not zeros
            p.grad = None
        self._armed = False

    @torch.no_grad()
    def finalize_backward(self):
        if not self._armed:
            return
        if self._core:
            return self._finalize_native()
        globally_unused = []
        # buckets that never became ready hold unused parameters
        for b in self.buckets[self.next_bucket:]:
            if not b.ready:
                if not self.find_unused:
                    missing = [p.name for p in b.params if self._is_missing(b, p)]
                    self._armed = False
                    raise RuntimeError(
                        "DataParallel: some parameters did not receive a gradient in this backward "
                        f"({missing[:4]}...). Construct DataParallel with find_unused_parameters=True.")
                if not b.sparse:
                    for i, p in enumerate(b.params):
                        if self._is_missing(b, p):
                            b.view(i).zero_()
                            p.grad = b.view(i)
                b.ready = True
        self._launch_ready()
        if self.find_unused:
            # after the last bucket so that every rank issues its collectives in the same order; every rank takes part
            # every step, because which parameters are unused may differ from rank to rank
            self._used_bitmap = self._reduce_used_bitmap()
            globally_unused = [self.params[i] for i, u in enumerate(self._used_bitmap.tolist()) if not u]
        for b in self.buckets:
            if b.task is not None:
                b.task.wait()
                b.task = None
        for p in globally_unused:            # nobody produced a gradient for it: keep "no gradient", not zeros
            p.grad = None
        self._armed = False

    def _is_missing(self, b, p):
        return id(p) not in b.marked

    def remove_hooks(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []

    def __del__(self):
        core, self._core = getattr(self, "_core", None), None
        if core:
            try:
                lib.runtime.pdb200_reducer_destroy(core)
            except Exception:  # noqa: BLE001  (interpreter shutdown)
                pass
