"""Symmetric-memory communication layer.

``SymmArena`` wraps the native arena (csrc/runtime/symm.cc: cuMemCreate + POSIX-FD exchange + peer mappings + an NVLS
multimem window) for one process group: an allocator hands out byte offsets that are identical on every rank, torch
tensors view the LOCAL mapping only, and the fused SIMT kernels (csrc/kernels/nvl_*.cu) receive the array of peer base
pointers plus the NVLS pointer.  Peer mappings and the NVLS window are registered with the kernel library as foreign
ranges: the tensor-core launchers (TMA tensor maps) refuse any operand inside them, so GEMM / attention operands are
always this GPU's own memory.  The first 4 KB of every arena is the signal pad used by the in-stream
cross-GPU barriers; epochs grow monotonically so the pad is never reset.

Every fused path has its NCCL composition behind the same call signature; ``FLAGS_use_fused_comm`` (and
availability: GPUs of one node, native libraries built, driver support) picks the transport.
"""
from __future__ import annotations

import ctypes
import os
import tempfile

import torch

from paddle_b200 import lib
from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm
from paddle_b200.framework.flags import flag
from paddle_b200.ops._util import dtype_code, stream_ptr

PAD_BYTES = 4096
_arenas = {}
_counter = [0]
_disabled_reason = [None]


class _Blob:
    """Expose raw device memory to torch through __cuda_array_interface__ (zero-copy view)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class SymmArena:
    def __init__(self, group, nbytes, want_multicast=True):
        rt = lib.runtime.require()
        self.group = group
        self.nranks, self.rank = group.nranks, group.rank
        dev = torch.cuda.current_device()
        _counter[0] += 1
        # rank 0 of the group picks a fresh rendezvous directory; everyone learns it through the group
        box = [tempfile.mkdtemp(prefix="pdb200_symm_") if self.rank == 0 else None]
        comm.broadcast_object_list(box, src=group.ranks[0], group=group)
        self._dir = box[0]
        tag = f"g{group.id}_{_counter[0]}"
        self._h = rt.pdb200_symm_create(self.rank, self.nranks, dev, int(nbytes) + PAD_BYTES, self._dir.encode(), tag.encode(),
                                        int(want_multicast), 60000)
        if not self._h:
            raise RuntimeError(f"symmetric arena: {rt.pdb200_rt_last_error().decode()}")
        self._rt = rt
        self.nbytes = rt.pdb200_symm_bytes(self._h)
        self.peer_ptrs = (ctypes.c_uint64 * self.nranks)(*[rt.pdb200_symm_peer_ptr(self._h, p) for p in range(self.nranks)])
        self.mc_ptr = rt.pdb200_symm_multicast_ptr(self._h)
        self.local_ptr = self.peer_ptrs[self.rank]
        # every mapping of somebody else's memory is off limits for TMA / tensor-core operands
        self._foreign = [int(self.peer_ptrs[p]) for p in range(self.nranks) if p != self.rank]
        if self.mc_ptr:
            self._foreign.append(int(self.mc_ptr))
        if lib.have_kernels():
            for base in self._foreign:
                lib.kernels.pdb200_register_foreign_range(base, self.nbytes)
        self._bytes = torch.as_tensor(_Blob(self.local_ptr, self.nbytes), device=f"cuda:{dev}")
        self._bytes[:PAD_BYTES].zero_()
        torch.cuda.synchronize()
        C.barrier(group)                       # pads are zero everywhere before the first in-stream barrier
        # offsets come from a native best-fit heap with coalescing free (csrc/runtime/heap.cc); it is a pure function of
        # the alloc/free call sequence, so ranks running the same program agree on every offset
        self._heap = rt.pdb200_heap_create(self.nbytes, PAD_BYTES)
        self._epoch = 1

    @property
    def peers_arg(self):
        return ctypes.cast(self.peer_ptrs, ctypes.c_void_p)

    def next_epoch(self):
        e = self._epoch
        self._epoch += 2
        return e

    def alloc(self, nbytes, align=256):
        off = self._rt.pdb200_heap_alloc(self._heap, int(nbytes), int(align))
        if off < 0:
            off = _evict_cached(self, int(nbytes), int(align))
        if off < 0:
            raise MemoryError(f"symmetric arena exhausted: need {nbytes} B, largest free block {self.largest_free()} B of "
                              f"{self.free_bytes()} B free, capacity {self.nbytes} B (PDB200_SYMM_ARENA_MB sets the size)")
        return off

    def free(self, t_or_offset):
        """Give a buffer back (the tensor from ``empty`` or its offset).  Collective in the same sense as ``empty``: every
        rank must free in the same order, and nobody - no kernel of this rank, no peer - may still be using the bytes."""
        off = t_or_offset._symm[1] if hasattr(t_or_offset, "_symm") else int(t_or_offset)
        if self._rt.pdb200_heap_free(self._heap, off) != 0:
            raise ValueError(f"symmetric arena: offset {off} is not a live allocation")

    def free_bytes(self):
        return self._rt.pdb200_heap_stat(self._heap, 0)

    def largest_free(self):
        return self._rt.pdb200_heap_stat(self._heap, 1)

    def stats(self):
        st = self._rt.pdb200_heap_stat
        return dict(capacity=self.nbytes, free=st(self._heap, 0), largest_free=st(self._heap, 1), live_blocks=st(self._heap, 2),
                    in_use=st(self._heap, 3), peak_in_use=st(self._heap, 4), free_blocks=st(self._heap, 5))

    def tensor(self, offset, shape, dtype):
        n = 1
        for s in shape:
            n *= int(s)
        nb = n * torch.empty(0, dtype=dtype).element_size()
        return self._bytes[offset:offset + nb].view(dtype).view(*shape)

    def empty(self, shape, dtype):
        n = 1
        for s in shape:
            n *= int(s)
        off = self.alloc(n * torch.empty(0, dtype=dtype).element_size())
        t = self.tensor(off, shape, dtype)
        t._symm = (self, off)
        return t

    def close(self):
        if getattr(self, "_h", None):
            torch.cuda.synchronize()
            if lib.have_kernels():
                for base in self._foreign:
                    lib.kernels.pdb200_unregister_foreign_range(base)
            self._rt.pdb200_symm_destroy(self._h)
            self._h = None
            self._rt.pdb200_heap_destroy(self._heap)
            self._heap = None
            for cache in _caches:
                for k in [k for k in cache if k[0] == id(self)]:
                    del cache[k]


def available(group=None) -> bool:
    """Fused NVLink paths are usable: flag on, GPUs, NCCL backend, one node, native libs, driver VMM support."""
    if _disabled_reason[0] is not None or not flag("use_fused_comm"):
        return False
    if not torch.cuda.is_available() or C._state.backend != "nccl":
        return False
    if group is not None and (group.nranks < 2 or group.nranks > 8 or group.rank < 0):
        return False
    if int(os.environ.get("PADDLE_NNODES", "1")) > 1:
        return False
    if not (lib.runtime.available() and lib.have_kernels()):
        return False
    if not hasattr(lib.kernels.require(), "pdb200_nvl_allreduce"):
        return False
    return lib.runtime.pdb200_symm_supported(torch.cuda.current_device()) >= 1


def get_arena(group, min_bytes=int(os.environ.get("PDB200_SYMM_ARENA_MB", "512")) << 20, key="default"):
    k = (group.id, key)
    a = _arenas.get(k)
    if a is None:
        try:
            a = SymmArena(group, min_bytes)
        except Exception as e:  # noqa: BLE001
            _disabled_reason[0] = str(e)
            raise
        _arenas[k] = a
    return a


def release_all():
    for a in _arenas.values():
        a.close()
    _arenas.clear()


# --------------------------------------------------------------------------------------------- fused collectives
def allreduce_(t, scale=1.0, max_ctas=0):
    """In-place scaled all-reduce of a tensor that lives in a symmetric arena (K1)."""
    arena, off = t._symm
    from paddle_b200.lib import check
    check(lib.kernels.pdb200_nvl_allreduce(arena.peers_arg, arena.mc_ptr, arena.nranks, arena.rank, off, t.numel(),
                                           dtype_code(t), float(scale), 0, arena.next_epoch(), int(max_ctas),
                                           stream_ptr(t.device)), "nvl_allreduce")
    return t


def barrier(arena):
    from paddle_b200.lib import check
    check(lib.kernels.pdb200_nvl_barrier(arena.peers_arg, arena.nranks, arena.rank, 0, arena.next_epoch(),
                                         stream_ptr(torch.device("cuda", torch.cuda.current_device()))), "nvl_barrier")


def rs_adamw_ag(grad_buf, param_buf, shard_start, shard_numel, master, m, v, lr, beta1, beta2, eps, weight_decay,
                beta_pows, grad_scale=None, found_inf=None, local_grad=None):
    """K2+K3: reduce-scatter(AVG) of ``grad_buf`` + AdamW on this rank's shard + all-gather into ``param_buf``.
    ``local_grad`` (fp32, shard-sized): the gradient was already reduced by ``reduce_scatter_norm`` (two-pass mode
    needed when a global-norm clip must see the reduced gradient first); the kernel then only updates + gathers."""
    arena, goff = grad_buf._symm
    arena2, poff = param_buf._symm
    assert arena is arena2, "gradient and parameter buffers must live in the same arena"
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    check(lib.kernels.pdb200_nvl_rs_adamw_ag(arena.peers_arg, arena.mc_ptr, arena.nranks, arena.rank, goff,
                                             dtype_code(grad_buf), poff, int(shard_start), int(shard_numel), ptr(master),
                                             ptr(m), ptr(v), float(lr), float(beta1), float(beta2), float(eps),
                                             float(weight_decay), ptr(beta_pows), ptr(grad_scale), ptr(found_inf),
                                             ptr(local_grad), 0, arena.next_epoch(), stream_ptr(grad_buf.device)),
          "nvl_rs_adamw_ag")


def reduce_scatter_norm(grad_buf, shard_start, shard_numel, out):
    """Pass 1 of the clipped update: AVG reduce-scatter into ``out`` (fp32 shard); returns the shard's squared norm
    as a device scalar."""
    arena, goff = grad_buf._symm
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    grid = lib.kernels.pdb200_nvl_rs_grid(int(shard_numel))
    partial = torch.empty(grid, dtype=torch.float32, device=grad_buf.device)
    check(lib.kernels.pdb200_nvl_reduce_scatter(arena.peers_arg, arena.mc_ptr, arena.nranks, arena.rank, goff,
                                                dtype_code(grad_buf), int(shard_start), int(shard_numel), ptr(out),
                                                ptr(partial), 0, arena.next_epoch(), stream_ptr(grad_buf.device)),
          "nvl_reduce_scatter")
    return partial.sum().reshape(1)


def all_gather_push(shard, full, max_ctas=0):
    """K3 stand-alone: ``full[r*len(shard):(r+1)*len(shard)] = rank r's shard`` on every rank.  Both tensors live in
    the same symmetric arena (``full`` at the same offset on every rank); runs on the current stream."""
    arena, soff = shard._symm
    arena2, doff = full._symm
    assert arena is arena2, "shard and gathered buffer must share one arena"
    from paddle_b200.lib import check
    check(lib.kernels.pdb200_nvl_allgather(arena.peers_arg, arena.mc_ptr, arena.nranks, arena.rank, soff, doff,
                                           shard.numel() * shard.element_size(), 0, arena.next_epoch(), int(max_ctas),
                                           stream_ptr(shard.device)), "nvl_allgather")
    return full


def _multi(arena, fn_name, segs, mid, max_ctas, stream):
    """Launch a multi-segment transport kernel, 16 segments per launch; each launch takes two pad epochs."""
    from paddle_b200.lib import check
    fn = getattr(lib.kernels, fn_name)
    for i in range(0, len(segs), 16):
        chunk = segs[i:i + 16]
        flat = (ctypes.c_int64 * (3 * len(chunk)))(*[int(v) for seg in chunk for v in seg])
        check(fn(arena.peers_arg, arena.mc_ptr, arena.nranks, arena.rank, ctypes.cast(flat, ctypes.c_void_p), len(chunk),
                 *mid, 0, arena.next_epoch(), int(max_ctas), stream), fn_name)


def allgather_multi(arena, segs, max_ctas=0, stream=None):
    """One launch all-gathers several shards: segs = [(src_off, dst_off, shard_bytes), ...] (byte offsets inside the
    arena, identical on all ranks); rank r's shard lands at ``dst_off + r * shard_bytes`` on every rank."""
    st = stream if stream is not None else stream_ptr(torch.device("cuda", torch.cuda.current_device()))
    _multi(arena, "pdb200_nvl_allgather_multi", segs, (), max_ctas, st)


def allgather_copy_engine(arena, segs, stream=None):
    """The same all-gather on the copy engines: one device-to-device copy per (segment, peer) through the arena's peer
    mappings, bracketed by the arena's cross-rank barriers.  No SM runs a copy loop, so the transfer takes neither
    issue slots nor L2 bandwidth from the kernels it overlaps; the price is 2 tiny barrier launches and
    ``len(segs) * nranks`` copy submissions (issued from native code)."""
    from paddle_b200.lib import check
    st = stream if stream is not None else stream_ptr(torch.device("cuda", torch.cuda.current_device()))
    e = arena.next_epoch()
    check(lib.kernels.pdb200_nvl_barrier(arena.peers_arg, arena.nranks, arena.rank, 0, e, st), "nvl_barrier")
    flat = (ctypes.c_int64 * (3 * len(segs)))(*[int(v) for seg in segs for v in seg])
    if arena._rt.pdb200_symm_allgather_ce(arena._h, ctypes.cast(flat, ctypes.c_void_p), len(segs), st) != 0:
        raise RuntimeError(f"symm_allgather_ce: {arena._rt.pdb200_rt_last_error().decode()}")
    check(lib.kernels.pdb200_nvl_barrier(arena.peers_arg, arena.nranks, arena.rank, 0, e + 1, st), "nvl_barrier")


def reduce_scatter_multi(arena, segs, grad_dtype, scale, accumulate=False, max_ctas=0, stream=None):
    """One launch reduce-scatters several staged buffers: segs = [(stage_off, shard_numel, out_ptr), ...];
    ``out (+)= scale * sum over ranks of stage[rank * shard_numel : (rank + 1) * shard_numel]`` (fp32, local)."""
    st = stream if stream is not None else stream_ptr(torch.device("cuda", torch.cuda.current_device()))
    code = {torch.float32: 0, torch.bfloat16: 1}[grad_dtype]
    _multi(arena, "pdb200_nvl_reduce_scatter_multi", segs, (code, float(scale), int(bool(accumulate))), max_ctas, st)


def reduce_scatter_copy_engine(arena, segs, grad_dtype, scale, accumulate=False, max_ctas=0, stream=None):
    """Reduce-scatter whose transfer runs on the copy engines: segs = [(stage_off, recv_off, shard_numel, out_ptr), ...];
    shard p of every staged buffer is copied into slot ``rank`` of rank p's receive buffer (``recv_off``, nranks slots),
    then a local kernel writes ``out (+)= scale * sum of the slots``.  Only that local sum runs on SMs."""
    from paddle_b200.lib import check
    st = stream if stream is not None else stream_ptr(torch.device("cuda", torch.cuda.current_device()))
    esz = 4 if grad_dtype == torch.float32 else 2
    e = arena.next_epoch()
    check(lib.kernels.pdb200_nvl_barrier(arena.peers_arg, arena.nranks, arena.rank, 0, e, st), "nvl_barrier")   # receive slots free
    flat = (ctypes.c_int64 * (3 * len(segs)))(*[int(v) for (so, ro, ne, _o) in segs for v in (so, ro, ne * esz)])
    if arena._rt.pdb200_symm_alltoall_ce(arena._h, ctypes.cast(flat, ctypes.c_void_p), len(segs), st) != 0:
        raise RuntimeError(f"symm_alltoall_ce: {arena._rt.pdb200_rt_last_error().decode()}")
    check(lib.kernels.pdb200_nvl_barrier(arena.peers_arg, arena.nranks, arena.rank, 0, e + 1, st), "nvl_barrier")   # all slots landed
    code = {torch.float32: 0, torch.bfloat16: 1}[grad_dtype]
    for i in range(0, len(segs), 16):
        chunk = segs[i:i + 16]
        tri = (ctypes.c_int64 * (3 * len(chunk)))(*[int(v) for (_so, ro, ne, o) in chunk for v in (arena.local_ptr + ro, ne, o)])
        check(lib.kernels.pdb200_reduce_slots_multi(ctypes.cast(tri, ctypes.c_void_p), len(chunk), arena.nranks, code, float(scale),
                                                    int(bool(accumulate)), int(max_ctas), st), "reduce_slots_multi")


def reduce_scatter_avg(grad_buf, shard_start, shard_numel, out, scratch):
    """AVG reduce-scatter of a symmetric gradient buffer into a local fp32 shard (no host sync, norm partials in
    ``scratch``, which must hold lib.kernels.pdb200_nvl_rs_grid(shard_numel) floats)."""
    arena, goff = grad_buf._symm
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    check(lib.kernels.pdb200_nvl_reduce_scatter(arena.peers_arg, arena.mc_ptr, arena.nranks, arena.rank, goff,
                                                dtype_code(grad_buf), int(shard_start), int(shard_numel), ptr(out),
                                                ptr(scratch), 0, arena.next_epoch(), stream_ptr(grad_buf.device)),
          "nvl_reduce_scatter")
    return out


_ce_scratch = {}


def ce_exchange_combine(stats, labels, group, ignore_index):
    """K8: publish this rank's [rows, 3] statistics in symmetric scratch, read every peer's copy in-kernel."""
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    arena = get_arena(group)
    rows = stats.shape[0]
    key = (group.id, rows)
    buf = _ce_scratch.get(key)
    if buf is None:
        buf = arena.empty((rows, 3), torch.float32)
        _ce_scratch[key] = buf
    buf.copy_(stats)
    lab = labels.reshape(-1).long().contiguous()
    loss = torch.empty(rows, dtype=torch.float32, device=stats.device)
    lse = torch.empty(rows, dtype=torch.float32, device=stats.device)
    check(lib.kernels.pdb200_nvl_ce_combine(arena.peers_arg, arena.nranks, arena.rank, buf._symm[1], ptr(lab), ptr(loss),
                                            ptr(lse), rows, int(ignore_index), 0, arena.next_epoch(),
                                            stream_ptr(stats.device)), "nvl_ce_combine")
    return loss, lse


class _ShapeCache(dict):
    """Per-shape staging state in an arena, least recently used first.  ``touch`` on every hit; ``_evict_cached`` pops
    from the front when the arena runs out.  Keys start with id(arena)."""

    def __init__(self):
        super().__init__()
        self._used = {}

    def touch(self, key):
        _tick[0] += 1
        self._used[key] = _tick[0]
        return self[key]

    def __setitem__(self, key, value):
        _tick[0] += 1
        self._used[key] = _tick[0]
        super().__setitem__(key, value)

    def __delitem__(self, key):
        self._used.pop(key, None)
        super().__delitem__(key)

    def pop(self, key, *default):
        self._used.pop(key, None)
        return super().pop(key, *default)

    def lookup(self, key):
        return self.touch(key) if key in self else None


_tick = [0]
_stage = _ShapeCache()
_caches = [_stage]


def _state_tensors(v):
    if isinstance(v, torch.Tensor):
        return [v] if hasattr(v, "_symm") else []
    if isinstance(v, dict):
        v = list(v.values())
    if isinstance(v, (list, tuple)):
        return [t for x in v for t in _state_tensors(x)]
    return []


def _evict_cached(arena, nbytes, align):
    """The arena is out of room: drop least-recently-used staging states (gathered buffers, partial-sum buffers, flag
    words cached per shape) one by one until the request fits, and return its offset (-1: nothing left to drop).  Every
    rank runs the same call sequence, so every rank evicts the same entries here; a peer may still be writing into a
    buffer of an earlier call, hence the device synchronisation and the group barrier before anything is reused."""
    victims = []
    for cache in _caches:
        for k in cache:
            if k[0] == id(arena):
                victims.append((cache._used.get(k, 0), id(cache), cache, k))
    if not victims:
        return -1
    victims = [(c, k) for _, _, c, k in sorted(victims, key=lambda v: v[:2])]      # least recently used first
    torch.cuda.synchronize()
    C.barrier(arena.group)
    for cache, k in victims:
        for t in _state_tensors(cache.pop(k)):
            arena.free(t)
        off = arena._rt.pdb200_heap_alloc(arena._heap, nbytes, align)
        if off >= 0:
            return off
    return -1


def _staging(arena, tag, shape, dtype):
    key = (id(arena), tag, tuple(shape), dtype)
    t = _stage.lookup(key)
    if t is not None:
        return t
    t = arena.empty(shape, dtype)
    _stage[key] = t
    return t


def _assert_local(arena, *tensors):
    """Tensor-core operands must be this GPU's memory: arena tensors are views of the LOCAL mapping by construction
    (SymmArena.tensor), anything else is ordinary device memory.  The native launchers re-check against the
    registered foreign ranges."""
    for t in tensors:
        if t is None:
            continue
        p = t.data_ptr()
        for base in arena._foreign:
            assert not (base <= p < base + arena.nbytes), "tensor-core operand inside a mapping of another GPU's memory"


_ag_state = _ShapeCache()
_caches.append(_ag_state)


def _worth_overlapping(M, N, K):
    """The overlapped K4/K5/K6 paths pay a fixed price per call (n-1 copy-engine transfers of a few microseconds each,
    stream memory operations, a second stream): below ~0.15 ms of GEMM the two-kernel path (one NVLS collective kernel +
    the GEMM) is faster.  Threshold in GFLOP per rank, PDB200_OVERLAP_MIN_GFLOP (default 200)."""
    return 2.0 * M * N * K >= float(os.environ.get("PDB200_OVERLAP_MIN_GFLOP", "200")) * 1e9


def _ag_buffers(arena, rows, K, n):
    """Per (arena, shape): two gathered buffers (alternating per call, so a peer may already push the next call's rows
    while this rank still reads the previous ones), one flag word per 256-row chunk and buffer, a side stream, a
    monotonic call counter that doubles as the flag epoch."""
    key = (id(arena), rows, K)
    st = _ag_state.lookup(key)
    if st is None:
        # chunks of >= 256 rows and ~4 MB or more (a copy-engine transfer has a few microseconds of fixed cost), at most
        # 4 per rank; the gate granularity of the GEMM is the chunk
        chunk = 0
        if rows % 256 == 0:
            chunk = rows
            for parts in (4, 2):
                if rows % (parts * 256) == 0 and (rows // parts) * K * 2 >= (4 << 20):
                    chunk = rows // parts
                    break
        nchunks = (rows // chunk) * n if chunk else 0
        st = dict(full=[arena.empty((rows * n, K), torch.bfloat16) for _ in range(2)],
                  flags=[arena.empty((max(nchunks, 1),), torch.int32) for _ in range(2)],
                  stream=torch.cuda.Stream(priority=-1), calls=0, chunk=chunk, per_rank=rows // chunk if chunk else 0)
        for f in st["flags"]:
            f.zero_()
        torch.cuda.current_stream().synchronize()
        C.barrier(arena.group)                     # flags are zero everywhere before anybody signals
        _ag_state[key] = st
    return st


def ag_gemm(x_shard, weight, group, bias=None, trans_b=False, act=None):
    """K5: ``all_gather(x_shard, dim 0) @ weight`` with the transfer hidden behind the GEMM.

    Every rank copies its shard, in 256-row chunks, into all ranks' gathered buffer with the copy engines (a side stream;
    no SM runs a copy loop) and announces each chunk with a system-scope flag write into the receivers' memory.  The
    CTA-pair tcgen05 GEMM starts at once on this rank's own rows and its TMA producer waits, chunk by chunk, for the flag
    of the rows it is about to load - from LOCAL memory, where the peers' copies land.  Two gathered buffers alternate,
    so no rank has to wait for the others before it starts pushing.  x_shard: [rows, K] bf16; returns [nranks*rows, N].
    Falls back to gather-then-GEMM (one NVLS ``multimem.st`` kernel, then the GEMM) when the shape cannot be chunked."""
    from paddle_b200.lib import check
    from paddle_b200.ops import gemm as G
    from paddle_b200.ops._util import ptr
    arena = get_arena(group)
    rows, K = x_shard.shape
    n = arena.nranks
    N = weight.shape[0] if trans_b else weight.shape[1]
    st = _ag_buffers(arena, rows, K, n)
    xs = x_shard.contiguous()
    gated = (st["chunk"] > 0 and flag("ag_gemm_overlap") and _worth_overlapping(rows * n, N, K) and hasattr(lib.kernels.require(), "pdb200_gemm_bf16_gated")
             and hasattr(arena._rt, "pdb200_symm_push_signal") and arena._rt.pdb200_symm_streamops_supported() == 1
             and K % 64 == 0 and N % 8 == 0 and weight.is_contiguous())
    if not gated:
        full = st["full"][0]
        check(lib.kernels.pdb200_nvl_allgather_from(arena.peers_arg, arena.mc_ptr, n, arena.rank, ptr(xs),
                                                    full._symm[1], xs.numel() * 2, 0, arena.next_epoch(), 0,
                                                    stream_ptr(xs.device)), "nvl_allgather_from")
        _assert_local(arena, full, weight)
        if G.supported(full, weight, False, trans_b):
            return G.gemm(full, weight, False, trans_b, bias=bias, act=act)
        out = torch.matmul(full, weight.t() if trans_b else weight)
        if bias is not None:
            out = out + bias
        return {"gelu": torch.nn.functional.gelu, "relu": torch.relu}.get(act, lambda t: t)(out)
    st["calls"] += 1
    epoch = st["calls"]
    buf = epoch & 1
    full, flags = st["full"][buf], st["flags"][buf]
    cur = torch.cuda.current_stream()
    side = st["stream"]
    ready = torch.cuda.Event()
    ready.record(cur)                                  # x_shard is final; earlier readers of this buffer are enqueued
    chunk, per_rank = st["chunk"], st["per_rank"]
    chunk_bytes = chunk * K * 2
    # the shard has to sit in the arena for the copy engines to reach the peers' mappings from it: stage it once
    stage = _staging(arena, "ag_src", (rows, K), torch.bfloat16)
    with torch.cuda.stream(side):
        side.wait_event(ready)
        stage.copy_(xs, non_blocking=True)
        xs.record_stream(side)
        # chunk c of rank r lands at row r * rows + c * chunk of every rank's gathered buffer, then its flag is raised
        if arena._rt.pdb200_symm_push_signal(arena._h, stage._symm[1], full._symm[1] + arena.rank * rows * K * 2, chunk_bytes, per_rank,
                                             flags._symm[1], arena.rank * per_rank, epoch, side.cuda_stream) != 0:
            raise RuntimeError(f"symm_push_signal: {arena._rt.pdb200_rt_last_error().decode()}")
    out = torch.empty(rows * n, N, dtype=torch.bfloat16, device=xs.device)
    _assert_local(arena, full, weight, out)
    check(lib.kernels.pdb200_gemm_bf16_gated(ptr(full), ptr(weight), ptr(out), ptr(bias), rows * n, N, K, int(trans_b),
                                             {"none": 0, None: 0, "gelu": 1, "relu": 2}[act], ptr(flags), chunk, epoch,
                                             (arena.rank * rows) // 256, stream_ptr(xs.device)), "gemm_bf16_gated")
    done = torch.cuda.Event()
    done.record(side)
    cur.wait_event(done)          # the side stream's work of this call is ordered before whatever follows the GEMM
    return out


def _ce_chunk(arena, src_off, dst_off, nbytes, stream):
    """Copy ``nbytes`` from ``src_off`` of this rank's segment to ``dst_off`` of EVERY rank's segment (copy engines)."""
    seg = (ctypes.c_int64 * 3)(src_off, dst_off, nbytes)
    if arena._rt.pdb200_symm_bcast_ce(arena._h, ctypes.cast(seg, ctypes.c_void_p), 1, stream) != 0:
        raise RuntimeError(f"symm_bcast_ce: {arena._rt.pdb200_rt_last_error().decode()}")


_rs_state = _ShapeCache()
_caches.append(_rs_state)


def _rs_buffers(arena, rows, N, n):
    key = (id(arena), rows, N)
    st = _rs_state.lookup(key)
    if st is None:
        st = dict(part=[arena.empty((rows * n, N), torch.bfloat16) for _ in range(2)],
                  recv=[arena.empty((n, rows, N), torch.bfloat16) for _ in range(2)],
                  counters=[arena.empty((max(n, 8),), torch.int32) for _ in range(2)],
                  flags=[arena.empty((max(n, 8),), torch.int32) for _ in range(2)],
                  stream=torch.cuda.Stream(priority=-1), calls=0, uses=[0, 0])
        for t in st["counters"] + st["flags"]:
            t.zero_()
        torch.cuda.current_stream().synchronize()
        C.barrier(arena.group)
        _rs_state[key] = st
    return st


def gemm_rs(x, weight, group, trans_b=False, _out=None):
    """K6: ``reduce_scatter(x @ weight, dim 0)`` with the transfer hidden behind the GEMM.

    The CTA-pair GEMM works through the owners' row blocks one after the other, its own block last, and counts every
    finished piece into a per-block counter.  A side stream waits for a block's counter, ships the block to its owner
    with the copy engines and raises the owner's arrival flag - while the GEMM is still computing the later blocks.
    When the GEMM ends only the last remote block is still in flight; a local kernel then sums the n partial blocks
    (its own and the n-1 received ones).  Falls back to GEMM + one NVLS pull-reduce kernel for shapes that cannot be
    blocked.  x: [nranks*rows, K] bf16; returns [rows, N] bf16."""
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    arena = get_arena(group)
    M, K = x.shape
    n = arena.nranks
    rows = M // n
    N = weight.shape[0] if trans_b else weight.shape[1]
    overlapped = (flag("gemm_rs_overlap") and (_out is not None or _worth_overlapping(M, N, K)) and rows % 256 == 0 and K % 8 == 0 and N % 8 == 0 and weight.is_contiguous()
                  and hasattr(lib.kernels.require(), "pdb200_gemm_bf16_counted") and hasattr(arena._rt, "pdb200_symm_ship_blocks")
                  and arena._rt.pdb200_symm_streamops_supported() == 1)
    if not overlapped:
        part = _gemm_into_symmetric(arena, x, weight, trans_b, "rs_part")
        out = _out if _out is not None else torch.empty(rows, N, dtype=torch.bfloat16, device=x.device)
        check(lib.kernels.pdb200_nvl_rs_sum_bf16(arena.peers_arg, arena.mc_ptr, n, arena.rank, part._symm[1],
                                                 arena.rank * rows * N, rows * N, ptr(out), 0, arena.next_epoch(),
                                                 stream_ptr(x.device)), "nvl_rs_sum_bf16")
        return out
    st = _rs_buffers(arena, rows, N, n)
    st["calls"] += 1
    epoch = st["calls"]
    b = epoch & 1
    st["uses"][b] += 1
    part, recv, counters, flags = st["part"][b], st["recv"][b], st["counters"][b], st["flags"][b]
    per_block = (rows // 256) * ((N + 255) // 256) * 8           # counter growth per finished block
    target = st["uses"][b] * per_block
    cur, side = torch.cuda.current_stream(), st["stream"]
    xc = x.contiguous()
    _assert_local(arena, xc, weight, part)
    start = torch.cuda.Event()
    start.record(cur)                                             # everything before this call is enqueued ahead
    r = arena.rank
    check(lib.kernels.pdb200_gemm_bf16_counted(ptr(xc), ptr(weight), ptr(part), M, N, K, int(trans_b), ptr(counters), rows,
                                               (((r + 1) % n) * rows) // 256, stream_ptr(x.device)), "gemm_bf16_counted")
    blk = rows * N * 2
    with torch.cuda.stream(side):
        side.wait_event(start)
        if arena._rt.pdb200_symm_ship_blocks(arena._h, counters._symm[1], target, part._symm[1], recv._symm[1], blk, flags._symm[1],
                                             epoch, side.cuda_stream) != 0:
            raise RuntimeError(f"symm_ship_blocks: {arena._rt.pdb200_rt_last_error().decode()}")
        sent = torch.cuda.Event()
        sent.record(side)
    # own block: after the GEMM (stream order) copy it next to the received ones, wait for the n-1 arrivals, sum
    recv[r].copy_(part[r * rows:(r + 1) * rows], non_blocking=True)
    if arena._rt.pdb200_symm_wait_flags(arena._h, flags._symm[1], epoch, r, stream_ptr(x.device)) != 0:
        raise RuntimeError(f"symm_wait_flags: {arena._rt.pdb200_rt_last_error().decode()}")
    out = _out if _out is not None else torch.empty(rows, N, dtype=torch.bfloat16, device=x.device)
    tri = (ctypes.c_int64 * 3)(recv.data_ptr(), rows * N, out.data_ptr())
    check(lib.kernels.pdb200_reduce_slots_multi2(ctypes.cast(tri, ctypes.c_void_p), 1, n, 1, 1.0, 0, 1, 0, stream_ptr(x.device)),
          "reduce_slots")
    cur.wait_event(sent)          # this rank's puts are done before part[b] can be written again (two calls from now)
    return out


def _gemm_into_symmetric(arena, x, weight, trans_b, tag):
    """x @ weight written straight into a (local) symmetric staging buffer by the tcgen05 GEMM (no copy)."""
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    M, K = x.shape
    N = weight.shape[0] if trans_b else weight.shape[1]
    part = _staging(arena, tag, (M, N), torch.bfloat16)
    xc = x.contiguous()
    _assert_local(arena, xc, weight, part)
    fn = lib.kernels.pdb200_gemm_bf16_pair if (M >= 512 and N >= 512 and flag("gemm_cta_pair")) \
        else lib.kernels.pdb200_gemm_bf16
    check(fn(ptr(xc), ptr(weight), ptr(part), None, M, N, K, 0, int(trans_b), 1, 0.0, 0, stream_ptr(x.device)), "gemm_bf16")
    return part


_ar_state = _ShapeCache()
_caches.append(_ar_state)


def gemm_ar(x, weight, group, trans_b=False):
    """K4: ``all_reduce(x @ weight)`` for the non-sequence-parallel TP linears = the overlapped GEMM -> reduce-scatter of
    ``gemm_rs`` (row blocks leave for their owners while the GEMM still runs) followed by a copy-engine all-gather of the
    reduced blocks, each announced by a flag.  x: [M, K] bf16, M divisible by nranks*256 for the overlapped path
    (otherwise: GEMM into a symmetric buffer + the NVLS two-shot all-reduce kernel).  Returns a private [M, N] tensor."""
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    arena = get_arena(group)
    M, K = x.shape
    n = arena.nranks
    N = weight.shape[0] if trans_b else weight.shape[1]
    rows = M // n
    if not (flag("gemm_rs_overlap") and _worth_overlapping(M, N, K) and M % n == 0 and rows % 256 == 0 and hasattr(arena._rt, "pdb200_symm_push_signal")
            and arena._rt.pdb200_symm_streamops_supported() == 1 and hasattr(lib.kernels.require(), "pdb200_gemm_bf16_counted")):
        part = _gemm_into_symmetric(arena, x, weight, trans_b, "ar_part")
        allreduce_(part)
        return part.clone()
    key = (id(arena), M, N)
    st = _ar_state.lookup(key)
    if st is None:
        st = dict(red=[arena.empty((rows, N), torch.bfloat16) for _ in range(2)],
                  full=[arena.empty((M, N), torch.bfloat16) for _ in range(2)],
                  flags=[arena.empty((max(n, 8),), torch.int32) for _ in range(2)], calls=0)
        for t in st["flags"]:
            t.zero_()
        torch.cuda.current_stream().synchronize()
        C.barrier(arena.group)
        _ar_state[key] = st
    st["calls"] += 1
    epoch = st["calls"]
    b = epoch & 1
    red, full, flags = st["red"][b], st["full"][b], st["flags"][b]
    gemm_rs(x, weight, group, trans_b, _out=red)                 # my reduced block, in the arena
    stream = stream_ptr(x.device)
    blk = rows * N * 2
    if arena._rt.pdb200_symm_push_signal(arena._h, red._symm[1], full._symm[1] + arena.rank * blk, blk, 1, flags._symm[1], arena.rank,
                                         epoch, stream) != 0 or \
            arena._rt.pdb200_symm_wait_flags(arena._h, flags._symm[1], epoch, -1, stream) != 0:
        raise RuntimeError(f"gemm_ar: {arena._rt.pdb200_rt_last_error().decode()}")
    return full.clone()


def vocab_embed(ids, weight, where):
    """K7: vocab-parallel embedding forward as a gather from the owner's shard; ``where`` = (arena, offset) of the
    symmetric storage of ``weight`` (see make_symmetric)."""
    from paddle_b200.lib import check
    from paddle_b200.ops._util import ptr
    arena, off = where
    ids64 = ids.reshape(-1).to(torch.int64).contiguous()
    out = torch.empty(ids64.numel(), weight.shape[1], dtype=weight.dtype, device=weight.device)
    check(lib.kernels.pdb200_nvl_vocab_embed(arena.peers_arg, arena.nranks, arena.rank, off, ptr(ids64), ptr(out),
                                             ids64.numel(), weight.shape[1], weight.shape[0], dtype_code(weight), 0,
                                             arena.next_epoch(), stream_ptr(weight.device)), "nvl_vocab_embed")
    return out.reshape(*ids.shape, weight.shape[1])


def make_symmetric(param, group):
    """Move a parameter's storage into the group's symmetric arena (once); returns False when that is not possible
    (views into flat buffers, no arena space, fused paths off)."""
    if getattr(param, "_symm", None) is not None and getattr(param, "_symm_ptr", None) == param.data_ptr():
        return True
    if not available(group) or param.data._is_view():
        return False
    try:
        arena = get_arena(group)
        if arena.free_bytes() < param.numel() * param.element_size() + (1 << 20):
            return False
        t = arena.empty(tuple(param.shape), param.dtype)
    except Exception:  # noqa: BLE001
        return False
    with torch.no_grad():
        t.copy_(param.data)
        param.data = t
    param._symm = t._symm
    param._symm_ptr = t.data_ptr()
    return True


def fused_gemm_ok(group, rows_per_rank, K, N, dtype, weight=None):
    if not (dtype == torch.bfloat16 and rows_per_rank % 128 == 0 and K % 64 == 0 and N % 64 == 0 and available(group)):
        return False
    if not hasattr(lib.kernels.require(), "pdb200_nvl_allgather_from"):
        return False
    try:
        arena = get_arena(group)
    except Exception:  # noqa: BLE001
        return False
    need = max(rows_per_rank * K, group.nranks * rows_per_rank * N) * 2
    return arena.free_bytes() > need or any(k[0] == id(arena) for k in _stage)


# --------------------------------------------------------------------------------------------- NCCL-shaped helpers
def all_gather_small(t, group):
    """[...]-> [nranks, ...] for small payloads (NCCL baseline of the CE statistics exchange)."""
    n = group.nranks
    out = torch.empty((n,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    comm.all_gather(out, t.contiguous(), group)
    return out


def all_gather_seq(x, group):
    """[s/n, ...] -> [s, ...] along dim 0 (sequence-parallel all-gather)."""
    n = group.nranks
    out = torch.empty((x.shape[0] * n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    comm.all_gather(out, x.contiguous(), group)
    return out
