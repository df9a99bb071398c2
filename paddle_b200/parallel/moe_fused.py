"""K9 / K10: MoE dispatch and combine over NVLink symmetric memory (csrc/kernels/moe_nvl.cu).

Replaces, for one NVSwitch node, the reference's count all-to-all + host-synchronous grouped send/recv
(python/paddle/distributed/utils/moe_utils.py:20-260, global_scatter_op.cu.cc:55-75) and the local
index_select / scatter around them (incubate/distributed/models/moe/moe_layer.py:441-496):

* ``dispatch(x, pos, lec)``: rows ``x[pos // topk]`` are stored straight into the destination ranks' expert-major
  slabs; routing offsets come from the counts every rank publishes in symmetric memory (no NCCL, no host trip).
  Returns this rank's received rows ``[R, d]`` and the per-expert row counts (ONE host read: the expert GEMMs need
  their shapes - the reference reads the same numbers).
* ``combine(y, meta)``: every (token, k) entry pulls its expert-output row from the rank that computed it
  -> ``[T * topk, d]`` in token order (zeros for dropped entries), ready for the gate-weighted sum.
Both are differentiable: dispatch.backward is a summing pull of the gradient slab, combine.backward a dispatch.
"""
from __future__ import annotations

import os

import torch

from paddle_b200 import lib
from paddle_b200.lib import check
from paddle_b200.ops._util import dtype_code, ptr, stream_ptr

from . import symm

_state = {}


class _Meta:
    """Per-layer routing record shared by dispatch and combine (and their backward passes)."""


def available(group, x):
    if group is None or group.nranks < 2 or not x.is_cuda or not symm.available(group):
        return False
    if not hasattr(lib.kernels.require(), "pdb200_moe_dispatch"):
        return False
    return x.dtype in (torch.bfloat16, torch.float32) and (x.shape[-1] * x.element_size()) % 16 == 0


def _buffers(group, rows, row_bytes, nb):
    """Dedicated arena for the MoE slabs of this group, grown when a layer needs more."""
    key = group.id
    st = _state.get(key)
    need = rows * row_bytes
    if st is None or st["slab_bytes"] < need or st["nb"] < nb:
        if st is not None:
            st["arena"].close()
        arena = symm.SymmArena(group, need + 4 * nb + (4 << 20))
        counts = arena.empty((max(nb, 64),), torch.int32)
        slab = arena.empty((need,), torch.uint8)
        st = _state[key] = dict(arena=arena, counts=counts, slab=slab, slab_bytes=need, nb=nb)
    return st


def release_all():
    for st in _state.values():
        st["arena"].close()
    _state.clear()


def _slab_rows(group, n_slots):
    # worst case is nranks * n_slots (everything lands on one rank); a factor of the balanced load is provisioned and
    # checked against the real counts after the dispatch
    f = float(os.environ.get("PDB200_MOE_SLAB_FACTOR", "2.0"))
    return int(max(n_slots * min(f, group.nranks), 64))


def _push(meta, src, div):
    a = meta.arena
    overflow = torch.zeros(1, dtype=torch.int32, device=src.device)
    slab_rows = meta.slab.numel() // meta.row_bytes
    check(lib.kernels.pdb200_moe_dispatch(a.peers_arg, a.nranks, a.rank, meta.counts._symm[1], meta.slab._symm[1], slab_rows,
                                          ptr(src), ptr(meta.pos), int(div), ptr(meta.lec), meta.E, meta.pos.numel(),
                                          meta.row_bytes, ptr(meta.route_rank), ptr(meta.route_row), ptr(meta.recv_counts),
                                          ptr(overflow), 0, a.next_epoch(), stream_ptr(src.device)), "moe_dispatch")
    return overflow


def _pull(meta, out, sum_over_k):
    a = meta.arena
    check(lib.kernels.pdb200_moe_pull(a.peers_arg, a.nranks, a.rank, meta.slab._symm[1], ptr(meta.route_rank),
                                      ptr(meta.route_row), ptr(meta.inv), meta.T * meta.topk, ptr(out), meta.row_bytes,
                                      int(sum_over_k), dtype_code(out), 0, a.next_epoch(), stream_ptr(out.device)),
          "moe_pull")
    return out


def _slab_view(meta, rows):
    return meta.slab[: rows * meta.row_bytes].view(meta.dtype).view(rows, meta.d)


class _Dispatch(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, meta):
        meta.overflow = _push(meta, x.contiguous(), meta.topk)
        slab_rows = meta.slab.numel() // meta.row_bytes
        if not meta.sync:
            # grouped-GEMM consumers take the slab at full capacity plus the device-side counts: no host read.  Rows
            # that did not fit were dropped by the kernel and flagged in meta.overflow (checked by check_overflow()).
            meta.R = slab_rows
            meta.counts_host = None
            ctx.meta = meta
            return _slab_view(meta, meta.R).clone()
        rc = meta.recv_counts.tolist()                      # the one host read of the layer (expert GEMM shapes)
        meta.R = int(rc[-1])
        meta.counts_host = rc[:-1]
        if meta.R > slab_rows:
            raise RuntimeError(f"MoE receive slab too small: {meta.R} rows routed to this rank, {slab_rows} provisioned; "
                               f"raise PDB200_MOE_SLAB_FACTOR (now {os.environ.get('PDB200_MOE_SLAB_FACTOR', '2.0')})")
        ctx.meta = meta
        return _slab_view(meta, meta.R).clone()

    @staticmethod
    def backward(ctx, dslab):
        meta = ctx.meta
        _slab_view(meta, meta.R).copy_(dslab)
        dx = torch.empty(meta.T, meta.d, dtype=meta.dtype, device=dslab.device)
        _pull(meta, dx, meta.topk)                          # dx[t] = sum_k dslab[slot(t, k)]
        return dx, None


class _Combine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, meta):
        _slab_view(meta, meta.R).copy_(y)
        out = torch.empty(meta.T * meta.topk, meta.d, dtype=meta.dtype, device=y.device)
        _pull(meta, out, 0)
        ctx.meta = meta
        return out

    @staticmethod
    def backward(ctx, dout):
        meta = ctx.meta
        _push(meta, dout.contiguous(), 1)                   # entry e = pos[s] itself
        return _slab_view(meta, meta.R).clone(), None


def dispatch(x, pos, lec, group, num_expert, topk, sync=True):
    """x: [T, d]; pos: [S] int64 (sorted slot -> token*topk + k); lec: [W*E] int64 send counts.  Returns
    (rows [R, d] expert-major, per-expert counts, meta for ``combine``).  ``sync=True``: R is the exact row count and
    the counts are a python list (one host read).  ``sync=False``: R is the slab capacity, the counts stay on the
    device (int64[num_expert]) and nothing is read back - for consumers that schedule on the device."""
    T, d = x.shape
    meta = _Meta()
    meta.sync = bool(sync)
    nb = group.nranks * num_expert
    meta.row_bytes = d * x.element_size()
    st = _buffers(group, _slab_rows(group, max(T * topk, 1)), meta.row_bytes, nb)
    meta.arena, meta.slab, meta.counts = st["arena"], st["slab"], st["counts"]
    meta.pos, meta.lec = pos.contiguous(), lec.to(torch.int64).contiguous()
    meta.E, meta.topk, meta.T, meta.d, meta.dtype = num_expert, topk, T, d, x.dtype
    S = pos.numel()
    # inverse map entry -> sorted slot; padded position lists carry -1 past the routed entries: those go to a dummy cell
    inv = torch.full((T * topk + 1,), -1, dtype=torch.int64, device=x.device)
    if S:
        safe = torch.where(pos >= 0, pos, torch.full_like(pos, T * topk))
        inv.scatter_(0, safe, torch.arange(S, dtype=torch.int64, device=x.device))
    meta.inv = inv[: T * topk].contiguous()
    meta.route_rank = torch.empty(max(S, 1), dtype=torch.int32, device=x.device)
    meta.route_row = torch.empty(max(S, 1), dtype=torch.int32, device=x.device)
    meta.recv_counts = torch.zeros(num_expert + 1, dtype=torch.int64, device=x.device)
    rows = _Dispatch.apply(x, meta)
    if not sync:
        return rows, meta.recv_counts[:num_expert], meta
    return rows, meta.counts_host, meta


def check_overflow(meta):
    """Host-side check (synchronises) that no routed row was dropped for lack of slab space; optional in the no-sync mode."""
    if int(meta.overflow.item()) != 0:
        raise RuntimeError("MoE receive slab too small: raise PDB200_MOE_SLAB_FACTOR "
                           f"(now {os.environ.get('PDB200_MOE_SLAB_FACTOR', '2.0')})")


def combine(y, meta):
    return _Combine.apply(y, meta)
