"""global_scatter / global_gather (reference: python/paddle/distributed/utils/moe_utils.py:20-260,
paddle/fluid/operators/collective/global_scatter_op.cu.cc:33-200, global_gather_op.cu.cc).

MoE token all-to-all with per-(expert, rank) counts.  ``local_count[i + j*E]`` = rows this rank sends to expert i of
rank j; ``global_count[i + j*E]`` = rows this rank receives from rank j for its local expert i.  Send buffer order is
(rank-major, expert-minor) = the order produced by ``assign_pos``; the receive buffer is expert-major
(Appendix B.10).  The reference copies both count tensors to the host and issues E x W grouped ncclSend/Recv; here
the counts are exchanged once as splits of a single all_to_all_single, and the fused NVLink path
(``parallel.symm_fused.moe_dispatch``) pushes rows straight into the destination expert's symmetric slab.
Both directions are differentiable (scatter backward = gather and vice versa)."""
from __future__ import annotations

import torch

from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm


def _split_sizes(count, n_expert, world):
    c = count.reshape(world, n_expert)
    return c.sum(1).tolist(), c


def _scatter_impl(x, local_count, global_count, group):
    g = group if group is not None else C._get_default_group()
    world = g.nranks
    n_expert = local_count.numel() // world
    lc, gc = local_count.long().cpu(), global_count.long().cpu()
    in_splits, _ = _split_sizes(lc, n_expert, world)
    out_splits, gcm = _split_sizes(gc, n_expert, world)
    recv = torch.empty((sum(out_splits),) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    if world == 1:
        recv.copy_(x[: recv.shape[0]])
    else:
        comm.alltoall_single(recv, x.contiguous(), in_splits, out_splits, g)
    # received layout: for rank j: for expert i: rows.  Reorder to expert-major (expert i: for rank j).
    if world > 1 and n_expert > 1:
        offs = torch.zeros(world, n_expert, dtype=torch.long)
        flat = gcm.reshape(-1)
        starts = torch.cumsum(flat, 0) - flat
        offs = starts.reshape(world, n_expert)
        pieces = []
        for i in range(n_expert):
            for j in range(world):
                s, n = int(offs[j, i]), int(gcm[j, i])
                if n:
                    pieces.append(recv[s:s + n])
        recv = torch.cat(pieces, 0) if pieces else recv[:0]
    return recv


def _gather_impl(x, local_count, global_count, group):
    g = group if group is not None else C._get_default_group()
    world = g.nranks
    n_expert = local_count.numel() // world
    lc, gc = local_count.long().cpu(), global_count.long().cpu()
    out_splits, _ = _split_sizes(lc, n_expert, world)
    in_splits, gcm = _split_sizes(gc, n_expert, world)
    # x is expert-major (expert i: for rank j); send order must be rank-major
    if world > 1 and n_expert > 1:
        sizes = gcm.t().reshape(-1)                 # [expert, rank] flattened
        starts = torch.cumsum(sizes, 0) - sizes
        starts = starts.reshape(n_expert, world)
        pieces = []
        for j in range(world):
            for i in range(n_expert):
                s, n = int(starts[i, j]), int(gcm[j, i])
                if n:
                    pieces.append(x[s:s + n])
        x = torch.cat(pieces, 0) if pieces else x[:0]
    out = torch.empty((sum(out_splits),) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    if world == 1:
        out.copy_(x[: out.shape[0]])
    else:
        comm.alltoall_single(out, x.contiguous(), in_splits, out_splits, g)
    return out


class _GlobalScatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, local_count, global_count, group):
        ctx.cfg = (local_count, global_count, group)
        return _scatter_impl(x, local_count, global_count, group)

    @staticmethod
    def backward(ctx, dy):
        lc, gc, group = ctx.cfg
        return _gather_impl(dy.contiguous(), lc, gc, group), None, None, None


class _GlobalGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, local_count, global_count, group):
        ctx.cfg = (local_count, global_count, group)
        return _gather_impl(x, local_count, global_count, group)

    @staticmethod
    def backward(ctx, dy):
        lc, gc, group = ctx.cfg
        return _scatter_impl(dy.contiguous(), lc, gc, group), None, None, None


def global_scatter(x, local_count, global_count, group=None, use_calc_stream=True):
    if group is not None and not group.is_member():
        return
    return _GlobalScatter.apply(x, local_count, global_count, group)


def global_gather(x, local_count, global_count, group=None, use_calc_stream=True):
    if group is not None and not group.is_member():
        return
    return _GlobalGather.apply(x, local_count, global_count, group)
