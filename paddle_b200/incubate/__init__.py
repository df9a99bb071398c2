"""paddle.incubate (reference: python/paddle/incubate/__init__.py)."""
import torch as _torch

from . import asp, autograd, distributed, nn, optimizer  # noqa: F401
from .optimizer import LookAhead, ModelAverage  # noqa: F401
from ..geometric import segment_max, segment_mean, segment_min, segment_sum  # noqa: F401
from ..geometric import send_u_recv as _send_u_recv


def softmax_mask_fuse(x, mask, name=None):
    """softmax(x + mask) over the last axis (reference: incubate/operators/softmax_mask_fuse.py)."""
    return _torch.softmax(x.float() + mask.float(), -1).to(x.dtype)


def softmax_mask_fuse_upper_triangle(x):
    """Causal softmax: the strict upper triangle of the last two axes is masked out."""
    s = x.shape[-1]
    m = _torch.ones(x.shape[-2], s, dtype=_torch.bool, device=x.device).tril()
    return _torch.softmax(x.float().masked_fill(~m, float("-inf")), -1).to(x.dtype)


def graph_send_recv(x, src_index, dst_index, pool_type="sum", out_size=None, name=None):
    return _send_u_recv(x, src_index, dst_index, pool_type, out_size)


def graph_reindex(x, neighbors, count, value_buffer=None, index_buffer=None, flag_buffer_hashtable=False, name=None):
    """Renumber the nodes of a sampled sub-graph: input nodes first, then new neighbours in order of appearance."""
    order = {}
    for v in x.tolist() + neighbors.tolist():
        order.setdefault(v, len(order))
    src = _torch.tensor([order[v] for v in neighbors.tolist()], dtype=x.dtype)
    dst = _torch.repeat_interleave(_torch.arange(x.numel(), dtype=x.dtype), count.long())
    return src, dst, _torch.tensor(list(order), dtype=x.dtype)


def graph_sample_neighbors(row, colptr, input_nodes, eids=None, perm_buffer=None, sample_size=-1, return_eids=False,
                           flag_perm_buffer=False, name=None):
    """Uniformly sample up to ``sample_size`` in-neighbours of every input node from a CSC graph."""
    outs, counts, oe = [], [], []
    for n in input_nodes.tolist():
        lo, hi = int(colptr[n]), int(colptr[n + 1])
        idx = _torch.arange(lo, hi)
        if 0 <= sample_size < idx.numel():
            idx = idx[_torch.randperm(idx.numel())[:sample_size]]
        outs.append(row[idx])
        counts.append(idx.numel())
        if return_eids and eids is not None:
            oe.append(eids[idx])
    nb = _torch.cat(outs) if outs else row[:0]
    cnt = _torch.tensor(counts, dtype=_torch.int32)
    return (nb, cnt, _torch.cat(oe)) if return_eids and eids is not None else (nb, cnt)


def graph_khop_sampler(row, colptr, input_nodes, sample_sizes, sorted_eids=None, return_eids=False, name=None):
    frontier, all_src, all_dst = input_nodes, [], []
    for k in sample_sizes:
        nb, cnt = graph_sample_neighbors(row, colptr, frontier, sample_size=k)
        all_src.append(nb)
        all_dst.append(_torch.repeat_interleave(frontier, cnt.long()))
        frontier = _torch.unique(nb)
    src, dst = _torch.cat(all_src), _torch.cat(all_dst)
    order = {}
    for v in input_nodes.tolist() + src.tolist():
        order.setdefault(v, len(order))
    re = lambda t: _torch.tensor([order[v] for v in t.tolist()], dtype=row.dtype)  # noqa: E731
    return re(src), re(dst), _torch.tensor(list(order), dtype=row.dtype), re(input_nodes)


def identity_loss(x, reduction="none"):
    """Marks ``x`` as a loss (IPU concept); reduces it as asked."""
    r = {0: "sum", 1: "mean", 2: "none"}.get(reduction, reduction)
    return x.sum() if r == "sum" else x.mean() if r == "mean" else x
