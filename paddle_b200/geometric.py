"""paddle.geometric (reference: python/paddle/geometric/{math,message_passing/send_recv}.py): segment reductions and
graph message passing, as index_reduce / scatter operations."""
from __future__ import annotations

import torch


def _segment(data, segment_ids, reduce):
    ids = segment_ids.long()
    n = int(ids.max().item()) + 1 if ids.numel() else 0
    out = torch.zeros((n,) + tuple(data.shape[1:]), dtype=data.dtype, device=data.device)
    if n == 0:
        return out
    idx = ids.reshape((-1,) + (1,) * (data.dim() - 1)).expand_as(data)
    if reduce == "sum":
        return out.scatter_add_(0, idx, data)
    if reduce == "mean":
        return out.scatter_reduce_(0, idx, data, "mean", include_self=False)
    return out.scatter_reduce_(0, idx, data, "amax" if reduce == "max" else "amin", include_self=False)


def segment_sum(data, segment_ids, name=None):
    return _segment(data, segment_ids, "sum")


def segment_mean(data, segment_ids, name=None):
    return _segment(data, segment_ids, "mean")


def segment_max(data, segment_ids, name=None):
    return _segment(data, segment_ids, "max")


def segment_min(data, segment_ids, name=None):
    return _segment(data, segment_ids, "min")


def _reduce_to(msg, dst_index, reduce_op, out_size, like):
    n = int(out_size) if out_size is not None else like.shape[0]
    out = torch.zeros((n,) + tuple(msg.shape[1:]), dtype=msg.dtype, device=msg.device)
    idx = dst_index.long().reshape((-1,) + (1,) * (msg.dim() - 1)).expand_as(msg)
    if reduce_op == "sum":
        return out.scatter_add_(0, idx, msg)
    return out.scatter_reduce_(0, idx, msg, {"mean": "mean", "max": "amax", "min": "amin"}[reduce_op], include_self=False)


def send_u_recv(x, src_index, dst_index, reduce_op="sum", out_size=None, name=None):
    """Gather node features along edges and reduce them at the destination nodes."""
    return _reduce_to(x[src_index.long()], dst_index, reduce_op, out_size, x)


def send_ue_recv(x, y, src_index, dst_index, message_op="add", reduce_op="sum", out_size=None, name=None):
    m = x[src_index.long()]
    m = {"add": m + y, "sub": m - y, "mul": m * y, "div": m / y}[message_op]
    return _reduce_to(m, dst_index, reduce_op, out_size, x)


def send_uv(x, y, src_index, dst_index, message_op="add", name=None):
    a, b = x[src_index.long()], y[dst_index.long()]
    return {"add": a + b, "sub": a - b, "mul": a * b, "div": a / b}[message_op]


# ------------------------------------------------------------------------------------- graph re-indexing and sampling
def _reindex(x, neighbor_lists, count_lists):
    """Shared numbering: the input nodes keep positions 0..n-1, every new neighbour gets the next free id in order of
    first appearance (scanning the edge types one after another)."""
    import numpy as np
    xs = x.cpu().numpy().tolist()
    table = {}
    for node in xs:
        if node in table:
            raise ValueError("the input nodes x should be unique")
        table[node] = len(table)
    srcs, dsts = [], []
    for neighbors, count in zip(neighbor_lists, count_lists):
        nb = neighbors.cpu().numpy().tolist()
        cnt = count.cpu().numpy().astype(np.int64)
        if int(cnt.sum()) != len(nb) or len(cnt) != len(xs):
            raise ValueError("count should have one entry per input node and sum to the number of neighbors")
        src = np.empty(len(nb), dtype=np.int64)
        for i, node in enumerate(nb):
            idx = table.get(node)
            if idx is None:
                idx = table[node] = len(table)
            src[i] = idx
        srcs.append(src)
        dsts.append(np.repeat(np.arange(len(xs), dtype=np.int64), cnt))
    out_nodes = np.fromiter(table.keys(), dtype=np.int64, count=len(table))
    dev, dt = x.device, x.dtype
    return (torch.from_numpy(np.concatenate(srcs)).to(dev, dt), torch.from_numpy(np.concatenate(dsts)).to(dev, dt),
            torch.from_numpy(out_nodes).to(dev, dt))


def reindex_graph(x, neighbors, count, value_buffer=None, index_buffer=None, name=None):
    """Renumber a sampled sub-graph from 0: returns (reindex_src, reindex_dst, out_nodes) where ``out_nodes`` lists the
    original ids (input nodes first), ``reindex_src[i]`` is the new id of ``neighbors[i]`` and ``reindex_dst`` repeats
    the new id of every input node ``count`` times (reference: python/paddle/geometric/reindex.py)."""
    return _reindex(x, [neighbors], [count])


def reindex_heter_graph(x, neighbors, count, value_buffer=None, index_buffer=None, name=None):
    """``reindex_graph`` for several edge types sharing one node numbering: ``neighbors`` / ``count`` are lists with one
    entry per edge type; the per-type results are concatenated."""
    if not isinstance(neighbors, (list, tuple)) or not isinstance(count, (list, tuple)) or len(neighbors) != len(count):
        raise ValueError("neighbors and count should be lists of the same length (one entry per edge type)")
    return _reindex(x, list(neighbors), list(count))


def _sample(row, colptr, input_nodes, sample_size, eids, return_eids, weights):
    import numpy as np
    if return_eids and eids is None:
        raise ValueError("`eids` should not be None if `return_eids` is True.")
    rows, ptr = row.cpu().numpy(), colptr.cpu().numpy()
    nodes = input_nodes.cpu().numpy()
    eid = eids.cpu().numpy() if eids is not None else None
    w = weights.cpu().numpy().astype(np.float64) if weights is not None else None
    seed = int(torch.randint(0, 2 ** 31 - 1, (1,)).item())                   # follows paddle.seed / torch.manual_seed
    rng = np.random.default_rng(seed)
    out_n, out_c, out_e = [], [], []
    for node in nodes.tolist():
        lo, hi = int(ptr[node]), int(ptr[node + 1])
        deg = hi - lo
        if sample_size < 0 or deg <= sample_size:
            pick = np.arange(lo, hi)
        elif w is None:
            pick = lo + rng.choice(deg, size=sample_size, replace=False)
        else:
            # weighted sampling without replacement (Efraimidis-Spirakis): keep the largest u^(1/w) keys
            keys = rng.random(deg) ** (1.0 / np.maximum(w[lo:hi], 1e-300))
            pick = lo + np.argpartition(-keys, sample_size - 1)[:sample_size]
        out_n.append(rows[pick])
        out_c.append(len(pick))
        if return_eids:
            out_e.append(eid[pick])
    dev = row.device
    neighbors = torch.from_numpy(np.concatenate(out_n) if out_n else np.zeros(0, dtype=rows.dtype)).to(dev)
    counts = torch.as_tensor(out_c, dtype=torch.int32, device=dev)
    if return_eids:
        return neighbors, counts, torch.from_numpy(np.concatenate(out_e) if out_e else np.zeros(0, dtype=eid.dtype)).to(dev)
    return neighbors, counts


def sample_neighbors(row, colptr, input_nodes, sample_size=-1, eids=None, return_eids=False, perm_buffer=None, name=None):
    """Uniformly sample up to ``sample_size`` in-neighbours (without replacement; all of them for -1 or small degrees)
    of every input node of a CSC graph (``row[colptr[v]:colptr[v + 1]]`` are v's neighbours).  Returns
    (out_neighbors, out_count[, out_eids])."""
    return _sample(row, colptr, input_nodes, sample_size, eids, return_eids, None)


def weighted_sample_neighbors(row, colptr, edge_weight, input_nodes, sample_size=-1, eids=None, return_eids=False, name=None):
    """``sample_neighbors`` with inclusion probabilities proportional to ``edge_weight`` (weighted reservoir sampling)."""
    return _sample(row, colptr, input_nodes, sample_size, eids, return_eids, edge_weight)
