"""MoE routing ops (K11).  Reference API: python/paddle/distributed/models/moe/utils.py:20-230
(_number_count, _assign_pos, _random_routing, _limit_by_capacity, _prune_gate_by_capacity)."""
from __future__ import annotations

import torch

from ._util import K, check, dtype_code, ptr, stream_ptr, use_native


def _cuda_i64(t):
    return t.is_cuda and t.dtype == torch.int64


def number_count(numbers, upper_range):
    nums = numbers.reshape(-1).long().contiguous()
    if nums.is_cuda and use_native(torch.empty(0, device=nums.device)):
        out = torch.empty(upper_range, dtype=torch.int64, device=nums.device)
        check(K().pdb200_number_count(ptr(nums), nums.numel(), ptr(out), int(upper_range), stream_ptr(nums.device)),
              "number_count")
        return out
    valid = nums[(nums >= 0) & (nums < upper_range)]
    return torch.bincount(valid, minlength=upper_range).to(torch.int64)


def assign_pos(x, cum_count, padded=False):
    """Positions of tokens grouped by expert (stable).  cum_count = inclusive cumsum of per-expert counts.
    ``padded=True``: no host read of the total - the result has one entry per input id and the entries past the number
    of routed ids are -1."""
    ids = x.reshape(-1).long().contiguous()
    cum = cum_count.reshape(-1).long().contiguous()
    if padded:
        total = ids.numel()
    else:
        total = int(cum[-1].item()) if cum.numel() else 0
    if ids.is_cuda and use_native(torch.empty(0, device=ids.device)):
        out = torch.full((total,), -1, dtype=torch.int64, device=ids.device) if padded else \
            torch.empty(total, dtype=torch.int64, device=ids.device)
        check(K().pdb200_assign_pos(ptr(ids), ids.numel(), ptr(cum), cum.numel(), ptr(out), stream_ptr(ids.device)),
              "assign_pos")
        return out
    order = torch.argsort(torch.where(ids >= 0, ids, torch.full_like(ids, cum.numel())), stable=True)
    if padded:
        n_valid = (ids >= 0).sum()
        return torch.where(torch.arange(total, device=ids.device) < n_valid, order, torch.full_like(order, -1))
    return order[:total]


def limit_by_capacity(expert_count, capacity, n_worker):
    ec = expert_count.reshape(-1).long().contiguous()
    cap = capacity.reshape(-1).long().contiguous()
    n_expert = ec.numel() // n_worker
    if ec.is_cuda and use_native(torch.empty(0, device=ec.device)):
        out = torch.empty_like(ec)
        check(K().pdb200_limit_by_capacity(ptr(ec), ptr(cap), ptr(out), n_expert, int(n_worker), stream_ptr(ec.device)),
              "limit_by_capacity")
        return out
    ecw = ec.reshape(n_worker, n_expert)
    before = torch.cumsum(ecw, 0) - ecw
    left = (cap[None, :] - before).clamp(min=0)
    return torch.minimum(ecw, left).reshape(-1)


def prune_gate_by_capacity(gate_idx, expert_count, n_expert, n_worker):
    g = gate_idx.reshape(-1).long().contiguous()
    ec = expert_count.reshape(-1).long().contiguous()
    if g.is_cuda and use_native(torch.empty(0, device=g.device)):
        out = torch.empty_like(g)
        work = torch.empty(ec.numel(), dtype=torch.int32, device=g.device)
        check(K().pdb200_prune_gate_by_capacity(ptr(g), ptr(out), ptr(ec), ptr(work), ec.numel(), g.numel(),
                                                stream_ptr(g.device)), "prune_gate_by_capacity")
        return out.reshape(gate_idx.shape)
    out = g.clone()
    # first-come-first-served in token order
    for e in range(ec.numel()):
        pos = (g == e).nonzero().flatten()
        out[pos[int(ec[e].item()):]] = -1
    return out.reshape(gate_idx.shape)


def random_routing(topk_idx, topk_value, prob, topk=2):
    if topk != 2:
        raise RuntimeError("only topk=2 is supported now")
    idx = topk_idx.long().contiguous().clone()
    if idx.is_cuda and use_native(topk_value):
        tv = topk_value.contiguous()
        pr = prob.to(tv.dtype).contiguous()
        check(K().pdb200_random_routing(ptr(idx), ptr(tv), ptr(pr), idx.shape[0], dtype_code(tv), stream_ptr(idx.device)),
              "random_routing")
        return idx
    drop = 2.0 * topk_value[:, 1].float() < prob.float()
    idx[:, 1] = torch.where(drop, torch.full_like(idx[:, 1], -1), idx[:, 1])
    return idx


def topk_softmax(logits, k, renorm=True, with_counts=True):
    """Fused router: softmax over experts + top-k + (optionally renormalised) weights + expert histogram."""
    E = logits.shape[-1]
    lg = logits.reshape(-1, E).contiguous()
    T = lg.shape[0]
    if use_native(lg) and E <= 256 and k <= 8:
        w = torch.empty(T, k, dtype=torch.float32, device=lg.device)
        idx = torch.empty(T, k, dtype=torch.int64, device=lg.device)
        cnt = torch.empty(E, dtype=torch.int64, device=lg.device) if with_counts else None
        check(K().pdb200_topk_softmax(ptr(lg), T, E, int(k), int(renorm), ptr(w), ptr(idx), ptr(cnt), dtype_code(lg),
                                      stream_ptr(lg.device)), "topk_softmax")
        return w, idx, cnt
    p = torch.softmax(lg.float(), -1)
    w, idx = torch.topk(p, k, dim=-1)
    if renorm:
        w = w / w.sum(-1, keepdim=True)
    cnt = torch.bincount(idx.reshape(-1), minlength=E) if with_counts else None
    return w, idx, cnt


# --------------------------------------------------------------------------------------------- grouped expert GEMMs
def _grouped_native(x, w):
    return (x.is_cuda and use_native(x) and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and w.dim() == 3
            and w.shape[0] <= 64 and w.shape[1] % 64 == 0 and w.shape[2] % 64 == 0 and x.is_contiguous() and w.is_contiguous()
            and hasattr(K(), "pdb200_gemm_grouped"))


def _grouped_ref(x, w, counts):
    """Per-expert loop (CPU / fallback): the numerical oracle of the grouped kernels.  Reads the counts on the host."""
    out = torch.zeros(x.shape[0], w.shape[2], dtype=x.dtype, device=x.device)
    o = 0
    for e, c in enumerate(counts.tolist()):
        c = int(min(max(c, 0), x.shape[0] - o))
        if c > 0:
            out[o:o + c] = x[o:o + c] @ w[e]
        o += c
    return out


class _GroupedLinear(torch.autograd.Function):
    """rows of expert e (expert-major slab ``x`` [cap, K], ``counts`` int64[E] on the device) times ``w[e]`` [K, N].
    One launch per GEMM (forward, dX, dW), scheduled on the device from ``counts``: no host read."""

    @staticmethod
    def forward(ctx, x, w, counts):
        cap, Kd = x.shape
        E, _, N = w.shape
        out = torch.empty(cap, N, dtype=x.dtype, device=x.device)
        check(K().pdb200_gemm_grouped(ptr(x), ptr(w), ptr(out), ptr(counts), E, cap, N, Kd, 0, 0, stream_ptr(x.device)),
              "gemm_grouped")
        ctx.save_for_backward(x, w, counts)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, w, counts = ctx.saved_tensors
        cap, Kd = x.shape
        E, _, N = w.shape
        dy = dy.contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            # dX = dY . W[e]^T : the stacked [E][K][N] weights are exactly the "[E][N'][K']" layout with N' = K, K' = N
            dx = torch.empty(cap, Kd, dtype=x.dtype, device=x.device)
            check(K().pdb200_gemm_grouped(ptr(dy), ptr(w), ptr(dx), ptr(counts), E, cap, Kd, N, 1, 0, stream_ptr(x.device)),
                  "gemm_grouped dgrad")
        if ctx.needs_input_grad[1]:
            dw = torch.empty_like(w)
            check(K().pdb200_gemm_grouped_wgrad(ptr(x), ptr(dy), ptr(dw), ptr(counts), E, cap, Kd, N, 1, 0.0,
                                                stream_ptr(x.device)), "gemm_grouped wgrad")
        return dx, dw, None


def grouped_linear(x, w, counts):
    """``out[r] = x[r] @ w[e]`` for the rows r of expert e.  x: [cap, K] expert-major slab (rows beyond sum(counts) are
    ignored and their outputs are unspecified), w: [E, K, N], counts: int64[E] (device tensor on GPU)."""
    counts = counts.to(torch.int64).contiguous()
    if _grouped_native(x, w):
        return _GroupedLinear.apply(x, w, counts)
    return _grouped_ref(x, w, counts)
