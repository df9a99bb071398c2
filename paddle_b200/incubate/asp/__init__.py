"""paddle.incubate.asp - automatic n:m structured sparsity (reference: python/paddle/incubate/asp/{asp,utils,
supported_layer_list}.py).  2:4 sparsity is what Blackwell's sparse tensor-core path consumes (the doubled peak figures);
this module produces and maintains the masks:

    asp.prune_model(model)                    # magnitude-based 2:4 masks on the supported layers' weights
    opt = asp.decorate(opt)                   # the masks are re-applied after every optimizer step

Masks are computed along the input (reduction) dimension of a weight, which for Paddle's [in, out] Linear layout is dim 0:
every group of m consecutive input channels of one output channel keeps its n largest magnitudes (``mask_1d``); ``mask_2d_*``
additionally constrains every m x m block to n per row AND n per column (needed when the transposed weight is used by the
backward GEMM as well)."""
from __future__ import annotations

import itertools
from enum import Enum

import torch

__all__ = ["calculate_density", "decorate", "prune_model", "set_excluded_layers", "reset_excluded_layers",
           "add_supported_layer"]


class MaskAlgo(Enum):
    MASK_1D = "get_mask_1d"
    MASK_2D_GREEDY = "get_mask_2d_greedy"
    MASK_2D_BEST = "get_mask_2d_best"


class CheckMethod(Enum):
    CHECK_1D = "check_mask_1d"
    CHECK_2D = "check_mask_2d"

    @staticmethod
    def get_checking_method(mask_algo):
        return CheckMethod.CHECK_1D if mask_algo == MaskAlgo.MASK_1D else CheckMethod.CHECK_2D


def calculate_density(x):
    x = torch.as_tensor(x)
    return float((x != 0).sum()) / max(x.numel(), 1)


def _groups(mat, m):
    """[rows, cols] -> [rows * ceil(cols / m), m] (zero padded), plus what is needed to undo it."""
    rows, cols = mat.shape
    pad = (-cols) % m
    if pad:
        mat = torch.cat([mat, mat.new_zeros(rows, pad)], 1)
    return mat.reshape(-1, m), (rows, cols, pad)


def get_mask_1d(mat, n, m):
    """Keep the n largest magnitudes of every m consecutive elements of each row."""
    mat = torch.as_tensor(mat)
    g, (rows, cols, pad) = _groups(mat.abs().float(), m)
    keep = g.topk(n, dim=1).indices
    mask = torch.zeros_like(g).scatter_(1, keep, 1.0)
    return mask.reshape(rows, cols + pad)[:, :cols].to(mat.dtype)


def check_mask_1d(mat, n, m):
    mat = torch.as_tensor(mat)
    g, _ = _groups((mat != 0).float(), m)
    return bool((g.sum(1) <= n).all())


def _blocks(mat, m):
    rows, cols = mat.shape
    pr, pc = (-rows) % m, (-cols) % m
    full = torch.zeros(rows + pr, cols + pc, dtype=mat.dtype)
    full[:rows, :cols] = mat
    b = full.reshape((rows + pr) // m, m, (cols + pc) // m, m).permute(0, 2, 1, 3).reshape(-1, m, m)
    return b, (rows, cols, pr, pc)


def _unblocks(b, meta, m):
    rows, cols, pr, pc = meta
    full = b.reshape((rows + pr) // m, (cols + pc) // m, m, m).permute(0, 2, 1, 3).reshape(rows + pr, cols + pc)
    return full[:rows, :cols]


def get_mask_2d_greedy(mat, n, m):
    """Every m x m block: visit elements by decreasing magnitude, keep one while its row and its column hold < n."""
    mat = torch.as_tensor(mat)
    b, meta = _blocks(mat.abs().float().cpu(), m)
    mask = torch.zeros_like(b)
    order = b.reshape(b.shape[0], -1).argsort(dim=1, descending=True)
    for k in range(b.shape[0]):
        rc, cc = [0] * m, [0] * m
        for idx in order[k].tolist():
            i, j = divmod(idx, m)
            if rc[i] < n and cc[j] < n:
                mask[k, i, j] = 1.0
                rc[i] += 1
                cc[j] += 1
    return _unblocks(mask, meta, m).to(mat.dtype).to(mat.device)


_patterns_cache = {}


def _valid_2d_patterns(n, m):
    key = (n, m)
    if key not in _patterns_cache:
        rows = [torch.tensor(p, dtype=torch.float32) for p in set(itertools.permutations([1.0] * n + [0.0] * (m - n)))]
        pats = []
        for combo in itertools.product(range(len(rows)), repeat=m):
            p = torch.stack([rows[i] for i in combo])
            if bool((p.sum(0) == n).all()):
                pats.append(p)
        _patterns_cache[key] = torch.stack(pats)
    return _patterns_cache[key]


def get_mask_2d_best(mat, n, m):
    """Every m x m block: the valid pattern (n per row and per column) that keeps the largest total magnitude."""
    mat = torch.as_tensor(mat)
    b, meta = _blocks(mat.abs().float().cpu(), m)
    pats = _valid_2d_patterns(n, m)                                   # [P, m, m]
    score = torch.einsum("bij,pij->bp", b, pats)
    mask = pats[score.argmax(1)]
    return _unblocks(mask, meta, m).to(mat.dtype).to(mat.device)


def check_mask_2d(mat, n, m):
    mat = torch.as_tensor(mat)
    b, _ = _blocks((mat != 0).float().cpu(), m)
    return bool((b.sum(1) <= n).all() and (b.sum(2) <= n).all())


def create_mask(tensor, func_name=MaskAlgo.MASK_1D, n=2, m=4):
    """Mask of a weight of any rank: trailing dims are flattened into the grouped dimension like the reference does
    (conv kernels [out, in, kh, kw] are masked along in * kh * kw)."""
    t = torch.as_tensor(tensor)
    shape = t.shape
    if t.dim() == 1:
        mat = t.reshape(1, -1)
    elif t.dim() == 2:
        mat = t.t()                     # Paddle Linear is [in, out]: group along the input channels
    else:
        mat = t.reshape(shape[0], -1)
    fn = globals()[func_name.value if isinstance(func_name, MaskAlgo) else func_name]
    mask = fn(mat, n, m)
    if t.dim() == 2:
        mask = mask.t()
    return mask.reshape(shape).contiguous()


def check_sparsity(tensor, func_name=CheckMethod.CHECK_1D, n=2, m=4):
    t = torch.as_tensor(tensor)
    mat = t.reshape(1, -1) if t.dim() == 1 else (t.t() if t.dim() == 2 else t.reshape(t.shape[0], -1))
    fn = globals()[func_name.value if isinstance(func_name, CheckMethod) else func_name]
    return fn(mat, n, m)


# ----------------------------------------------------------------------------------------------- model-level workflow
_excluded = set()
_supported = {}
_masks = {}                    # id(param) -> (param, mask)


def _default_supported():
    from ... import nn
    return (nn.Linear, nn.Conv2D) + tuple(_supported)


def add_supported_layer(layer, pruning_func=None):
    """Register another layer type (or a parameter-name substring) whose ``weight`` is prunable."""
    _supported[layer] = pruning_func


def set_excluded_layers(param_names, main_program=None):
    _excluded.update(param_names)


def reset_excluded_layers(main_program=None):
    _excluded.clear()


def prune_model(model, n=2, m=4, mask_algo="mask_1d", with_mask=True):
    """Prune the weights of the supported layers to n:m and (``with_mask``) remember the masks for ``decorate``d
    optimizers.  Returns ``{parameter name: mask}``."""
    algo = {"mask_1d": MaskAlgo.MASK_1D, "mask_2d_greedy": MaskAlgo.MASK_2D_GREEDY, "mask_2d_best": MaskAlgo.MASK_2D_BEST}[mask_algo]
    types = tuple(t for t in _default_supported() if isinstance(t, type))
    names = [t for t in _supported if isinstance(t, str)]
    out = {}
    for lname, layer in model.named_sublayers(include_self=True):
        w = getattr(layer, "weight", None)
        if w is None or not isinstance(w, torch.Tensor) or w.dim() < 2:
            continue
        pname = getattr(w, "name", lname + ".weight")
        if not (isinstance(layer, types) or any(s in pname or s in lname for s in names)):
            continue
        if pname in _excluded or lname in _excluded or (lname + ".weight") in _excluded:
            continue
        grouped = w.shape[0] if w.dim() == 2 else w[0].numel()
        if grouped % m:
            continue                                   # the grouped dimension must be a multiple of m
        fn = _supported.get(type(layer))
        mask = fn(w.detach(), n, m, algo) if callable(fn) else create_mask(w.detach(), algo, n, m)
        with torch.no_grad():
            w.mul_(mask.to(w.dtype).to(w.device))
        if with_mask:
            _masks[id(w)] = (w, mask.to(w.dtype).to(w.device))
        out[pname] = mask
    return out


class OptimizerWithSparsityGuarantee:
    """``decorate`` result (reference: asp.py OptimizerWithSparsityGuarantee): after each step the pruned weights are
    multiplied by their masks again, so that the update cannot revive a pruned element."""

    def __init__(self, optimizer):
        self.__dict__["_optimizer"] = optimizer

    def __getattr__(self, k):
        return getattr(self._optimizer, k)

    def __setattr__(self, k, v):
        setattr(self._optimizer, k, v)

    @torch.no_grad()
    def step(self):
        self._optimizer.step()
        for w, mask in _masks.values():
            w.mul_(mask)

    def minimize(self, loss, *a, **k):
        out = self._optimizer.minimize(loss, *a, **k)
        with torch.no_grad():
            for w, mask in _masks.values():
                w.mul_(mask)
        return out


def decorate(optimizer):
    return OptimizerWithSparsityGuarantee(optimizer)
