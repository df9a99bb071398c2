"""paddle.sparse (reference: python/paddle/sparse/{creation,unary,binary,multiary}.py).

Sparse tensors are torch COO / CSR tensors; the Paddle accessor names (``crows`` / ``cols`` / ``nnz`` /
``is_sparse_coo`` / ``is_sparse_csr`` / ``to_sparse_coo``) are added to ``torch.Tensor`` below.  Element-wise functions
act on the stored values only and keep the sparsity pattern (they are all zero-preserving); results of COO operations
are coalesced, as the reference's kernels produce them.
"""
from __future__ import annotations

import math
import warnings

import torch

warnings.filterwarnings("ignore", message="Sparse CSR tensor support is in beta state")
torch.sparse.check_sparse_tensor_invariants.disable()          # explicit choice: the constructors below validate their inputs

__all__ = [
    "sparse_coo_tensor", "sparse_csr_tensor", "sin", "tan", "asin", "atan", "sinh", "tanh", "asinh", "atanh", "sqrt", "square",
    "log1p", "abs", "pow", "pca_lowrank", "cast", "neg", "deg2rad", "rad2deg", "expm1", "mv", "matmul", "mask_as", "masked_matmul",
    "addmm", "add", "subtract", "transpose", "sum", "multiply", "divide", "coalesce", "is_same_shape", "reshape", "isnan", "slice",
]


def _dtype(dtype):
    if dtype is None or isinstance(dtype, torch.dtype):
        return dtype
    from ..framework.dtype import to_torch_dtype
    return to_torch_dtype(dtype)


def _is_coo(x):
    return isinstance(x, torch.Tensor) and x.layout == torch.sparse_coo


def _is_csr(x):
    return isinstance(x, torch.Tensor) and x.layout == torch.sparse_csr


def _is_sparse(x):
    return _is_coo(x) or _is_csr(x)


# --------------------------------------------------------------------------------------------------------- creation
def sparse_coo_tensor(indices, values, shape=None, dtype=None, place=None, stop_gradient=True):
    """``indices`` [sparse_dim, nnz], ``values`` [nnz, *dense dims]; the shape defaults to the smallest that holds every
    index.  The result is coalesced (sorted indices, duplicates summed)."""
    i = torch.as_tensor(indices)
    if i.dtype not in (torch.int32, torch.int64):
        i = i.to(torch.int64)
    v = values if isinstance(values, torch.Tensor) else torch.as_tensor(values)
    if not isinstance(values, torch.Tensor) and v.dtype == torch.float64:
        v = v.to(torch.get_default_dtype())
    if dtype is not None:
        v = v.to(_dtype(dtype))
    if i.dim() != 2:
        raise ValueError("'indices' must be a 2-D tensor.")
    if i.shape[1] != v.shape[0]:
        raise ValueError(f"the indices and values must have same number of non-zero, but get {i.shape[1]} and {v.shape[0]}")
    if shape is None:
        lead = (i.max(1).values + 1).tolist() if i.numel() else [0] * i.shape[0]
        shape = list(lead) + list(v.shape[1:])
    elif len(shape) != i.shape[0] + v.dim() - 1:
        raise ValueError(f"the number of dimensions(len(shape) must be sparse_dim({i.shape[0]}) + dense_dim({v.dim() - 1}), but get {len(shape)}")
    t = torch.sparse_coo_tensor(i.to(torch.int64), v, size=tuple(int(s) for s in shape)).coalesce()
    if not stop_gradient and v.dtype.is_floating_point:
        t.requires_grad_(True)
    return t


def sparse_csr_tensor(crows, cols, values, shape, dtype=None, place=None, stop_gradient=True):
    """2-D ``[M, N]`` or batched 3-D ``[B, M, N]`` CSR: ``crows`` has M + 1 entries per batch (B * (M + 1) when
    batched, concatenated as the reference stores them)."""
    cr, co = torch.as_tensor(crows).to(torch.int64), torch.as_tensor(cols).to(torch.int64)
    v = values if isinstance(values, torch.Tensor) else torch.as_tensor(values)
    if not isinstance(values, torch.Tensor) and v.dtype == torch.float64:
        v = v.to(torch.get_default_dtype())
    if dtype is not None:
        v = v.to(_dtype(dtype))
    shape = tuple(int(s) for s in shape)
    if len(shape) not in (2, 3):
        raise ValueError(f"SparseCsrTensor only support 2-D or 3-D matrix. but get shape {list(shape)}")
    if co.shape[0] != v.shape[0]:
        raise ValueError("the length of cols must be same as length of values")
    if len(shape) == 2:
        if cr.shape[0] != shape[0] + 1:
            raise ValueError(f"The length({cr.shape[0]}) of crows must be equal to the rows({shape[0]})+1 of matrix.")
        if int(cr[0]) != 0 or int(cr[-1]) != v.shape[0]:
            raise ValueError("the 0th value of crows must be 0 and the last value must be the number of non-zeros")
        t = torch.sparse_csr_tensor(cr, co, v, size=shape)
    else:
        b, m = shape[0], shape[1]
        if cr.shape[0] != b * (m + 1):
            raise ValueError(f"The length({cr.shape[0]}) of crows must be equal to batch * (rows + 1) = {b * (m + 1)}")
        cr2 = cr.reshape(b, m + 1)
        per = (cr2[:, -1]).tolist()
        if len(set(per)) == 1:
            t = torch.sparse_csr_tensor(cr2, co.reshape(b, -1), v.reshape(b, -1), size=shape)
        else:
            # ragged batches: torch's batched CSR needs equal counts, so go through COO
            rows = torch.cat([torch.repeat_interleave(torch.arange(m), cr2[i, 1:] - cr2[i, :-1]) for i in range(b)])
            batch = torch.repeat_interleave(torch.arange(b), torch.as_tensor(per))
            return torch.sparse_coo_tensor(torch.stack([batch, rows, co]), v, size=shape).coalesce()
    if not stop_gradient and v.dtype.is_floating_point:
        t.requires_grad_(True)
    return t


# ---------------------------------------------------------------------------------------------- accessors on Tensor
def _install_accessors():
    T = torch.Tensor
    if not hasattr(T, "crows"):
        T.crows = lambda self: self.crow_indices().reshape(-1)
        T.cols = lambda self: self.col_indices().reshape(-1)
    if not hasattr(T, "nnz"):
        T.nnz = lambda self: int(self._nnz()) if self.layout != torch.strided else int((self != 0).sum())
    if not hasattr(T, "is_sparse_coo"):
        T.is_sparse_coo = lambda self: self.layout == torch.sparse_coo
        T.is_sparse_csr = lambda self: self.layout == torch.sparse_csr
    if not getattr(T.to_sparse_coo, "_pd_wrapped", False):
        original = T.to_sparse_coo

        def _to_sparse_coo(self, sparse_dim=None):               # torch's takes no argument; Paddle's takes sparse_dim
            if self.layout == torch.sparse_csr:
                return original(self)
            return to_sparse_coo(self, sparse_dim)
        _to_sparse_coo._pd_wrapped = True
        T.to_sparse_coo = _to_sparse_coo


def to_dense(x):
    return x.to_dense()


def to_sparse_coo(x, sparse_dim=None):
    if _is_coo(x):
        return x
    if _is_csr(x):
        return x.to_sparse(layout=torch.sparse_coo).coalesce()
    return (x.to_sparse() if sparse_dim is None else x.to_sparse(int(sparse_dim))).coalesce()


def to_sparse_csr(x):
    return x if _is_csr(x) else x.to_sparse_csr()


_install_accessors()


def _rebuild(x, values):
    """A tensor with ``x``'s pattern and new values."""
    if _is_coo(x):
        x = x.coalesce()
        return torch.sparse_coo_tensor(x.indices(), values, x.shape, is_coalesced=True)
    return torch.sparse_csr_tensor(x.crow_indices(), x.col_indices(), values, size=x.shape)


def _values(x):
    return x.coalesce().values() if _is_coo(x) else x.values()


# -------------------------------------------------------------------------------------------------------------- unary
def _unary(fn, name):
    def op(x, name=None):
        if not _is_sparse(x):
            raise TypeError(f"paddle.sparse.{op.__name__} expects a sparse tensor")
        return _rebuild(x, fn(_values(x)))
    op.__name__ = name
    op.__doc__ = f"Element-wise ``{name}`` of the stored values (the pattern is kept: {name}(0) = 0)."
    return op


sin, tan, asin, atan = _unary(torch.sin, "sin"), _unary(torch.tan, "tan"), _unary(torch.asin, "asin"), _unary(torch.atan, "atan")
sinh, tanh, asinh, atanh = (_unary(torch.sinh, "sinh"), _unary(torch.tanh, "tanh"), _unary(torch.asinh, "asinh"),
                            _unary(torch.atanh, "atanh"))
sqrt, square, log1p, expm1 = (_unary(torch.sqrt, "sqrt"), _unary(torch.square, "square"), _unary(torch.log1p, "log1p"),
                              _unary(torch.expm1, "expm1"))
abs, neg = _unary(torch.abs, "abs"), _unary(torch.neg, "neg")                                   # noqa: A001
deg2rad, rad2deg = _unary(torch.deg2rad, "deg2rad"), _unary(torch.rad2deg, "rad2deg")
isnan = _unary(torch.isnan, "isnan")


def pow(x, factor, name=None):  # noqa: A001
    return _rebuild(x, torch.pow(_values(x), factor))


def cast(x, index_dtype=None, value_dtype=None, name=None):
    """New index and / or value dtypes (None keeps the current one)."""
    v = _values(x) if value_dtype is None else _values(x).to(_dtype(value_dtype))
    idt = _dtype(index_dtype)
    if _is_coo(x):
        x = x.coalesce()
        i = x.indices() if idt is None else x.indices().to(idt)
        out = torch.sparse_coo_tensor(i.to(torch.int64), v, x.shape, is_coalesced=True)       # torch keeps COO indices in int64
        out._pd_index_dtype = idt or torch.int64
        return out
    cr, co = x.crow_indices(), x.col_indices()
    if idt is not None:
        cr, co = cr.to(idt), co.to(idt)
    return torch.sparse_csr_tensor(cr, co, v, size=x.shape)


def coalesce(x, name=None):
    """Sort the indices and sum duplicate entries."""
    return x.coalesce() if _is_coo(x) else x


def transpose(x, perm, name=None):
    perm = [int(p) % x.dim() for p in perm]
    if sorted(perm) != list(range(x.dim())):
        raise ValueError(f"perm {perm} is not a permutation of the {x.dim()} dims")
    if _is_coo(x):
        x = x.coalesce()
        sd = x.sparse_dim()
        if any(p >= sd for p in perm[:sd]):
            return x.to_dense().permute(*perm).to_sparse(sd).coalesce()
        new_idx = x.indices()[perm[:sd]]
        vals = x.values().permute(0, *[p - sd + 1 for p in perm[sd:]]) if x.dense_dim() else x.values()
        return torch.sparse_coo_tensor(new_idx, vals, [x.shape[p] for p in perm]).coalesce()
    if _is_csr(x):
        return transpose(x.to_sparse_coo(), perm).to_sparse_csr()
    return x.permute(*perm)


def reshape(x, shape, name=None):
    """Reshape the SPARSE dims of a COO tensor (dense dims must keep their shape); 0 copies a dim, -1 is inferred."""
    if _is_csr(x):
        return reshape(x.to_sparse_coo(), shape).to_sparse_csr()
    x = x.coalesce()
    shape = [int(s) for s in shape]
    shape = [x.shape[i] if s == 0 else s for i, s in enumerate(shape)]
    if -1 in shape:
        known = math.prod(s for s in shape if s != -1)
        shape[shape.index(-1)] = x.numel() // known if known else 0
    if math.prod(shape) != x.numel():
        raise ValueError(f"cannot reshape a tensor of {x.numel()} elements to {shape}")
    dd = x.dense_dim()
    dense_shape = list(x.shape[x.dim() - dd:]) if dd else []
    if dd and shape[len(shape) - dd:] != dense_shape:
        raise ValueError("reshape of a sparse tensor must keep its dense dims")
    old_sparse = list(x.shape[:x.sparse_dim()])
    new_sparse = shape[:len(shape) - dd] if dd else shape
    strides = [math.prod(old_sparse[i + 1:]) for i in range(len(old_sparse))]
    flat = (x.indices() * torch.tensor(strides, dtype=torch.int64)[:, None]).sum(0)
    idx = []
    for i in range(len(new_sparse)):
        step = math.prod(new_sparse[i + 1:])
        idx.append((flat // step) % new_sparse[i])
    return torch.sparse_coo_tensor(torch.stack(idx), x.values(), shape).coalesce()


def slice(x, axes, starts, ends, name=None):  # noqa: A001
    """``x[..., starts[i]:ends[i], ...]`` along every ``axes[i]`` (negative positions count from the end)."""
    if _is_csr(x):
        return slice(x.to_sparse_coo(), axes, starts, ends).to_sparse_csr()
    x = x.coalesce()
    idx, vals, shape = x.indices(), x.values(), list(x.shape)
    keep = torch.ones(idx.shape[1], dtype=torch.bool)
    offsets = {}
    for ax, st, en in zip(axes, starts, ends):
        ax = int(ax) % x.dim()
        n = shape[ax]
        st, en = int(st), int(en)
        st = max(st + n, 0) if st < 0 else min(st, n)
        en = max(en + n, 0) if en < 0 else min(en, n)
        en = max(en, st)
        if ax < x.sparse_dim():
            keep &= (idx[ax] >= st) & (idx[ax] < en)
            offsets[ax] = st
        else:
            vals = vals.narrow(ax - x.sparse_dim() + 1, st, en - st)
        shape[ax] = en - st
    idx = idx[:, keep].clone()
    for ax, st in offsets.items():
        idx[ax] -= st
    return torch.sparse_coo_tensor(idx, vals[keep], shape).coalesce()


def sum(x, axis=None, dtype=None, keepdim=False, name=None):  # noqa: A001
    """Sum over ``axis`` (all elements when None); the result is sparse, as in the reference."""
    layout_csr = _is_csr(x)
    c = x.to_sparse_coo().coalesce() if layout_csr else x.coalesce()
    td = _dtype(dtype)
    if axis is None:
        total = c.values().sum() if td is None else c.values().sum(dtype=td)
        shape = [1] * x.dim() if keepdim else [1]
        out = torch.sparse_coo_tensor(torch.zeros((len(shape), 1), dtype=torch.int64), total.reshape(1), shape).coalesce()
        return out
    axis = int(axis) % x.dim()
    out = torch.sparse.sum(c, dim=axis) if td is None else torch.sparse.sum(c, dim=axis, dtype=td)
    if not out.is_sparse:
        out = out.to_sparse()
    out = out.coalesce()
    if keepdim:
        sd = out.sparse_dim()
        if axis <= sd:
            idx = torch.cat([out.indices()[:axis], torch.zeros((1, out._nnz()), dtype=torch.int64), out.indices()[axis:]])
            shape = list(out.shape[:axis]) + [1] + list(out.shape[axis:])
            out = torch.sparse_coo_tensor(idx, out.values(), shape).coalesce()
    return out.to_sparse_csr() if layout_csr and out.dim() in (2, 3) else out


# ------------------------------------------------------------------------------------------------------------- binary
def _same_layout(x, y, op):
    if not (_is_sparse(x) and _is_sparse(y)):
        raise TypeError(f"paddle.sparse.{op} expects two sparse tensors")
    if x.layout != y.layout:
        raise ValueError(f"paddle.sparse.{op}: both operands must have the same sparse layout")
    if tuple(x.shape) != tuple(y.shape):
        raise ValueError(f"paddle.sparse.{op}: shapes {tuple(x.shape)} and {tuple(y.shape)} differ")


def _binary(x, y, fn, op, union):
    """Element-wise ``fn`` over the union (add / subtract) or the intersection-as-union (multiply / divide: computed
    over the union with zeros filled in, as the reference's kernels do - so divide gives inf / nan where y has no entry)."""
    _same_layout(x, y, op)
    csr = _is_csr(x)
    a, b = (x.to_sparse_coo().coalesce(), y.to_sparse_coo().coalesce()) if csr else (x.coalesce(), y.coalesce())
    # union pattern through an all-ones add
    ones_a = torch.sparse_coo_tensor(a.indices(), torch.ones(a._nnz(), dtype=torch.float32), a.shape[:a.sparse_dim()])
    ones_b = torch.sparse_coo_tensor(b.indices(), torch.ones(b._nnz(), dtype=torch.float32), b.shape[:b.sparse_dim()])
    pattern = (ones_a + ones_b).coalesce().indices()
    va, vb = _gather_at(a, pattern), _gather_at(b, pattern)
    out = torch.sparse_coo_tensor(pattern, fn(va, vb), a.shape, is_coalesced=True)
    return out.to_sparse_csr() if csr else out


def _gather_at(x, pattern):
    """Values of coalesced COO ``x`` at the index columns ``pattern`` (zeros where x has no entry)."""
    sd = x.sparse_dim()
    sizes = list(x.shape[:sd])
    strides = torch.tensor([math.prod(sizes[i + 1:]) for i in range(sd)], dtype=torch.int64)[:, None]
    key_x = (x.indices() * strides).sum(0)
    key_p = (pattern * strides).sum(0)
    pos = torch.searchsorted(key_x, key_p)
    pos_c = pos.clamp(max=max(key_x.numel() - 1, 0))
    hit = (key_x[pos_c] == key_p) if key_x.numel() else torch.zeros_like(key_p, dtype=torch.bool)
    vals = x.values()
    out = torch.zeros((pattern.shape[1],) + tuple(vals.shape[1:]), dtype=vals.dtype)
    if key_x.numel():
        picked = vals[pos_c]
        mask = hit.reshape((-1,) + (1,) * (vals.dim() - 1))
        out = torch.where(mask, picked, out)
    return out


def add(x, y, name=None):
    if _is_sparse(x) and not _is_sparse(y):
        return x.to_dense() + y
    return _binary(x, y, torch.add, "add", True)


def subtract(x, y, name=None):
    return _binary(x, y, torch.sub, "subtract", True)


def multiply(x, y, name=None):
    if isinstance(y, (int, float)):
        return _rebuild(x, _values(x) * y)
    return _binary(x, y, torch.mul, "multiply", False)


def divide(x, y, name=None):
    if isinstance(y, (int, float)):
        return _rebuild(x, _values(x) / y)
    return _binary(x, y, torch.div, "divide", False)


def is_same_shape(x, y):
    return tuple(x.shape) == tuple(y.shape)


# ----------------------------------------------------------------------------------------------------------- products
def _to_mm_layout(x):
    return x.coalesce() if _is_coo(x) else x


def matmul(x, y, name=None):
    """sparse @ dense -> dense, sparse @ sparse -> sparse, dense @ dense -> dense; batched 3-D operands are multiplied
    batch by batch."""
    if not _is_sparse(x) and not _is_sparse(y):
        return torch.matmul(x, y)
    if x.dim() == 2:
        if _is_sparse(y):
            out = torch.sparse.mm(_to_mm_layout(x), _to_mm_layout(y))
            return out.coalesce() if _is_coo(out) else out
        return torch.sparse.mm(_to_mm_layout(x), y)
    if x.dim() == 3:
        xs = x.to_sparse_coo().coalesce() if _is_csr(x) else x.coalesce()
        outs = [matmul(_select_batch(xs, b), y[b] if y.dim() == 3 else y) for b in range(x.shape[0])]
        return torch.stack([o.to_dense() if _is_sparse(o) else o for o in outs])
    raise ValueError("paddle.sparse.matmul supports 2-D and batched 3-D operands")


def _select_batch(x, b):
    sel = x.indices()[0] == b
    return torch.sparse_coo_tensor(x.indices()[1:, sel], x.values()[sel], x.shape[1:]).coalesce()


def masked_matmul(x, y, mask, name=None):
    """(dense x @ dense y) evaluated only at the positions stored in the sparse ``mask``; the result has its layout."""
    csr = mask if _is_csr(mask) else mask.to_sparse_csr()
    pattern = torch.sparse_csr_tensor(csr.crow_indices(), csr.col_indices(), torch.zeros_like(csr.values(), dtype=x.dtype), size=csr.shape)
    out = torch.sparse.sampled_addmm(pattern, x, y, beta=0.0)
    return out if _is_csr(mask) else out.to_sparse_coo().coalesce()


def mv(x, vec, name=None):
    """sparse [M, N] @ dense [N] -> dense [M]."""
    if vec.dim() != 1:
        raise ValueError("vec should be a 1-D tensor")
    return torch.sparse.mm(_to_mm_layout(x), vec.unsqueeze(1)).squeeze(1)


def addmm(input, x, y, beta=1.0, alpha=1.0, name=None):
    """beta * input + alpha * (x @ y)."""
    prod = matmul(x, y)
    if _is_sparse(input) or _is_sparse(prod):
        a = input.to_dense() if _is_sparse(input) else input
        p = prod.to_dense() if _is_sparse(prod) else prod
        out = beta * a + alpha * p
        if _is_sparse(input) and _is_sparse(prod):
            return out.to_sparse_csr() if _is_csr(input) else out.to_sparse().coalesce()
        return out
    return beta * input + alpha * prod


def mask_as(x, mask, name=None):
    """The entries of DENSE ``x`` at the positions stored in sparse ``mask``, in ``mask``'s layout."""
    if _is_sparse(x):
        raise TypeError("mask_as expects a dense x")
    if _is_coo(mask):
        m = mask.coalesce()
        idx = m.indices()
        vals = x[tuple(idx[i] for i in range(idx.shape[0]))]
        return torch.sparse_coo_tensor(idx, vals, x.shape, is_coalesced=True)
    coo = mask.to_sparse_coo().coalesce()
    idx = coo.indices()
    vals = x[tuple(idx[i] for i in range(idx.shape[0]))]
    return torch.sparse_coo_tensor(idx, vals, x.shape).coalesce().to_sparse_csr()


def pca_lowrank(x, q=None, center=True, niter=2, name=None):
    """Randomised PCA of a sparse 2-D matrix: (U, S, V) with ``x - mean ~ U diag(S) V^T`` (Halko et al.)."""
    if not _is_sparse(x):
        raise ValueError(f"Input must be sparse, but got {type(x)} instead")
    m, n = x.shape[-2:]
    q = min(6, m, n) if q is None else q
    if not 0 <= q <= min(m, n):
        raise ValueError(f"q(={q}) must be non-negative integer and not greater than min(m, n)={min(m, n)}")
    if not niter >= 0:
        raise ValueError(f"niter(={niter}) must be non-negative integer")
    dense = x.to_dense()
    if center:
        dense = dense - dense.mean(-2, keepdim=True)
    g = torch.randn(n, q, dtype=dense.dtype)
    basis = torch.linalg.qr(dense @ g).Q
    for _ in range(niter):
        basis = torch.linalg.qr(dense.transpose(-1, -2) @ basis).Q
        basis = torch.linalg.qr(dense @ basis).Q
    small = basis.transpose(-1, -2) @ dense
    u, s, vh = torch.linalg.svd(small, full_matrices=False)
    return basis @ u, s, vh.transpose(-1, -2)


from . import nn  # noqa: E402,F401
