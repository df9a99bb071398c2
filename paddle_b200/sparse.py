"""paddle.sparse (reference: python/paddle/sparse/*.py): COO / CSR creation and the core ops on torch.sparse."""
from __future__ import annotations

import torch


def sparse_coo_tensor(indices, values, shape=None, dtype=None, place=None, stop_gradient=True):
    i, v = torch.as_tensor(indices), torch.as_tensor(values)
    if dtype is not None:
        from .framework.dtype import to_torch_dtype as convert_dtype
        v = v.to(convert_dtype(dtype))
    t = torch.sparse_coo_tensor(i, v, size=tuple(shape) if shape is not None else None).coalesce()
    t.requires_grad_(not stop_gradient and v.is_floating_point())
    return t


def sparse_csr_tensor(crows, cols, values, shape, dtype=None, place=None, stop_gradient=True):
    v = torch.as_tensor(values)
    if dtype is not None:
        from .framework.dtype import to_torch_dtype as convert_dtype
        v = v.to(convert_dtype(dtype))
    return torch.sparse_csr_tensor(torch.as_tensor(crows), torch.as_tensor(cols), v, size=tuple(shape))


def matmul(x, y, name=None):
    return torch.sparse.mm(x, y) if x.is_sparse or x.layout == torch.sparse_csr else torch.matmul(x, y)


def masked_matmul(x, y, mask, name=None):
    return torch.sparse.sampled_addmm(mask.to_sparse_csr() if mask.layout != torch.sparse_csr else mask, x, y, beta=0.0)


def add(x, y, name=None):
    return x + y


def subtract(x, y, name=None):
    return x - y


def multiply(x, y, name=None):
    return x * y


def transpose(x, perm, name=None):
    return x.permute(*perm) if not x.is_sparse else x.transpose(perm[0], perm[1]) if len(perm) == 2 else x.to_dense().permute(*perm).to_sparse()


def sum(x, axis=None, dtype=None, keepdim=False, name=None):  # noqa: A001
    return torch.sparse.sum(x) if axis is None else torch.sparse.sum(x, dim=axis)


def to_dense(x):
    return x.to_dense()


def to_sparse_coo(x, sparse_dim=None):
    return x.to_sparse() if sparse_dim is None else x.to_sparse(sparse_dim)


def to_sparse_csr(x):
    return x.to_sparse_csr()


def is_same_shape(x, y):
    return tuple(x.shape) == tuple(y.shape)


class nn:  # noqa: N801  (namespace, as paddle.sparse.nn)
    class ReLU(torch.nn.Module):
        def forward(self, x):
            return torch.sparse_coo_tensor(x.indices(), torch.relu(x.values()), x.shape) if x.is_sparse else torch.relu(x)

    class Softmax(torch.nn.Module):
        def __init__(self, axis=-1):
            super().__init__()
            self.axis = axis

        def forward(self, x):
            return torch.sparse.softmax(x, dim=self.axis)
