"""paddle.sparse.nn.functional (reference: python/paddle/sparse/nn/functional/{activation,conv,pooling,transformer}.py).

Sparse convolution / pooling take COO tensors in channel-last layout (``[N, D, H, W, C]`` / ``[N, H, W, C]``) whose
first dims are sparse and whose channel dim is dense, and weights shaped ``[*kernel, C_in / groups, C_out]``.  They are
computed on the densified input with the library convolution and read back at the active output sites:
  * ordinary convolution: an output site is active when at least one active input site falls into its window;
  * submanifold convolution (stride 1): the output sites are exactly the input sites.
Contributions of inactive sites are zero either way, so the values equal the rulebook-based reference's.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as TF

__all__ = ["conv2d", "conv3d", "subm_conv2d", "subm_conv2d_igemm", "subm_conv3d", "subm_conv3d_igemm", "max_pool3d", "relu",
           "relu6", "leaky_relu", "softmax", "attention"]


def _vals(x):
    return x.coalesce().values() if x.layout == torch.sparse_coo else x.values()


def _with_values(x, v):
    if x.layout == torch.sparse_coo:
        x = x.coalesce()
        return torch.sparse_coo_tensor(x.indices(), v, x.shape, is_coalesced=True)
    return torch.sparse_csr_tensor(x.crow_indices(), x.col_indices(), v, size=x.shape)


def relu(x, name=None):
    return _with_values(x, torch.relu(_vals(x)))


def relu6(x, name=None):
    return _with_values(x, TF.relu6(_vals(x)))


def leaky_relu(x, negative_slope=0.01, name=None):
    return _with_values(x, TF.leaky_relu(_vals(x), negative_slope))


def softmax(x, axis=-1, name=None):
    """Softmax over the STORED entries of every row (missing entries count as -inf, i.e. probability 0)."""
    if axis not in (-1, x.dim() - 1):
        raise ValueError("sparse softmax only supports the last axis")
    if x.layout == torch.sparse_csr:
        coo = x.to_sparse_coo().coalesce()
        out = torch.sparse.softmax(coo, dim=coo.dim() - 1).coalesce()
        return out.to_sparse_csr()
    return torch.sparse.softmax(x.coalesce(), dim=x.dim() - 1).coalesce()


def attention(query, key, value, sparse_mask, key_padding_mask=None, attn_mask=None, name=None):
    """softmax(Q K^T / sqrt(d)) V with the softmax restricted to the pattern of ``sparse_mask`` (CSR or COO, shape
    [batch * heads, seq, seq]; only WHICH entries are stored matters, not their values); ``key_padding_mask`` [batch, seq] and ``attn_mask`` [seq, seq] zero entries out further
    (a 0 there masks the position)."""
    b, h, s, d = query.shape
    scores = torch.matmul(query, key.transpose(-1, -2)) / math.sqrt(d)
    stored = sparse_mask.to_sparse(layout=torch.sparse_coo).coalesce() if sparse_mask.layout == torch.sparse_csr else sparse_mask.coalesce()
    pattern = torch.zeros(b * h, s, s, dtype=torch.bool)
    idx = stored.indices()
    pattern[idx[0], idx[1], idx[2]] = True
    pattern = pattern.reshape(b, h, s, s)
    if key_padding_mask is not None:
        pattern = pattern & (key_padding_mask.reshape(b, 1, 1, s) != 0)
    if attn_mask is not None:
        pattern = pattern & (attn_mask.reshape(1, 1, s, s) != 0)
    scores = scores.masked_fill(~pattern, float("-inf"))
    probs = torch.softmax(scores, -1)
    probs = torch.where(pattern.any(-1, keepdim=True), probs, torch.zeros_like(probs))
    return torch.matmul(probs, value)


# ------------------------------------------------------------------------------------------------------ conv / pooling
def _tuple(v, nd):
    return tuple(int(e) for e in v) if isinstance(v, (list, tuple)) else (int(v),) * nd


def _check_input(x, nd, fmt):
    if x.layout != torch.sparse_coo:
        raise TypeError("sparse convolution / pooling expects a SparseCooTensor")
    if x.dim() != nd + 2:
        raise ValueError(f"expected a {nd + 2}-D input in {fmt} layout, got shape {tuple(x.shape)}")
    x = x.coalesce()
    if x.sparse_dim() != nd + 1:
        raise ValueError(f"the input should have {nd + 1} sparse dims (batch + spatial) and a dense channel dim")
    return x


def _occupancy(x, nd):
    idx = x.indices()
    occ = torch.zeros((x.shape[0], 1) + tuple(x.shape[1:nd + 1]), dtype=x.values().dtype)
    occ[(idx[0], torch.zeros_like(idx[0])) + tuple(idx[i] for i in range(1, nd + 1))] = 1.0
    return occ


def _conv(nd, x, weight, bias, stride, padding, dilation, groups, subm, data_format):
    fmt = "NDHWC" if nd == 3 else "NHWC"
    if data_format != fmt:
        raise ValueError(f"Attr(data_format) should be '{fmt}'. Received Attr(data_format): {data_format}.")
    x = _check_input(x, nd, fmt)
    stride, padding, dilation = _tuple(stride, nd), _tuple(padding, nd), _tuple(dilation, nd)
    if weight.dim() != nd + 2:
        raise ValueError(f"weight should be [*kernel, C_in / groups, C_out], got shape {tuple(weight.shape)}")
    if subm and any(s != 1 for s in stride):
        raise ValueError("submanifold convolution needs stride 1")
    ksize = tuple(weight.shape[:nd])
    if subm:
        padding = tuple((k - 1) // 2 * d for k, d in zip(ksize, dilation))           # output sites = input sites
    dense = x.to_dense().permute(0, nd + 1, *range(1, nd + 1))                       # N C spatial
    w = weight.permute(nd + 1, nd, *range(nd))                                        # C_out, C_in / g, *kernel
    conv = TF.conv3d if nd == 3 else TF.conv2d
    out = conv(dense, w, bias, stride, padding, dilation, groups)
    if subm:
        sites = x.indices()
    else:
        hit = conv(_occupancy(x, nd), torch.ones((1, 1) + ksize, dtype=dense.dtype), None, stride, padding, dilation)
        sites = torch.nonzero(hit[:, 0] > 0.5).t()
    vals = out.permute(0, *range(2, nd + 2), 1)[tuple(sites[i] for i in range(nd + 1))]
    shape = (out.shape[0],) + tuple(out.shape[2:]) + (out.shape[1],)
    return torch.sparse_coo_tensor(sites, vals, shape).coalesce()


def conv3d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, data_format="NDHWC", name=None):
    return _conv(3, x, weight, bias, stride, padding, dilation, groups, False, data_format)


def conv2d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, data_format="NHWC", name=None):
    return _conv(2, x, weight, bias, stride, padding, dilation, groups, False, data_format)


def subm_conv3d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, data_format="NDHWC", key=None, name=None):
    return _conv(3, x, weight, bias, stride, padding, dilation, groups, True, data_format)


def subm_conv2d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, data_format="NHWC", key=None, name=None):
    return _conv(2, x, weight, bias, stride, padding, dilation, groups, True, data_format)


# the implicit-GEMM variants differ from the plain ones only in the kernel the reference picks
subm_conv3d_igemm = subm_conv3d
subm_conv2d_igemm = subm_conv2d


def max_pool3d(x, kernel_size, stride=None, padding=0, ceil_mode=False, data_format="NDHWC", name=None):
    """Maximum over the ACTIVE sites of every window (inactive sites do not count as zeros); an output site is active
    when its window holds an active input site."""
    if data_format != "NDHWC":
        raise ValueError("Currently, sparse.max_pool3d only support data format of 'NDHWC'")
    x = _check_input(x, 3, "NDHWC")
    k = _tuple(kernel_size, 3)
    s = k if stride is None else _tuple(stride, 3)
    p = _tuple(padding, 3)
    occ = _occupancy(x, 3)
    dense = x.to_dense().permute(0, 4, 1, 2, 3)
    dense = torch.where(occ > 0.5, dense, torch.full_like(dense, float("-inf")))
    out = TF.max_pool3d(dense, k, s, p, ceil_mode=ceil_mode)
    hit = TF.max_pool3d(occ, k, s, p, ceil_mode=ceil_mode)
    sites = torch.nonzero(hit[:, 0] > 0.5).t()
    vals = out.permute(0, 2, 3, 4, 1)[tuple(sites[i] for i in range(4))]
    shape = (out.shape[0],) + tuple(out.shape[2:]) + (out.shape[1],)
    return torch.sparse_coo_tensor(sites, vals, shape).coalesce()
