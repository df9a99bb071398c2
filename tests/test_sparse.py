"""paddle.sparse against dense restatements (reference: python/paddle/sparse/*.py, python/paddle/sparse/nn/*)."""
import itertools
import math

import numpy as np
import pytest
import torch

import paddle
import paddle.sparse as S


def _rand_sparse(shape, density=0.4, seed=0, dense_dims=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(*shape, generator=g)
    lead = shape[:len(shape) - dense_dims]
    mask = torch.rand(*lead, generator=g) < density
    x = x * mask.reshape(lead + (1,) * dense_dims)
    return x, x.to_sparse(len(lead)).coalesce()


def test_creation_accessors_and_conversions():
    s = S.sparse_coo_tensor([[0, 1, 2, 0], [1, 0, 2, 1]], [1.0, -2.0, 3.0, 4.0])        # duplicate (0, 1) is summed
    assert tuple(s.shape) == (3, 3) and s.nnz() == 3 and s.is_sparse_coo() and not s.is_sparse_csr()
    assert s.indices().tolist() == [[0, 1, 2], [1, 0, 2]] and s.values().tolist() == [5.0, -2.0, 3.0]
    assert s.dtype == torch.float32
    d = S.sparse_coo_tensor([[0, 1]], [[1.0, 2.0], [3.0, 4.0]], [3, 2])                   # one sparse dim, one dense dim
    assert d.to_dense().tolist() == [[1.0, 2.0], [3.0, 4.0], [0.0, 0.0]]
    with pytest.raises(ValueError):
        S.sparse_coo_tensor([[0, 1]], [1.0])
    with pytest.raises(ValueError):
        S.sparse_coo_tensor([[0, 1]], [1.0, 2.0], [3, 3])
    c = S.sparse_csr_tensor([0, 2, 3, 5], [1, 3, 2, 0, 1], [1, 2, 3, 4, 5], [3, 4], dtype="float32")
    assert c.is_sparse_csr() and c.crows().tolist() == [0, 2, 3, 5] and c.cols().tolist() == [1, 3, 2, 0, 1] and c.nnz() == 5
    assert c.to_dense().tolist() == [[0, 1, 0, 2], [0, 0, 3, 0], [4, 5, 0, 0]]
    b = S.sparse_csr_tensor([0, 1, 2, 0, 1, 2], [0, 1, 1, 0], [1.0, 2.0, 3.0, 4.0], [2, 2, 2])
    assert b.to_dense().tolist() == [[[1.0, 0.0], [0.0, 2.0]], [[0.0, 3.0], [4.0, 0.0]]]
    ragged = S.sparse_csr_tensor([0, 1, 1, 0, 1, 2], [0, 1, 0], [1.0, 3.0, 4.0], [2, 2, 2])
    assert ragged.to_dense().tolist() == [[[1.0, 0.0], [0.0, 0.0]], [[0.0, 3.0], [4.0, 0.0]]]
    with pytest.raises(ValueError):
        S.sparse_csr_tensor([0, 2, 3], [1, 3, 2], [1, 2, 3], [3, 4])
    dense = torch.tensor([[0.0, 1.0], [2.0, 0.0]])
    assert dense.to_sparse_coo(2).nnz() == 2 and S.to_sparse_csr(dense).cols().tolist() == [1, 0]
    torch.testing.assert_close(c.to_sparse_coo().to_dense(), c.to_dense())
    assert S.is_same_shape(s, torch.zeros(3, 3)) and not S.is_same_shape(s, c)
    g = S.sparse_coo_tensor([[0], [0]], [1.0], [2, 2], stop_gradient=False)
    assert g.requires_grad


def test_unary_cast_and_shape_ops():
    dense, sp = _rand_sparse((5, 6), seed=1)
    dense = dense.clamp(-0.9, 0.9)
    sp = dense.to_sparse().coalesce()
    for name in ("sin", "tan", "asin", "atan", "sinh", "tanh", "asinh", "atanh", "square", "log1p", "abs", "neg", "deg2rad", "rad2deg", "expm1"):
        got = getattr(S, name)(sp)
        assert got.is_sparse_coo() and got.indices().tolist() == sp.indices().tolist(), name
        torch.testing.assert_close(got.to_dense(), getattr(torch, name)(dense), rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(getattr(S, name)(sp.to_sparse_csr()).to_dense(), getattr(torch, name)(dense), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(S.sqrt(S.abs(sp)).to_dense(), dense.abs().sqrt())
    torch.testing.assert_close(S.pow(sp, 3).to_dense(), dense ** 3)
    assert S.isnan(sp).values().dtype == torch.bool and not S.isnan(sp).values().any()
    with pytest.raises(TypeError):
        S.sin(dense)
    casted = S.cast(sp, "int32", "float64")
    assert casted.values().dtype == torch.float64
    ccsr = S.cast(sp.to_sparse_csr(), "int32", "float16")
    assert ccsr.crows().dtype == torch.int32 and ccsr.values().dtype == torch.float16
    # transpose / reshape / slice / sum
    d3, s3 = _rand_sparse((3, 4, 5), seed=2)
    torch.testing.assert_close(S.transpose(s3, [2, 0, 1]).to_dense(), d3.permute(2, 0, 1))
    torch.testing.assert_close(S.transpose(sp.to_sparse_csr(), [1, 0]).to_dense(), dense.t())
    torch.testing.assert_close(S.reshape(s3, [6, 10]).to_dense(), d3.reshape(6, 10))
    torch.testing.assert_close(S.reshape(s3, [0, -1]).to_dense(), d3.reshape(3, 20))
    torch.testing.assert_close(S.reshape(sp.to_sparse_csr(), [10, 3]).to_dense(), dense.reshape(10, 3))
    dd, sd = _rand_sparse((3, 4, 2), seed=3, dense_dims=1)
    torch.testing.assert_close(S.reshape(sd, [2, 6, 2]).to_dense(), dd.reshape(2, 6, 2))
    with pytest.raises(ValueError):
        S.reshape(s3, [7, 9])
    torch.testing.assert_close(S.slice(s3, [0, 2], [1, -3], [3, 5]).to_dense(), d3[1:3, :, 2:5])
    torch.testing.assert_close(S.slice(sp.to_sparse_csr(), [1], [2], [100]).to_dense(), dense[:, 2:])
    torch.testing.assert_close(S.slice(sd, [2], [1], [2]).to_dense(), dd[:, :, 1:2])
    torch.testing.assert_close(S.sum(s3).to_dense().reshape(()), d3.sum())
    torch.testing.assert_close(S.sum(s3, axis=1).to_dense(), d3.sum(1))
    torch.testing.assert_close(S.sum(s3, axis=-1, keepdim=True).to_dense(), d3.sum(-1, keepdim=True))
    assert S.sum(s3, axis=0, dtype="float64").values().dtype == torch.float64
    assert S.sum(sp.to_sparse_csr(), axis=1).to_dense().shape == (5,)
    dup = torch.sparse_coo_tensor([[0, 0], [1, 1]], [1.0, 2.0], (2, 2))
    assert S.coalesce(dup).values().tolist() == [3.0]


def test_binary_and_products():
    da, a = _rand_sparse((4, 5), seed=4)
    db, b = _rand_sparse((4, 5), seed=5)
    for layout in ("coo", "csr"):
        x, y = (a, b) if layout == "coo" else (a.to_sparse_csr(), b.to_sparse_csr())
        torch.testing.assert_close(S.add(x, y).to_dense(), da + db)
        torch.testing.assert_close(S.subtract(x, y).to_dense(), da - db)
        torch.testing.assert_close(S.multiply(x, y).to_dense(), da * db)
        assert S.add(x, y).layout == x.layout
        q = S.divide(x, y).to_dense()
        both = (da != 0) & (db != 0)
        torch.testing.assert_close(q[both], (da / db)[both])
        assert torch.isinf(q[(da != 0) & (db == 0)]).all()
    torch.testing.assert_close(S.multiply(a, 2.0).to_dense(), da * 2)
    torch.testing.assert_close(S.divide(a, 4.0).to_dense(), da / 4)
    with pytest.raises(ValueError):
        S.add(a, b.to_sparse_csr())
    with pytest.raises(ValueError):
        S.add(a, _rand_sparse((4, 6))[1])
    dense = torch.randn(5, 3)
    torch.testing.assert_close(S.matmul(a, dense), da @ dense, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(S.matmul(a.to_sparse_csr(), dense), da @ dense, rtol=1e-5, atol=1e-6)
    dc, c = _rand_sparse((5, 4), seed=6)
    ss = S.matmul(a, c)
    assert ss.is_sparse_coo()
    torch.testing.assert_close(ss.to_dense(), da @ dc, rtol=1e-5, atol=1e-6)
    d3, s3 = _rand_sparse((2, 4, 5), seed=7)
    torch.testing.assert_close(S.matmul(s3, torch.stack([dense, dense * 2])), torch.stack([d3[0] @ dense, d3[1] @ dense * 2]), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(S.mv(a, dense[:, 0]), da @ dense[:, 0], rtol=1e-5, atol=1e-6)
    with pytest.raises(ValueError):
        S.mv(a, dense)
    inp = torch.randn(4, 3)
    torch.testing.assert_close(S.addmm(inp, a, dense, beta=0.5, alpha=2.0), 0.5 * inp + 2.0 * (da @ dense), rtol=1e-5, atol=1e-6)
    dm, m = _rand_sparse((4, 4), seed=8)
    sp_out = S.addmm(m, a, c, beta=2.0, alpha=1.0)
    assert sp_out.is_sparse_coo()
    torch.testing.assert_close(sp_out.to_dense(), 2.0 * dm + da @ dc, rtol=1e-5, atol=1e-6)
    # masked_matmul / mask_as keep the mask's pattern
    x, y = torch.randn(4, 6), torch.randn(6, 4)
    for mask in (m, m.to_sparse_csr()):
        out = S.masked_matmul(x, y, mask)
        assert out.layout == mask.layout
        torch.testing.assert_close(out.to_dense(), (x @ y) * (dm != 0), rtol=1e-5, atol=1e-6)
        picked = S.mask_as(x[:, :4], mask)
        assert picked.layout == mask.layout
        torch.testing.assert_close(picked.to_dense(), x[:, :4] * (dm != 0))
    with pytest.raises(TypeError):
        S.mask_as(m, m)
    # randomised PCA recovers a low-rank matrix
    torch.manual_seed(0)
    low = (torch.randn(30, 3) @ torch.randn(3, 12)) * (torch.rand(30, 12) < 2).float()
    u, sv, v = S.pca_lowrank(low.to_sparse(), q=5, center=False, niter=3)
    assert u.shape == (30, 5) and sv.shape == (5,) and v.shape == (12, 5)
    torch.testing.assert_close(u @ torch.diag(sv) @ v.t(), low, rtol=1e-3, atol=1e-3)
    assert float(sv[3]) < 1e-3 * float(sv[0])
    with pytest.raises(ValueError):
        S.pca_lowrank(low, q=5)


def _conv3d_rulebook(sp, weight, bias, stride, padding, subm):
    """Loop restatement of sparse convolution: for every active input site and kernel offset, scatter
    value @ weight[offset] to the output site it lands on."""
    sp = sp.coalesce()
    n, d, h, w, cin = sp.shape
    kd, kh, kw, _, cout = weight.shape
    if subm:
        padding = ((kd - 1) // 2, (kh - 1) // 2, (kw - 1) // 2)
        out_shape = (d, h, w)
    else:
        out_shape = tuple((s + 2 * p - k) // st + 1 for s, p, k, st in zip((d, h, w), padding, (kd, kh, kw), stride))
    acc = {}
    active_in = {tuple(c) for c in sp.indices().t().tolist()}
    for (b, z, y, x), val in zip(sp.indices().t().tolist(), sp.values()):
        for oz, oy, ox in itertools.product(range(kd), range(kh), range(kw)):
            pz, py, px = z + padding[0] - oz, y + padding[1] - oy, x + padding[2] - ox
            if pz % stride[0] or py % stride[1] or px % stride[2]:
                continue
            site = (b, pz // stride[0], py // stride[1], px // stride[2])
            if not all(0 <= site[i + 1] < out_shape[i] for i in range(3)):
                continue
            if subm and site not in active_in:
                continue
            acc[site] = acc.get(site, torch.zeros(cout)) + val @ weight[oz, oy, ox]
    dense = torch.zeros((n,) + out_shape + (cout,))
    for site, v in acc.items():
        dense[site] = v + (bias if bias is not None else 0)
    return dense, set(acc)


def test_sparse_conv_pool_norm_activations():
    torch.manual_seed(0)
    x = torch.randn(2, 5, 4, 6, 3)
    x[torch.rand(2, 5, 4, 6) < 0.75] = 0
    sp = x.to_sparse(4).coalesce()
    w, bias = torch.randn(3, 3, 3, 3, 4) * 0.3, torch.randn(4)
    out = S.nn.functional.conv3d(sp, w, bias, stride=[2, 1, 2], padding=[1, 0, 1])
    want, sites = _conv3d_rulebook(sp, w, bias, (2, 1, 2), (1, 0, 1), False)
    assert tuple(out.shape) == tuple(want.shape) and {tuple(c) for c in out.indices().t().tolist()} == sites
    torch.testing.assert_close(out.to_dense(), want, rtol=1e-4, atol=1e-5)
    sub = S.nn.functional.subm_conv3d(sp, w, bias)
    want, sites = _conv3d_rulebook(sp, w, bias, (1, 1, 1), None, True)
    assert sub.indices().tolist() == sp.indices().tolist()                          # the pattern does not dilate
    torch.testing.assert_close(sub.to_dense(), want, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(S.nn.functional.subm_conv3d_igemm(sp, w, bias).to_dense(), sub.to_dense())
    # 2-D layers, gradients reach weights and input values
    x2 = torch.randn(2, 6, 7, 3)
    x2[torch.rand(2, 6, 7) < 0.6] = 0
    v2 = x2.to_sparse(3).coalesce()
    vals = v2.values().clone().requires_grad_(True)
    sp2 = torch.sparse_coo_tensor(v2.indices(), vals, v2.shape, is_coalesced=True)
    layer = S.nn.Conv2D(3, 5, 3, stride=2, padding=1)
    assert layer.weight.shape == (3, 3, 3, 5) and layer.bias.shape == (5,)
    out2 = layer(sp2)
    dense_ref = torch.nn.functional.conv2d(x2.permute(0, 3, 1, 2), layer.weight.permute(3, 2, 0, 1), layer.bias, 2, 1).permute(0, 2, 3, 1)
    occ = torch.nn.functional.conv2d((x2.abs().sum(-1, keepdim=True) > 0).float().permute(0, 3, 1, 2), torch.ones(1, 1, 3, 3), None, 2, 1)[:, 0] > 0
    torch.testing.assert_close(out2.to_dense(), dense_ref * occ[..., None], rtol=1e-4, atol=1e-5)
    out2.coalesce().values().square().sum().backward()
    assert vals.grad is not None and layer.weight.grad is not None and float(layer.weight.grad.abs().sum()) > 0
    subm2 = S.nn.SubmConv2D(3, 5, 3)
    assert subm2(v2).indices().tolist() == v2.indices().tolist()
    with pytest.raises(ValueError):
        S.nn.functional.subm_conv2d(v2, subm2.weight, stride=2)
    with pytest.raises(ValueError):
        S.nn.Conv3D(3, 4, 3, data_format="NCDHW")
    with pytest.raises(TypeError):
        S.nn.functional.conv3d(x, w)
    # max pooling over ACTIVE sites: negative maxima survive
    neg = -x.abs()
    spn = neg.to_sparse(4).coalesce()
    pooled = S.nn.MaxPool3D(2, stride=2)(spn)
    want = torch.zeros(2, 2, 2, 3, 3)
    for b, z, y, xx in itertools.product(range(2), range(2), range(2), range(3)):
        win = neg[b, 2 * z:2 * z + 2, 2 * y:2 * y + 2, 2 * xx:2 * xx + 2].reshape(-1, 3)
        act = win[(win != 0).any(-1)]
        if len(act):
            want[b, z, y, xx] = act.max(0).values
    torch.testing.assert_close(pooled.to_dense(), want)
    assert float(pooled.values().max()) <= 0
    # batch norm over the active sites only
    bn = S.nn.BatchNorm(3, momentum=0.8)
    bn.train()
    y = bn(sp)
    v = sp.values()
    torch.testing.assert_close(y.values(), (v - v.mean(0)) / torch.sqrt(v.var(0, unbiased=False) + 1e-5), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(bn._mean, 0.2 * v.mean(0))
    torch.testing.assert_close(bn._variance, 0.8 + 0.2 * v.var(0, unbiased=False))
    assert y.indices().tolist() == sp.indices().tolist() and sorted(bn.state_dict()) == ["_mean", "_variance", "bias", "weight"]
    bn.eval()
    torch.testing.assert_close(bn(sp).values(), (v - bn._mean) / torch.sqrt(bn._variance + 1e-5), rtol=1e-4, atol=1e-5)
    sync = S.nn.SyncBatchNorm.convert_sync_batchnorm(paddle.nn.Sequential(S.nn.BatchNorm(3), S.nn.ReLU()))
    assert isinstance(sync[0], S.nn.SyncBatchNorm)
    sync.train()
    torch.testing.assert_close(sync(sp).values(), torch.relu((v - v.mean(0)) / torch.sqrt(v.var(0, unbiased=False) + 1e-5)), rtol=1e-4, atol=1e-5)
    # activations and softmax act on stored entries
    d, s = _rand_sparse((4, 6), seed=9)
    torch.testing.assert_close(S.nn.ReLU()(s).to_dense(), torch.relu(d))
    torch.testing.assert_close(S.nn.ReLU6()(S.multiply(s, 10.0)).to_dense(), torch.nn.functional.relu6(d * 10))
    torch.testing.assert_close(S.nn.LeakyReLU(0.2)(s.to_sparse_csr()).to_dense(), torch.nn.functional.leaky_relu(d, 0.2))
    for layout in (s, s.to_sparse_csr()):
        sm = S.nn.Softmax()(layout)
        assert sm.layout == layout.layout
        want = torch.softmax(d.masked_fill(d == 0, float("-inf")), -1)
        want = torch.where(torch.isnan(want), torch.zeros_like(want), want)
        torch.testing.assert_close(sm.to_dense(), want, rtol=1e-5, atol=1e-6)
    with pytest.raises(ValueError):
        S.nn.functional.softmax(s, axis=0)


def test_sparse_attention():
    torch.manual_seed(0)
    b, h, s, d = 2, 2, 6, 4
    q, k, v = (torch.randn(b, h, s, d) for _ in range(3))
    pattern = torch.tril(torch.ones(s, s)).bool() & (torch.rand(b * h, s, s) < 0.8)
    pattern |= torch.eye(s, dtype=torch.bool)
    mask = pattern.float().to_sparse()                                   # ragged per-batch counts: COO
    kp = torch.ones(b, s)
    kp[1, 0] = 0
    am = torch.ones(s, s)
    am[3, 1] = 0
    out = S.nn.functional.attention(q, k, v, mask, kp, am)
    allowed = pattern.reshape(b, h, s, s) & (kp.reshape(b, 1, 1, s) != 0) & (am.reshape(1, 1, s, s) != 0)
    scores = (q @ k.transpose(-1, -2) / math.sqrt(d)).masked_fill(~allowed, float("-inf"))
    want = torch.nan_to_num(torch.softmax(scores, -1), nan=0.0) @ v        # a row with nothing allowed gives zeros
    torch.testing.assert_close(out, want, rtol=1e-5, atol=1e-6)
    assert float(out[1, :, 0].abs().sum()) == 0
    plain = S.nn.functional.attention(q, k, v, mask)
    scores = (q @ k.transpose(-1, -2) / math.sqrt(d)).masked_fill(~pattern.reshape(b, h, s, s), float("-inf"))
    torch.testing.assert_close(plain, torch.softmax(scores, -1) @ v, rtol=1e-5, atol=1e-6)
    causal = torch.tril(torch.ones(b * h, s, s)).to_sparse_csr()         # equal counts per batch: batched CSR
    got = S.nn.functional.attention(q, k, v, causal)
    torch.testing.assert_close(got, torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True), rtol=1e-5, atol=1e-6)
