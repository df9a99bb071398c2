"""Paddle semantics of the tensor API where they differ from torch's (argument names, defaults, return conventions),
checked against numpy restatements of the reference's documented behaviour (python/paddle/tensor/*.py)."""
import numpy as np
import pytest
import torch

import paddle_b200 as paddle


def T(a, dtype=None):
    return paddle.to_tensor(np.asarray(a), dtype=dtype)


def eq(t, ref, tol=1e-6):
    np.testing.assert_allclose(np.asarray(t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else t), np.asarray(ref), rtol=tol, atol=tol)


x = np.arange(24, dtype=np.float32).reshape(2, 3, 4)


def test_shape_manipulation():
    eq(paddle.transpose(T(x), perm=[2, 0, 1]), x.transpose(2, 0, 1))
    eq(paddle.reshape(T(x), [0, -1]), x.reshape(2, -1))                  # 0 copies the input dim
    eq(paddle.flatten(T(x), start_axis=1, stop_axis=2), x.reshape(2, 12))
    eq(paddle.squeeze(T(x[:, :1]), axis=[1]), x[:, 0])
    eq(paddle.unsqueeze(T(x), axis=[0, 2]), x[None][:, :, None])
    eq(paddle.concat([T(x), T(x)], axis=-1), np.concatenate([x, x], -1))
    eq(paddle.stack([T(x), T(x)], axis=1), np.stack([x, x], 1))
    a, b, c = paddle.split(T(x), num_or_sections=[1, 2, -1], axis=2)     # -1: the rest
    eq(a, x[:, :, :1]); eq(b, x[:, :, 1:3]); eq(c, x[:, :, 3:])
    parts = paddle.split(T(x), 3, axis=1)
    assert len(parts) == 3 and tuple(parts[0].shape) == (2, 1, 4)
    assert [tuple(p.shape) for p in paddle.chunk(T(x), 2, axis=2)] == [(2, 3, 2)] * 2
    eq(paddle.tile(T(x[0]), repeat_times=[2, 1]), np.tile(x[0], (2, 1)))
    eq(paddle.expand(T(x[:, :1]), shape=[2, 3, 4]), np.broadcast_to(x[:, :1], (2, 3, 4)))
    eq(paddle.expand(T(x[:, :1]), shape=[-1, 3, -1]), np.broadcast_to(x[:, :1], (2, 3, 4)))
    eq(paddle.broadcast_to(T(x[0, 0]), [3, 4]), np.broadcast_to(x[0, 0], (3, 4)))
    eq(paddle.roll(T(x), shifts=1, axis=2), np.roll(x, 1, 2))
    eq(paddle.roll(T(x), shifts=2), np.roll(x.reshape(-1), 2).reshape(x.shape))    # no axis: flattened roll
    eq(paddle.flip(T(x), axis=[0, 2]), x[::-1, :, ::-1])
    eq(paddle.slice(T(x), axes=[1, 2], starts=[1, 0], ends=[3, 2]), x[:, 1:3, 0:2])
    eq(paddle.strided_slice(T(x), axes=[2], starts=[0], ends=[4], strides=[2]), x[:, :, 0:4:2])
    eq(paddle.unbind(T(x), axis=1)[2], x[:, 2])
    eq(paddle.moveaxis(T(x), 0, 2), np.moveaxis(x, 0, 2))
    eq(paddle.t(T(x[0])), x[0].T)
    eq(paddle.cast(T(x), "int32"), x.astype(np.int32))
    eq(paddle.repeat_interleave(T(x[0]), 2, axis=0), np.repeat(x[0], 2, 0))
    eq(paddle.unstack(T(x), axis=0)[1], x[1])
    eq(paddle.crop(T(x), shape=[1, 2, 2], offsets=[1, 1, 1]), x[1:2, 1:3, 1:3])


def test_indexing_and_search():
    idx = np.array([2, 0])
    eq(paddle.gather(T(x), T(idx), axis=1), x[:, idx])
    eq(paddle.index_select(T(x), T(idx), axis=2), x[:, :, idx])
    eq(paddle.gather_nd(T(x), T([[0, 1], [1, 2]])), x[[0, 1], [1, 2]])
    base = np.zeros((4, 3), np.float32)
    upd = np.ones((3, 3), np.float32)
    i3 = np.array([1, 1, 3])
    ow = base.copy(); ow[i3] = upd
    eq(paddle.scatter(T(base), T(i3), T(upd), overwrite=True), ow)
    acc = base.copy(); np.add.at(acc, i3, upd)
    eq(paddle.scatter(T(base), T(i3), T(upd), overwrite=False), acc)    # accumulate mode: rows zeroed first, then summed
    eq(paddle.scatter_nd_add(T(base), T([[1], [1]]), T(np.ones((2, 3), np.float32))), acc * 0 + np.array([[0] * 3, [2] * 3, [0] * 3, [0] * 3]))
    eq(paddle.take_along_axis(T(x), T(np.zeros((2, 3, 1), np.int64)), axis=2), x[:, :, :1])
    eq(paddle.put_along_axis(T(x), T(np.zeros((2, 3, 1), np.int64)), 9.0, axis=2)[:, :, 0], np.full((2, 3), 9.0))
    eq(paddle.masked_select(T(x), T(x > 20)), x[x > 20])
    eq(paddle.where(T(x > 5), T(x), T(-x)), np.where(x > 5, x, -x))
    nz = paddle.nonzero(T(x[0] > 5))
    eq(nz, np.argwhere(x[0] > 5))                                          # [n, rank] matrix, not a tuple
    assert isinstance(paddle.nonzero(T(x[0] > 5), as_tuple=True), tuple)
    v, i = paddle.topk(T(x), k=2, axis=1, largest=False)
    eq(v, np.sort(x, 1)[:, :2]); eq(i, np.argsort(x, 1)[:, :2])
    eq(paddle.argsort(T(x), axis=-1, descending=True), np.argsort(-x, -1))
    eq(paddle.sort(T(-x), axis=1), np.sort(-x, 1))
    eq(paddle.argmax(T(x), axis=1, keepdim=True), x.argmax(1)[:, None])
    eq(paddle.argmax(T(x)), x.argmax())                                    # flattened when axis is None
    eq(paddle.searchsorted(T([1., 3., 5.]), T([0., 3., 6.])), [0, 1, 3])
    eq(paddle.bincount(T([1, 1, 3])), [0, 2, 0, 1])
    u, inv, cnt = paddle.unique(T([3, 1, 3, 2]), return_inverse=True, return_counts=True)
    eq(u, [1, 2, 3]); eq(inv, [2, 0, 2, 1]); eq(cnt, [1, 1, 2])
    eq(paddle.index_sample(T(x[0]), T([[0, 3], [1, 1], [2, 0]])), np.take_along_axis(x[0], np.array([[0, 3], [1, 1], [2, 0]]), 1))
    eq(paddle.diagonal(T(x), offset=1, axis1=1, axis2=2), np.diagonal(x, 1, 1, 2))
    eq(paddle.tril(T(x[0]), diagonal=-1), np.tril(x[0], -1)); eq(paddle.triu(T(x[0]), 1), np.triu(x[0], 1))
    eq(paddle.one_hot(T([0, 2]), num_classes=3) if hasattr(paddle, "one_hot") else paddle.nn.functional.one_hot(T([0, 2]), 3), np.eye(3)[[0, 2]])


def test_math_reductions_and_defaults():
    eq(paddle.sum(T(x), axis=[0, 2], keepdim=True), x.sum((0, 2), keepdims=True))
    eq(paddle.mean(T(x), axis=-1), x.mean(-1))
    eq(paddle.std(T(x), axis=1), x.std(1, ddof=1)); eq(paddle.std(T(x), axis=1, unbiased=False), x.std(1))
    eq(paddle.var(T(x)), x.var(ddof=1))
    eq(paddle.max(T(x), axis=1), x.max(1)); eq(paddle.amin(T(x), axis=[1, 2]), x.min((1, 2)))
    eq(paddle.prod(T(x[0, :, :2] + 1), axis=0), (x[0, :, :2] + 1).prod(0))
    eq(paddle.cumsum(T(x)), x.reshape(-1).cumsum()); eq(paddle.cumsum(T(x), axis=1), x.cumsum(1))
    eq(paddle.cumprod(T(x[0] + 1), dim=1), (x[0] + 1).cumprod(1))
    eq(paddle.logsumexp(T(x), axis=2), np.log(np.exp(x).sum(2)), 1e-5)
    eq(paddle.norm(T(x), p=2, axis=1), np.sqrt((x ** 2).sum(1)), 1e-5); eq(paddle.norm(T(x)), np.sqrt((x ** 2).sum()), 1e-5)
    eq(paddle.norm(T(x), p="fro", axis=[1, 2]), np.sqrt((x ** 2).sum((1, 2))), 1e-5)
    eq(paddle.clip(T(x), min=3, max=10), np.clip(x, 3, 10))
    eq(paddle.pow(T(x), 2), x ** 2); eq(paddle.square(T(x)), x ** 2); eq(paddle.rsqrt(T(x + 1)), 1 / np.sqrt(x + 1), 1e-5)
    eq(paddle.floor_divide(T([7, -7]), T([2, 2])), [3, -4])
    eq(paddle.remainder(T([7., -7.]), T([3., 3.])), [1., 2.]); eq(paddle.mod(T([7, -7]), T([3, 3])), [1, 2])
    eq(paddle.scale(T(x), scale=2.0, bias=1.0), 2 * x + 1); eq(paddle.scale(T(x), 2.0, 1.0, bias_after_scale=False), 2 * (x + 1))
    eq(paddle.add_n([T(x), T(x), T(x)]), 3 * x)
    eq(paddle.matmul(T(x[0]), T(x[1]), transpose_y=True), x[0] @ x[1].T); eq(paddle.matmul(T(x[0]), T(x[1]), transpose_x=True), x[0].T @ x[1])
    eq(paddle.bmm(T(x), T(x.transpose(0, 2, 1))), x @ x.transpose(0, 2, 1))
    eq(paddle.dot(T(x[0]), T(x[1])), (x[0] * x[1]).sum(-1))                  # batched: row-wise dot
    eq(paddle.mv(T(x[0]), T(x[0, 0])), x[0] @ x[0, 0])
    eq(paddle.outer(T([1., 2.]), T([3., 4., 5.])), np.outer([1, 2], [3, 4, 5]))
    eq(paddle.kron(T([[1., 2.]]), T([[1.], [3.]])), np.kron([[1, 2]], [[1], [3]]))
    eq(paddle.einsum("bij,bkj->bik", T(x), T(x)), np.einsum("bij,bkj->bik", x, x))
    eq(paddle.trace(T(x[0]), offset=1), np.trace(x[0], 1))
    eq(paddle.diff(T(x), axis=2), np.diff(x, axis=2))
    eq(paddle.lerp(T([0., 10.]), T([10., 20.]), 0.5), [5., 15.])
    eq(paddle.isclose(T([1.0, 2.0]), T([1.0, 2.1])), [True, False])
    assert bool(paddle.allclose(T(x), T(x + 1e-9))) and bool(paddle.equal_all(T(x), T(x)))
    eq(paddle.logical_and(T([True, False]), T([True, True])), [True, False]); eq(paddle.bitwise_xor(T([5]), T([3])), [6])
    eq(paddle.sign(T([-2., 0., 3.])), [-1, 0, 1]); eq(paddle.reciprocal(T([2., 4.])), [0.5, 0.25])
    eq(paddle.maximum(T([1., 5.]), T([3., 2.])), [3, 5]); eq(paddle.fmax(T([np.nan, 1.]), T([2., np.nan])), [2, 1])
    eq(paddle.stanh(T([1.]), scale_a=0.67, scale_b=1.7159), 1.7159 * np.tanh(0.67), 1e-5)
    eq(paddle.increment(T([1.]), 2.0), [3.])
    eq(paddle.median(T([[1., 3., 2.], [4., 6., 5.]]), axis=1), [2, 5]); eq(paddle.quantile(T([1., 2., 3., 4.]), 0.5), 2.5)
    eq(paddle.kthvalue(T([3., 1., 2.]), 2)[0], 2.0); eq(paddle.mode(T([1, 2, 2, 3]))[0], 2)
    eq(paddle.count_nonzero(T(x), axis=1), (x != 0).sum(1)); eq(paddle.numel(T(x)), 24)
    eq(paddle.rank(T(x)), 3); eq(paddle.shape(T(x)), [2, 3, 4])


def test_creation_defaults():
    assert paddle.zeros([2, 3]).dtype == torch.float32 and paddle.ones([2], dtype="int64").dtype == torch.int64
    eq(paddle.full([2, 2], 3.5), np.full((2, 2), 3.5)); eq(paddle.full_like(T(x), 2), np.full_like(x, 2))
    eq(paddle.arange(1, 7, 2), [1, 3, 5]); assert paddle.arange(5).dtype == torch.int64 and paddle.arange(0, 1, 0.5).dtype == torch.float32
    eq(paddle.linspace(0, 1, 5), np.linspace(0, 1, 5)); eq(paddle.eye(2, 3), np.eye(2, 3))
    eq(paddle.diag(T([1., 2.]), offset=1), np.diag([1., 2.], 1)); eq(paddle.diag(T(x[0])), np.diag(x[0]))
    eq(paddle.meshgrid(T([1, 2]), T([3, 4, 5]))[0], np.meshgrid([1, 2], [3, 4, 5], indexing="ij")[0])
    eq(paddle.tril_indices(3, 3, 0), np.stack(np.tril_indices(3)))
    assert paddle.to_tensor([1.0, 2.0]).dtype == torch.float32 and paddle.to_tensor([1, 2]).dtype == torch.int64
    assert paddle.to_tensor(np.float64(1.0)).dtype == torch.float64 and paddle.to_tensor(1.0).dtype == torch.float32
    assert paddle.to_tensor([1.0], stop_gradient=False).requires_grad
    paddle.seed(3); a = paddle.rand([4]); paddle.seed(3); b = paddle.rand([4]); eq(a, b)
    assert tuple(paddle.randint(0, 5, [3, 2]).shape) == (3, 2) and paddle.randint(0, 5, [3]).dtype == torch.int64
    assert tuple(paddle.randperm(5).shape) == (5,) and tuple(paddle.normal(0.0, 1.0, [2, 2]).shape) == (2, 2)
    assert tuple(paddle.uniform([2, 3], min=-1, max=1).shape) == (2, 3) and tuple(paddle.empty_like(T(x)).shape) == x.shape
    eq(paddle.assign(T(x)), x); eq(paddle.clone(T(x)), x)
    eq(paddle.complex(T([1.]), T([2.])), [1 + 2j]); eq(paddle.real(T([1 + 2j])), [1.]); eq(paddle.angle(T([1j])), [np.pi / 2], 1e-6)


def test_tensor_methods_follow_the_functions():
    t = T(x)
    eq(t.transpose([1, 0, 2]), x.transpose(1, 0, 2)); eq(t.reshape([6, 4]), x.reshape(6, 4)); eq(t.astype("float64"), x.astype(np.float64))
    eq(t.sum(axis=1), x.sum(1)); eq(t.max(axis=2, keepdim=True), x.max(2, keepdims=True)); eq(t.unsqueeze(0), x[None])
    eq(t.flatten(), x.reshape(-1)); eq(t.numpy(), x); assert t.numel() == 24 and t.ndim == 3 and tuple(t.shape) == (2, 3, 4)
    eq(t.expand([2, 2, 3, 4])[1], x); eq(t.tile([1, 1, 2]), np.tile(x, (1, 1, 2))); eq(t.cast("int64"), x.astype(np.int64))
    eq(t.clip(2, 5), np.clip(x, 2, 5)); eq(t.matmul(T(x.transpose(0, 2, 1))), x @ x.transpose(0, 2, 1)); eq(t.norm(), np.sqrt((x ** 2).sum()), 1e-5)
    eq(t.gather(T([1]), axis=1), x[:, [1]]); eq(t.split(2, axis=2)[1], x[:, :, 2:]); eq(t.argmax(axis=-1), x.argmax(-1))
    g = T(x); g.stop_gradient = False
    (g * 2).sum().backward()
    eq(g.grad, np.full_like(x, 2)); g.clear_gradient(); assert g.grad is None or float(g.grad.abs().sum()) == 0
    assert g.detach().stop_gradient and not g.stop_gradient and t.place is not None and t.is_contiguous()
    eq(t.add_(T(x)), 2 * x)                                                 # in place
    eq(t.scale_(0.5), x)


def test_every_reference_tensor_method_exists_and_top_p_sampling():
    import os
    import re
    path = "/root/reference/python/paddle/tensor/__init__.py"
    if os.path.exists(path):
        names = re.findall(r"['\"]([^'\"]+)['\"]", re.search(r"tensor_method_func = \[(.*?)\]", open(path).read(), re.S).group(1))
        assert [n for n in names if not hasattr(torch.Tensor, n)] == []
    x = torch.rand(3, 3) + torch.eye(3)
    torch.testing.assert_close(x.pinv(), torch.linalg.pinv(x))
    assert x.atleast_3d().shape == (3, 3, 1) and x.tensordot(x, axes=1).shape == (3, 3) and bool(x.is_tensor())
    assert paddle.create_tensor("int64").dtype == torch.int64 and paddle.create_tensor("float32").numel() == 0
    y = torch.zeros(2, 3)
    assert y.put_along_axis_(torch.tensor([[0], [2]]), torch.tensor([[5.0], [7.0]]), 1) is y and y.tolist() == [[5.0, 0, 0], [0, 0, 7.0]]
    # nucleus sampling: only the tokens needed to reach p can be drawn
    probs = torch.tensor([[0.05, 0.6, 0.3, 0.05], [0.4, 0.3, 0.2, 0.1]])
    seen = [set(), set()]
    for s in range(200):
        v, i = paddle.top_p_sampling(probs, torch.tensor([0.5, 0.65]), seed=s)
        assert v.shape == (2, 1) and i.dtype == torch.int64
        torch.testing.assert_close(v, probs.gather(1, i))
        seen[0].add(int(i[0]))
        seen[1].add(int(i[1]))
    assert seen[0] == {1} and seen[1] == {0, 1}            # row 0: the top token alone covers 0.5; row 1: 0.4 + 0.3 crosses 0.65
    a = paddle.top_p_sampling(probs, torch.tensor([0.99, 0.99]), seed=7)[1]
    b = paddle.top_p_sampling(probs, torch.tensor([0.99, 0.99]), seed=7)[1]
    assert a.tolist() == b.tolist()
    only_top = {int(paddle.top_p_sampling(probs, torch.tensor([0.99, 0.99]), threshold=torch.tensor([0.5, 0.35]), seed=s)[1][1]) for s in range(50)}
    assert only_top == {0}
