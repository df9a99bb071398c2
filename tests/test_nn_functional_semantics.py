"""Paddle semantics of paddle.nn.functional where they differ from torch.nn.functional (reference:
python/paddle/nn/functional/*.py), against numpy / explicit torch restatements."""
import numpy as np
import pytest
import torch
import torch.nn.functional as TF

import paddle_b200 as paddle
import paddle_b200.nn.functional as F


def T(a, dtype=None):
    return paddle.to_tensor(np.asarray(a), dtype=dtype)


def eq(t, ref, tol=1e-5):
    r = ref.detach().cpu().numpy() if isinstance(ref, torch.Tensor) else np.asarray(ref)
    np.testing.assert_allclose(t.detach().cpu().numpy(), r, rtol=tol, atol=tol)


rng = np.random.RandomState(0)
x = rng.randn(2, 3, 4).astype(np.float32)


def test_activations_and_softmax_axis():
    eq(F.softmax(T(x), axis=1), TF.softmax(torch.tensor(x), 1)); eq(F.log_softmax(T(x), axis=0), TF.log_softmax(torch.tensor(x), 0))
    eq(F.gelu(T(x), approximate=True), TF.gelu(torch.tensor(x), approximate="tanh")); eq(F.gelu(T(x)), TF.gelu(torch.tensor(x)))
    eq(F.leaky_relu(T(x), negative_slope=0.2), np.where(x > 0, x, 0.2 * x)); eq(F.relu6(T(x * 5)), np.clip(x * 5, 0, 6))
    eq(F.elu(T(x), alpha=0.5), np.where(x > 0, x, 0.5 * (np.exp(x) - 1))); eq(F.silu(T(x)), x / (1 + np.exp(-x)))
    eq(F.hardswish(T(x)), x * np.clip(x + 3, 0, 6) / 6); eq(F.hardsigmoid(T(x)), np.clip(x / 6 + 0.5, 0, 1))      # slope 1/6 (paddle: 0.1666667)
    eq(F.swish(T(x)), x / (1 + np.exp(-x))); eq(F.softplus(T(x), beta=2, threshold=20), np.log1p(np.exp(2 * x)) / 2)
    eq(F.glu(T(x), axis=-1), x[..., :2] / (1 + np.exp(-x[..., 2:]))); eq(F.normalize(T(x), p=2, axis=1), x / np.sqrt((x ** 2).sum(1, keepdims=True)))
    eq(F.softsign(T(x)), x / (1 + np.abs(x))); eq(F.mish(T(x)), x * np.tanh(np.log1p(np.exp(x)))); eq(F.tanhshrink(T(x)), x - np.tanh(x))
    eq(F.thresholded_relu(T(x), threshold=0.5), np.where(x > 0.5, x, 0)); eq(F.hardtanh(T(x), min=-0.3, max=0.4), np.clip(x, -0.3, 0.4))
    eq(F.prelu(T(x), T(np.array([0.25], np.float32))), np.where(x > 0, x, 0.25 * x)); eq(F.log_sigmoid(T(x)), -np.log1p(np.exp(-x)))
    eq(F.maxout(T(x.reshape(2, 6, 2)[:, :, :, None]), groups=2, axis=1), x.reshape(2, 3, 2, 2, 1).max(2))


def test_linear_embedding_norms():
    w, b = rng.randn(4, 5).astype(np.float32), rng.randn(5).astype(np.float32)
    eq(F.linear(T(x), T(w), T(b)), x @ w + b)                                  # weight is [in, out]
    table = rng.randn(6, 3).astype(np.float32)
    ids = np.array([[0, 2], [5, 2]])
    out = F.embedding(T(ids), T(table), padding_idx=2)
    ref = table[ids].copy(); ref[ids == 2] = 0
    eq(out, ref)                                                               # padding rows come out as zeros
    eq(F.one_hot(T(ids), num_classes=6), np.eye(6, dtype=np.float32)[ids])
    g, be = rng.rand(4).astype(np.float32), rng.randn(4).astype(np.float32)
    mu, var = x.mean(-1, keepdims=True), x.var(-1, keepdims=True)
    eq(F.layer_norm(T(x), 4, weight=T(g), bias=T(be), epsilon=1e-5), (x - mu) / np.sqrt(var + 1e-5) * g + be)
    xb = rng.randn(4, 3, 2, 2).astype(np.float32)
    rm, rv = np.zeros(3, np.float32), np.ones(3, np.float32)
    out = F.batch_norm(T(xb), T(rm), T(rv), weight=T(np.ones(3, np.float32)), bias=T(np.zeros(3, np.float32)), training=True, momentum=0.9)
    m, v = xb.mean((0, 2, 3), keepdims=True), xb.var((0, 2, 3), keepdims=True)
    eq(out, (xb - m) / np.sqrt(v + 1e-5), 1e-4)
    xl = np.moveaxis(xb, 1, -1).copy()
    out = F.batch_norm(T(xl), T(rm), T(rv), T(np.ones(3, np.float32)), T(np.zeros(3, np.float32)), training=True, data_format="NHWC")
    eq(out, np.moveaxis((xb - m) / np.sqrt(v + 1e-5), 1, -1), 1e-4)
    eq(F.instance_norm(T(xb)), (xb - xb.mean((2, 3), keepdims=True)) / np.sqrt(xb.var((2, 3), keepdims=True) + 1e-5), 1e-4)
    eq(F.group_norm(T(xb), num_groups=1) if hasattr(F, "group_norm") else T(xb), TF.group_norm(torch.tensor(xb), 1) if hasattr(F, "group_norm") else xb, 1e-4)
    eq(F.label_smooth(T(np.eye(3, dtype=np.float32)), epsilon=0.1), np.eye(3) * 0.9 + 0.1 / 3)
    eq(F.dropout(T(x), p=0.5, training=False), x)
    eq(F.dropout(T(x), p=0.5, training=False, mode="downscale_in_infer"), x * 0.5)        # the other mode scales at inference
    d = F.dropout(T(np.ones((64, 8), np.float32)), p=0.5, axis=0).numpy()                 # whole rows dropped together
    assert set(np.unique(d)) <= {0.0, 2.0} and all(len(set(r)) == 1 for r in d)


def test_losses():
    logits = rng.randn(5, 4).astype(np.float32)
    lab = np.array([0, 3, 1, 2, 3])
    p = np.exp(logits) / np.exp(logits).sum(1, keepdims=True)
    ce = -np.log(p[np.arange(5), lab])
    eq(F.cross_entropy(T(logits), T(lab)), ce.mean()); eq(F.cross_entropy(T(logits), T(lab), reduction="none"), ce)
    eq(F.cross_entropy(T(logits), T(lab[:, None])), ce.mean())                           # labels may carry a trailing 1
    soft = np.eye(4, dtype=np.float32)[lab] * 0.9 + 0.025
    eq(F.cross_entropy(T(logits), T(soft), soft_label=True), (-(soft * np.log(p)).sum(1)).mean())
    eq(F.cross_entropy(T(p), T(lab), use_softmax=False), ce.mean(), 1e-4)                # input already probabilities
    w = np.array([1., 2., 0.5, 1.], np.float32)
    eq(F.cross_entropy(T(logits), T(lab), weight=T(w)), (ce * w[lab]).sum() / w[lab].sum())
    lab2 = lab.copy(); lab2[1] = -100
    eq(F.cross_entropy(T(logits), T(lab2), ignore_index=-100), ce[lab2 != -100].mean())
    loss, sm = F.softmax_with_cross_entropy(T(logits), T(lab[:, None]), return_softmax=True)
    eq(loss, ce[:, None]); eq(sm, p)
    eq(F.nll_loss(T(np.log(p)), T(lab)), ce.mean())
    z, y = rng.randn(6).astype(np.float32), (rng.rand(6) > 0.5).astype(np.float32)
    sg = 1 / (1 + np.exp(-z))
    eq(F.binary_cross_entropy_with_logits(T(z), T(y)), (-(y * np.log(sg) + (1 - y) * np.log(1 - sg))).mean())
    eq(F.binary_cross_entropy_with_logits(T(z), T(y), pos_weight=T([2.0] * 6)), (-(2 * y * np.log(sg) + (1 - y) * np.log(1 - sg))).mean())
    eq(F.binary_cross_entropy(T(sg), T(y), reduction="sum"), (-(y * np.log(sg) + (1 - y) * np.log(1 - sg))).sum(), 1e-4)
    a, b = rng.randn(3, 4).astype(np.float32), rng.randn(3, 4).astype(np.float32)
    eq(F.mse_loss(T(a), T(b)), ((a - b) ** 2).mean()); eq(F.l1_loss(T(a), T(b), reduction="sum"), np.abs(a - b).sum())
    d = np.abs(a - b)
    eq(F.smooth_l1_loss(T(a), T(b), delta=0.5), np.where(d < 0.5, 0.5 * d ** 2, 0.5 * (d - 0.25)).mean())   # huber form: delta * (|d| - delta / 2)
    q = np.abs(b) / np.abs(b).sum(1, keepdims=True)
    eq(F.kl_div(T(np.log(np.abs(a) / np.abs(a).sum(1, keepdims=True))), T(q), reduction="batchmean"),
       (q * (np.log(q) - np.log(np.abs(a) / np.abs(a).sum(1, keepdims=True)))).sum() / 3, 1e-4)
    eq(F.margin_ranking_loss(T(a), T(b), T(np.ones_like(a)), margin=0.1), np.maximum(0, -(a - b) + 0.1).mean())
    eq(F.cosine_similarity(T(a), T(b), axis=1), (a * b).sum(1) / np.sqrt((a ** 2).sum(1) * (b ** 2).sum(1)))
    eq(F.square_error_cost(T(a), T(b)), (a - b) ** 2); eq(F.log_loss(T(sg[:, None]), T(y[:, None])), -(y * np.log(sg + 1e-4) + (1 - y) * np.log(1 - sg + 1e-4))[:, None], 1e-4)
    fl = F.sigmoid_focal_loss(T(z), T(y), alpha=0.25, gamma=2.0, reduction="sum")
    pt = np.where(y == 1, sg, 1 - sg); at = np.where(y == 1, 0.25, 0.75)
    eq(fl, (-at * (1 - pt) ** 2 * np.log(pt)).sum(), 1e-4)


def test_conv_pool_pad_interpolate():
    img = rng.randn(1, 2, 6, 6).astype(np.float32)
    k = rng.randn(3, 2, 3, 3).astype(np.float32)
    eq(F.conv2d(T(img), T(k), padding="SAME"), TF.conv2d(torch.tensor(img), torch.tensor(k), padding=1), 1e-4)
    eq(F.conv2d(T(img), T(k), padding=[1, 2]), TF.conv2d(torch.tensor(img), torch.tensor(k), padding=(1, 2)), 1e-4)
    nhwc = np.moveaxis(img, 1, -1).copy()
    eq(F.conv2d(T(nhwc), T(k), padding=1, data_format="NHWC"), np.moveaxis(TF.conv2d(torch.tensor(img), torch.tensor(k), padding=1).numpy(), 1, -1), 1e-4)
    kt = rng.randn(2, 3, 3, 3).astype(np.float32)                                          # transpose conv weight: [in, out, kh, kw]
    eq(F.conv2d_transpose(T(img), T(kt), stride=2, padding=1, output_padding=1), TF.conv_transpose2d(torch.tensor(img), torch.tensor(kt), stride=2, padding=1, output_padding=1), 1e-4)
    eq(F.max_pool2d(T(img), kernel_size=2, stride=2), TF.max_pool2d(torch.tensor(img), 2, 2))
    out, mask = F.max_pool2d(T(img), kernel_size=2, stride=2, return_mask=True)
    ref_o, ref_i = TF.max_pool2d(torch.tensor(img), 2, 2, return_indices=True)
    eq(out, ref_o); eq(mask, ref_i)
    eq(F.avg_pool2d(T(img), kernel_size=3, stride=2, padding=1, exclusive=False), TF.avg_pool2d(torch.tensor(img), 3, 2, 1, count_include_pad=True))
    eq(F.avg_pool2d(T(img), kernel_size=3, stride=2, padding=1), TF.avg_pool2d(torch.tensor(img), 3, 2, 1, count_include_pad=False))   # exclusive=True is the default
    eq(F.adaptive_avg_pool2d(T(img), output_size=[2, 3]), TF.adaptive_avg_pool2d(torch.tensor(img), (2, 3)))
    # pad: the list runs from the LAST dim backwards when it is shorter than 2 * rank, otherwise dim by dim from the first
    eq(F.pad(T(img), [1, 2, 0, 1]), np.pad(img, ((0, 0), (0, 0), (0, 1), (1, 2))))
    eq(F.pad(T(img), [0, 0, 0, 0, 1, 0, 2, 1]), np.pad(img, ((0, 0), (0, 0), (1, 0), (2, 1))))
    eq(F.pad(T(img), [1, 1, 1, 1], mode="reflect"), np.pad(img, ((0, 0), (0, 0), (1, 1), (1, 1)), mode="reflect"))
    eq(F.pad(T(img), [1, 1, 1, 1], mode="replicate"), np.pad(img, ((0, 0), (0, 0), (1, 1), (1, 1)), mode="edge"))
    eq(F.pad(T(img), [1, 0, 0, 0], value=3.0), np.pad(img, ((0, 0), (0, 0), (0, 0), (1, 0)), constant_values=3.0))
    eq(F.interpolate(T(img), scale_factor=2, mode="nearest"), TF.interpolate(torch.tensor(img), scale_factor=2, mode="nearest"))
    eq(F.interpolate(T(img), size=[9, 9], mode="bilinear", align_corners=True), TF.interpolate(torch.tensor(img), size=(9, 9), mode="bilinear", align_corners=True), 1e-4)
    eq(F.interpolate(T(img), size=[9, 9], mode="bilinear", align_corners=False, align_mode=0), TF.interpolate(torch.tensor(img), size=(9, 9), mode="bilinear", align_corners=False), 1e-4)
    assert tuple(F.pixel_shuffle(T(rng.randn(1, 8, 2, 2).astype(np.float32)), 2).shape) == (1, 2, 4, 4)
    eq(F.unfold(T(img), kernel_sizes=[2, 2], strides=2), TF.unfold(torch.tensor(img), 2, stride=2))
    seq = rng.randn(2, 3, 8).astype(np.float32)
    eq(F.conv1d(T(seq), T(rng.randn(4, 3, 3).astype(np.float32) * 0 + 1), padding=1), TF.conv1d(torch.tensor(seq), torch.ones(4, 3, 3), padding=1), 1e-4)
    eq(F.max_pool1d(T(seq), 2), TF.max_pool1d(torch.tensor(seq), 2)); eq(F.adaptive_max_pool1d(T(seq), 4), TF.adaptive_max_pool1d(torch.tensor(seq), 4))


def test_layer_parameter_names_and_layouts_match_the_reference():
    """state_dict keys and weight layouts are the .pdparams contract (reference: python/paddle/nn/layer/{common,conv,norm,
    rnn,transformer}.py)."""
    nn = paddle.nn

    def keys(layer):
        return {k: tuple(v.shape) for k, v in layer.state_dict().items()}

    assert keys(nn.Linear(3, 5)) == {"weight": (3, 5), "bias": (5,)}                      # [in, out]
    assert keys(nn.Conv2D(4, 8, 3, groups=2)) == {"weight": (8, 2, 3, 3), "bias": (8,)}
    assert keys(nn.Conv2DTranspose(4, 8, 3)) == {"weight": (4, 8, 3, 3), "bias": (8,)}    # [in, out / groups, kh, kw]
    assert keys(nn.BatchNorm2D(4)) == {"weight": (4,), "bias": (4,), "_mean": (4,), "_variance": (4,)}
    assert keys(nn.LayerNorm(6)) == {"weight": (6,), "bias": (6,)} and keys(nn.Embedding(5, 3)) == {"weight": (5, 3)}
    lstm = keys(nn.LSTM(4, 8, num_layers=2, direction="bidirect"))
    assert lstm["weight_ih_l0"] == (32, 4) and lstm["weight_hh_l0_reverse"] == (32, 8) and lstm["weight_ih_l1"] == (32, 16)
    assert set(keys(nn.GRU(4, 8))) == {"weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0"} and keys(nn.GRU(4, 8))["weight_ih_l0"] == (24, 4)
    assert set(keys(nn.MultiHeadAttention(8, 2))) == {f"{p}_proj.{w}" for p in ("q", "k", "v", "out") for w in ("weight", "bias")}
    enc = set(keys(nn.TransformerEncoderLayer(8, 2, 16)))
    assert {"self_attn.q_proj.weight", "linear1.weight", "linear2.bias", "norm1.weight", "norm2.bias"} <= enc
    xs = torch.randn(2, 5, 4)
    out, (h, c) = nn.LSTM(4, 8, direction="bidirect")(xs)
    assert tuple(out.shape) == (2, 5, 16) and tuple(h.shape) == (2, 2, 8) == tuple(c.shape)
    out, h = nn.GRU(4, 8, time_major=True)(xs)
    assert tuple(out.shape) == (2, 5, 8) and tuple(h.shape) == (1, 5, 8)
    out, _ = nn.LSTM(4, 8)(xs, sequence_length=torch.tensor([5, 3]))
    assert float(out[1, 3:].abs().sum()) == 0.0                                            # padded steps produce zeros
    assert set(keys(nn.LSTMCell(4, 8))) == {"weight_ih", "weight_hh", "bias_ih", "bias_hh"} == set(keys(nn.GRUCell(4, 8)))
    assert set(keys(nn.Conv1DTranspose(4, 8, 3))) == {"weight", "bias"} == set(keys(nn.Bilinear(4, 8, 3)))
    import inspect
    for name in dir(nn):                                   # no layer leaks a wrapper prefix into its state dict
        cls = getattr(nn, name)
        if inspect.isclass(cls) and issubclass(cls, torch.nn.Module):
            for args in ((), (4,), (4, 8), (4, 8, 3), (8, 2), (8, 2, 16)):
                try:
                    layer = cls(*args)
                except Exception:  # noqa: BLE001  (constructor needs other arguments)
                    continue
                assert not any(k.startswith("_m.") or k.startswith("_cell.") or k.startswith("_rnn.") or "._m." in k for k in layer.state_dict()), name
                break
    m = nn.Sequential(nn.Linear(3, 4), nn.LSTM(4, 8))
    assert "1.weight_ih_l0" in m.state_dict() and len(list(m.parameters())) == 6


def test_default_initializers_follow_the_reference():
    """Linear / Embedding: Xavier uniform over [fan_in, fan_out] = weight shape; Conv: N(0, sqrt(2 / (in * kh * kw)));
    biases zero (reference: LayerHelperBase.create_parameter, nn/layer/conv.py _get_default_param_initializer)."""
    import math
    nn, I = paddle.nn, paddle.nn.initializer
    paddle.seed(0)
    w = nn.Linear(400, 600).weight.detach()
    assert abs(float(w.abs().max()) - math.sqrt(6 / 1000)) < 1e-3 and float(nn.Linear(4, 4).bias.detach().abs().sum()) == 0
    assert abs(float(nn.Embedding(1000, 64).weight.detach().abs().max()) - math.sqrt(6 / 1064)) < 1e-3
    assert abs(float(nn.Conv2D(16, 32, 3).weight.detach().std()) - math.sqrt(2 / (16 * 9))) < 5e-3
    mk = lambda shape, init: paddle.create_parameter(shape, "float32", default_initializer=init).detach()   # noqa: E731
    assert abs(float(mk([300, 500], I.KaimingNormal()).std()) - math.sqrt(2 / 300)) < 2e-3
    assert abs(float(mk([300, 500], I.KaimingUniform()).abs().max()) - math.sqrt(6 / 300)) < 1e-3
    assert abs(float(mk([64, 32, 3, 3], I.XavierNormal()).std()) - math.sqrt(2 / (32 * 9 + 64 * 9))) < 2e-3
    assert float(mk([300, 500], I.TruncatedNormal(0.0, 1.0)).abs().max()) <= 2.0 + 1e-6
    t = mk([6, 6], I.Orthogonal())
    assert float((t @ t.t() - torch.eye(6)).abs().max()) < 1e-5
    assert float((mk([4, 4], I.Assign(torch.eye(4))) - torch.eye(4)).abs().sum()) == 0
    assert float(mk([3], I.Constant(2.5)).sum()) == 7.5 and abs(float(mk([2000], I.Uniform(-0.5, 0.5)).abs().max()) - 0.5) < 5e-3


def test_weight_norm_parameter_layout_and_copy():
    """Reference: python/paddle/nn/utils/weight_norm_hook.py — ``weight_g`` is 1-D (``shape[dim]`` entries, one for
    ``dim=None``), ``weight_v`` keeps the weight's shape, the weight is rebuilt before every forward."""
    import copy
    import paddle.nn as nn
    from paddle.nn.utils import weight_norm, remove_weight_norm
    lin = nn.Linear(4, 3)
    w0 = lin.weight.detach().clone()
    weight_norm(lin, dim=1)
    assert sorted(n for n, _ in lin.named_parameters()) == ["bias", "weight_g", "weight_v"]
    assert tuple(lin.weight_g.shape) == (3,) and tuple(lin.weight_v.shape) == (4, 3)
    torch.testing.assert_close(lin.weight_g.detach(), w0.norm(dim=0))
    x = torch.randn(5, 4)
    y = lin(x)
    torch.testing.assert_close(y, x @ w0 + lin.bias)
    y.sum().backward()
    assert lin.weight_g.grad is not None and lin.weight_v.grad is not None
    with torch.no_grad():
        lin.weight_g.mul_(2.0)
    torch.testing.assert_close(lin(x), x @ (2 * w0) + lin.bias)
    assert sorted(lin.state_dict()) == ["bias", "weight_g", "weight_v"]
    remove_weight_norm(lin)
    assert sorted(n for n, _ in lin.named_parameters()) == ["bias", "weight"]
    torch.testing.assert_close(lin.weight.detach(), 2 * w0)
    conv = nn.Conv2D(3, 6, 3)
    weight_norm(conv)                                   # dim=0
    assert tuple(conv.weight_g.shape) == (6,)
    whole = nn.Conv2D(3, 6, 3)
    k0 = whole.weight.detach().clone()
    weight_norm(whole, dim=None)
    assert tuple(whole.weight_g.shape) == (1,)
    torch.testing.assert_close(whole.weight_g.detach(), k0.norm().reshape(1))
    img = torch.randn(1, 3, 5, 5)
    twin = copy.deepcopy(conv)
    torch.testing.assert_close(twin(img), conv(img))
    with pytest.raises(ValueError):
        remove_weight_norm(nn.Linear(2, 2))
