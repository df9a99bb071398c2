"""Convolution / pooling / padding / resampling / normalisation argument algebra against restatements of the reference's
documented rules (python/paddle/nn/functional/{conv,pooling,common,norm}.py; paddle/phi/kernels/cpu/conv_util.h:23-72;
paddle/phi/kernels/cpu/batch_norm_kernel.cc:100-157)."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as TF

import paddle
import paddle.nn as nn
import paddle.nn.functional as F


def _conv2d_naive(x, w, stride, pads, dilation):
    """Direct convolution in numpy: x [N, C, H, W], w [O, C, kh, kw], pads ((top, bottom), (left, right))."""
    x = np.pad(x, ((0, 0), (0, 0), pads[0], pads[1]))
    n, c, h, wd = x.shape
    o, _, kh, kw = w.shape
    oh = (h - (dilation[0] * (kh - 1) + 1)) // stride[0] + 1
    ow = (wd - (dilation[1] * (kw - 1) + 1)) // stride[1] + 1
    out = np.zeros((n, o, oh, ow), dtype=np.float64)
    for i in range(oh):
        for j in range(ow):
            patch = x[:, :, i * stride[0]: i * stride[0] + dilation[0] * (kh - 1) + 1: dilation[0],
                      j * stride[1]: j * stride[1] + dilation[1] * (kw - 1) + 1: dilation[1]]
            out[:, :, i, j] = np.einsum("nchw,ochw->no", patch, w)
    return out


def test_conv_padding_forms():
    rng = np.random.RandomState(0)
    x = rng.randn(2, 3, 7, 8)
    w = rng.randn(4, 3, 3, 2)
    xt, wt = torch.from_numpy(x), torch.from_numpy(w)
    cases = [
        (0, ((0, 0), (0, 0))),
        (1, ((1, 1), (1, 1))),
        ([1, 2], ((1, 1), (2, 2))),
        ([1, 0, 2, 1], ((1, 0), (2, 1))),                                  # (top, bottom, left, right)
        ([[0, 0], [0, 0], [2, 1], [0, 3]], ((2, 1), (0, 3))),              # one pair per tensor dim
        ("VALID", ((0, 0), (0, 0))),
    ]
    for padding, pads in cases:
        got = F.conv2d(xt, wt, stride=[2, 1], padding=padding, dilation=[1, 2]).numpy()
        np.testing.assert_allclose(got, _conv2d_naive(x, w, (2, 1), pads, (1, 2)), atol=1e-10, err_msg=str(padding))
    # SAME: ceil(in / stride) outputs, the odd pixel goes after, dilation is reset to 1
    got = F.conv2d(xt, wt, stride=2, padding="SAME", dilation=3).numpy()
    assert got.shape == (2, 4, 4, 4)
    # H: out 4 -> total = 3*2+3-7 = 2 -> (1, 1); W: out 4 -> total = 3*2+2-8 = 0
    np.testing.assert_allclose(got, _conv2d_naive(x, w, (2, 2), ((1, 1), (0, 0)), (1, 1)), atol=1e-10)
    got = F.conv2d(xt, wt, stride=1, padding="same").numpy()                # W: total 1 -> (0, 1)
    np.testing.assert_allclose(got, _conv2d_naive(x, w, (1, 1), ((1, 1), (0, 1)), (1, 1)), atol=1e-10)
    # channel-last input, same weights
    got = F.conv2d(xt.permute(0, 2, 3, 1), wt, padding=[1, 0, 2, 1], data_format="NHWC")
    np.testing.assert_allclose(got.permute(0, 3, 1, 2).numpy(), _conv2d_naive(x, w, (1, 1), ((1, 0), (2, 1)), (1, 1)), atol=1e-10)
    with pytest.raises(ValueError):
        F.conv2d(xt, wt, padding=[[0, 0], [1, 1], [1, 1], [1, 1]])          # padding on the channel dim
    with pytest.raises(ValueError):
        F.conv2d(xt, wt, padding="FULL")
    with pytest.raises(ValueError):
        F.conv2d(xt, wt, padding=-1)
    with pytest.raises(ValueError):
        F.conv2d(xt, wt, data_format="NCWH")
    # 1-D and 3-D share the algebra
    x1, w1 = torch.randn(2, 3, 9), torch.randn(5, 3, 4)
    torch.testing.assert_close(F.conv1d(x1, w1, padding=[2, 1]), TF.conv1d(TF.pad(x1, (2, 1)), w1))
    torch.testing.assert_close(F.conv1d(x1.transpose(1, 2), w1, padding="SAME", stride=2, data_format="NLC"),
                               TF.conv1d(TF.pad(x1, (1, 2)), w1, stride=2).transpose(1, 2))
    x3, w3 = torch.randn(1, 2, 5, 6, 7), torch.randn(4, 1, 2, 3, 2)
    torch.testing.assert_close(F.conv3d(x3, w3, groups=2, padding=[1, 0, 0, 2, 1, 1]),
                               TF.conv3d(TF.pad(x3, (1, 1, 0, 2, 1, 0)), w3, groups=2))
    torch.testing.assert_close(F.conv3d(x3.permute(0, 2, 3, 4, 1), w3, groups=2, padding=1, data_format="NDHWC"),
                               TF.conv3d(x3, w3, groups=2, padding=1).permute(0, 2, 3, 4, 1))


def test_conv_transpose_output_size_and_padding():
    x, w = torch.randn(2, 3, 5, 6), torch.randn(3, 4, 3, 3)
    base = F.conv2d_transpose(x, w, stride=2, padding=1)
    assert base.shape == (2, 4, 9, 11)                                      # (in - 1) * 2 - 2 + 3
    for size in ([9, 11], [10, 12], [10, 11]):
        y = F.conv2d_transpose(x, w, stride=2, padding=1, output_size=size)
        assert list(y.shape[2:]) == size
        torch.testing.assert_close(y[:, :, :9, :11], base)                  # the extra row / column is appended
    torch.testing.assert_close(F.conv2d_transpose(x, w, stride=2, padding=1, output_padding=1),
                               F.conv2d_transpose(x, w, stride=2, padding=1, output_size=[10, 12]))
    with pytest.raises(ValueError):
        F.conv2d_transpose(x, w, stride=2, padding=1, output_size=[11, 12])
    with pytest.raises(ValueError):
        F.conv2d_transpose(x, w, stride=2, padding=1, output_size=[10, 12], output_padding=1)
    # it is the adjoint of conv2d with the same padding, asymmetric pads included
    for padding in (1, [1, 0, 2, 1], "VALID"):
        y = F.conv2d_transpose(x, w, stride=2, padding=padding)
        z = torch.randn_like(y)
        lhs = (y * z).sum()
        rhs = (x * F.conv2d(z, w, stride=2, padding=padding)[:, :, :5, :6]).sum()
        torch.testing.assert_close(lhs, rhs, rtol=1e-4, atol=1e-3)
    y = F.conv2d_transpose(x.permute(0, 2, 3, 1), w, stride=2, padding=1, data_format="NHWC")
    torch.testing.assert_close(y.permute(0, 3, 1, 2), base)
    x1, w1 = torch.randn(2, 3, 5), torch.randn(3, 2, 4)
    assert F.conv1d_transpose(x1, w1, stride=3, output_size=[18]).shape == (2, 2, 18)
    x3, w3 = torch.randn(1, 2, 3, 4, 5), torch.randn(2, 3, 2, 2, 2)
    assert F.conv3d_transpose(x3, w3, stride=2, output_size=[7, 8, 11]).shape == (1, 3, 7, 8, 11)
    # layers: signature order, output_size at call time, weight layout
    layer = nn.Conv2DTranspose(3, 4, 3, 2, 1, 0, 1, 1)                      # ..., output_padding, dilation, groups
    assert layer.weight.shape == (3, 4, 3, 3) and layer(x, output_size=[10, 12]).shape == (2, 4, 10, 12)
    layer = nn.Conv1DTranspose(3, 4, 3, 2, 1, 0, 1, 2)                      # ..., output_padding, groups, dilation
    assert layer._groups == 1 and layer._dilation == (2,)
    layer = nn.Conv3DTranspose(2, 4, 2, stride=2, groups=2)
    assert layer.weight.shape == (2, 2, 2, 2, 2) and layer(x3).shape == (1, 4, 6, 8, 10)


def test_conv_layers_padding_mode_and_format():
    x = torch.randn(2, 3, 6, 7)
    for mode, tmode in (("reflect", "reflect"), ("replicate", "replicate"), ("circular", "circular")):
        c = nn.Conv2D(3, 5, 3, padding=[1, 2], padding_mode=mode)
        want = TF.conv2d(TF.pad(x, (2, 2, 1, 1), mode=tmode), c.weight, c.bias)
        torch.testing.assert_close(c(x), want)
    c = nn.Conv2D(3, 5, 3, padding=1, data_format="NHWC")
    assert c.weight.shape == (5, 3, 3, 3)
    torch.testing.assert_close(c(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2), TF.conv2d(x, c.weight, c.bias, padding=1))
    c = nn.Conv1D(3, 4, 3, padding=2, padding_mode="reflect", data_format="NLC")
    x1 = torch.randn(2, 9, 3)
    want = TF.conv1d(TF.pad(x1.transpose(1, 2), (2, 2), mode="reflect"), c.weight, c.bias).transpose(1, 2)
    torch.testing.assert_close(c(x1), want)
    c = nn.Conv2D(4, 6, 3, groups=2, bias_attr=False)
    assert c.weight.shape == (6, 2, 3, 3) and c.bias is None and sorted(c.state_dict()) == ["weight"]
    # default initializer: Normal(0, sqrt(2 / (kh * kw * in_channels)))
    paddle.seed(0)
    c = nn.Conv2D(64, 128, 3)
    assert abs(float(c.weight.detach().std()) - math.sqrt(2.0 / (9 * 64))) < 2e-3 and float(c.bias.abs().max()) == 0.0
    for bad in (dict(padding_mode="mirror"), dict(groups=2), dict(data_format="NCWH")):
        with pytest.raises(ValueError):
            nn.Conv2D(3, 4, 3, **bad)


def _pool2d_naive(x, k, s, pads, kind, exclusive=True, ceil_mode=False):
    n, c, h, w = x.shape
    (pt, pb), (pl, pr) = pads

    def osize(n_in, kk, ss, p0, p1):
        v = (n_in + p0 + p1 - kk) / ss
        return int(math.ceil(v) if ceil_mode else math.floor(v)) + 1
    oh, ow = osize(h, k[0], s[0], pt, pb), osize(w, k[1], s[1], pl, pr)
    out = np.zeros((n, c, oh, ow))
    for i in range(oh):
        for j in range(ow):
            h0, w0 = i * s[0] - pt, j * s[1] - pl
            h1, w1 = min(h0 + k[0], h + pb), min(w0 + k[1], w + pr)
            hh0, ww0, hh1, ww1 = max(h0, 0), max(w0, 0), min(h1, h), min(w1, w)
            win = x[:, :, hh0:hh1, ww0:ww1]
            if kind == "max":
                out[:, :, i, j] = win.max((2, 3))
            else:
                div = (hh1 - hh0) * (ww1 - ww0) if exclusive else (h1 - h0) * (w1 - w0)
                out[:, :, i, j] = win.sum((2, 3)) / div
    return out


def test_pooling_padding_forms_and_flags():
    rng = np.random.RandomState(1)
    x = rng.randn(2, 3, 7, 9)
    xt = torch.from_numpy(x)
    for padding, pads in [(0, ((0, 0), (0, 0))), (1, ((1, 1), (1, 1))), ([1, 0], ((1, 1), (0, 0))),
                          ([1, 0, 0, 1], ((1, 0), (0, 1))), ([[0, 0], [0, 0], [0, 1], [1, 1]], ((0, 1), (1, 1)))]:
        for ceil in (False, True):
            got = F.max_pool2d(xt, 3, 2, padding, ceil_mode=ceil).numpy()
            np.testing.assert_allclose(got, _pool2d_naive(x, (3, 3), (2, 2), pads, "max", ceil_mode=ceil), err_msg=f"{padding} {ceil}")
            for excl in (True, False):
                got = F.avg_pool2d(xt, 3, 2, padding, ceil_mode=ceil, exclusive=excl).numpy()
                want = _pool2d_naive(x, (3, 3), (2, 2), pads, "avg", exclusive=excl, ceil_mode=ceil)
                np.testing.assert_allclose(got, want, atol=1e-12, err_msg=f"{padding} ceil={ceil} excl={excl}")
    # SAME: ceil(in / stride) outputs; VALID: no padding; stride defaults to the kernel
    assert F.max_pool2d(xt, 3, 2, "SAME").shape == (2, 3, 4, 5)
    np.testing.assert_allclose(F.avg_pool2d(xt, 3, 2, "SAME").numpy(),
                               _pool2d_naive(x, (3, 3), (2, 2), ((1, 1), (1, 1)), "avg"), atol=1e-12)
    np.testing.assert_allclose(F.max_pool2d(xt, [2, 3], 2, "SAME").numpy(),           # H total 1 -> (0, 1); W total 2
                               _pool2d_naive(x, (2, 3), (2, 2), ((0, 1), (1, 1)), "max"))
    assert F.avg_pool2d(xt, 2).shape == (2, 3, 3, 4) and F.max_pool2d(xt, 3, padding="VALID").shape == (2, 3, 2, 3)
    # channel-last, divisor_override, masks
    torch.testing.assert_close(F.avg_pool2d(xt.permute(0, 2, 3, 1), 3, 2, 1, data_format="NHWC").permute(0, 3, 1, 2),
                               F.avg_pool2d(xt, 3, 2, 1))
    torch.testing.assert_close(F.avg_pool2d(xt, 2, divisor_override=3), TF.avg_pool2d(xt, 2) * 4 / 3)
    y, mask = F.max_pool2d(xt, 2, return_mask=True)
    torch.testing.assert_close(xt.flatten(2).gather(2, mask.flatten(2)).reshape(y.shape), y)
    y, mask = F.max_pool2d(xt, 3, 2, [1, 0, 0, 1], return_mask=True)                 # asymmetric pads: input indices
    assert int(mask.max()) < 7 * 9 and int(mask.min()) >= 0
    torch.testing.assert_close(xt.flatten(2).gather(2, mask.flatten(2)).reshape(y.shape), y)
    with pytest.raises(ValueError):
        F.max_pool2d(xt.permute(0, 2, 3, 1), 2, return_mask=True, data_format="NHWC")
    # unpool puts the maxima back; output_size picks among the sizes that pool to the same shape
    ye, me = F.max_pool2d(xt[:, :, :6, :8], 2, return_mask=True)
    up = F.max_unpool2d(ye, me, 2)
    assert up.shape == (2, 3, 6, 8)
    torch.testing.assert_close(up, TF.max_unpool2d(ye, me, 2))
    y, mask = F.max_pool2d(xt, 2, return_mask=True)
    up = F.max_unpool2d(y, mask, 2, output_size=[7, 9])
    assert up.shape == (2, 3, 7, 9) and int((up != 0).sum()) == y.numel()
    torch.testing.assert_close(up, TF.max_unpool2d(y, mask, 2, output_size=[7, 9]))
    assert nn.MaxUnPool2D(2, output_size=[7, 9])(y, mask).shape == (2, 3, 7, 9)
    with pytest.raises(ValueError):
        F.max_unpool2d(y, mask, 2, output_size=[12, 9])
    # 1-D / 3-D / lp / adaptive
    x1 = torch.randn(2, 3, 10)
    torch.testing.assert_close(F.avg_pool1d(x1, 3, 2, [1, 0], exclusive=True),
                               torch.stack([x1[..., max(i * 2 - 1, 0):i * 2 + 2].mean(-1) for i in range(5)], -1))
    x3 = torch.randn(1, 2, 4, 5, 6)
    assert F.max_pool3d(x3, 2, padding="SAME").shape == (1, 2, 2, 3, 3)
    torch.testing.assert_close(F.avg_pool3d(x3.permute(0, 2, 3, 4, 1), 2, data_format="NDHWC").permute(0, 4, 1, 2, 3), F.avg_pool3d(x3, 2))
    torch.testing.assert_close(F.lp_pool2d(xt.abs(), 2, 2), (TF.avg_pool2d(xt.abs() ** 2, 2) * 4).sqrt())
    torch.testing.assert_close(F.lp_pool1d(x1.abs(), float("inf"), 2), F.max_pool1d(x1.abs(), 2))
    assert F.adaptive_avg_pool2d(xt, [None, 4]).shape == (2, 3, 7, 4)
    assert F.adaptive_avg_pool2d(xt.permute(0, 2, 3, 1), 2, data_format="NHWC").shape == (2, 2, 2, 3)
    assert F.adaptive_avg_pool3d(x3, [2, None, 3]).shape == (1, 2, 2, 5, 3)
    y, m = F.adaptive_max_pool2d(xt, 3, return_mask=True)
    torch.testing.assert_close(xt.flatten(2).gather(2, m.flatten(2)).reshape(y.shape), y)
    # layers carry every option through
    assert nn.AvgPool2D(3, 2, 1, ceil_mode=True, exclusive=False)(xt).shape == F.avg_pool2d(xt, 3, 2, 1, True, False).shape
    assert nn.MaxPool3D(2, data_format="NDHWC")(x3.permute(0, 2, 3, 4, 1)).shape == (1, 2, 2, 3, 2)
    assert nn.AdaptiveAvgPool2D((None, 2), data_format="NHWC")(xt.permute(0, 2, 3, 1)).shape == (2, 7, 2, 3)
    assert nn.LPPool2D(2, 2)(xt.abs()).shape == (2, 3, 3, 4)


def test_pad_forms():
    x = torch.arange(2 * 3 * 4 * 5, dtype=torch.float32).reshape(2, 3, 4, 5)
    xn = x.numpy()
    # spatial form: (left, right, top, bottom)
    for mode, npmode in (("constant", "constant"), ("reflect", "reflect"), ("replicate", "edge"), ("circular", "wrap")):
        kw = dict(constant_values=1.5) if mode == "constant" else {}
        want = np.pad(xn, ((0, 0), (0, 0), (2, 0), (1, 3)), mode=npmode, **kw)
        np.testing.assert_array_equal(F.pad(x, [1, 3, 2, 0], mode=mode, value=1.5).numpy(), want)
        got = F.pad(x.permute(0, 2, 3, 1), [1, 3, 2, 0], mode=mode, value=1.5, data_format="NHWC")
        np.testing.assert_array_equal(got.permute(0, 3, 1, 2).numpy(), want)
        np.testing.assert_array_equal(nn.Pad2D([1, 3, 2, 0], mode=mode, value=1.5)(x).numpy(), want)
    # every-dim form (constant only): pairs from the FIRST dim
    want = np.pad(xn, ((1, 0), (0, 2), (3, 0), (0, 1)))
    np.testing.assert_array_equal(F.pad(x, [1, 0, 0, 2, 3, 0, 0, 1]).numpy(), want)
    np.testing.assert_array_equal(F.pad(x, [0, 1, 3, 0, 0, 2, 1, 0], pad_from_left_axis=False).numpy(), want)
    np.testing.assert_array_equal(F.pad(torch.zeros(2, 3), [1, 2, 3, 4]).shape, (5, 10))
    # 1-D / 3-D layers
    x1 = torch.randn(2, 3, 6)
    torch.testing.assert_close(nn.Pad1D([1, 2], mode="reflect")(x1), TF.pad(x1, (1, 2), mode="reflect"))
    torch.testing.assert_close(nn.Pad1D(2, data_format="NLC")(x1), TF.pad(x1, (0, 0, 2, 2)))
    x3 = torch.randn(1, 2, 3, 4, 5)
    torch.testing.assert_close(nn.Pad3D([1, 0, 0, 2, 1, 1], mode="replicate")(x3), TF.pad(x3, (1, 0, 0, 2, 1, 1), mode="replicate"))
    assert nn.ZeroPad2D([1, 2, 3, 4])(x).shape == (2, 3, 11, 8) and nn.ZeroPad1D(1)(x1).shape == (2, 3, 8)
    assert nn.ZeroPad3D(1, data_format="NDHWC")(x3).shape == (1, 4, 5, 6, 5)
    assert F.zeropad2d(x, [1, 1, 0, 0]).shape == (2, 3, 4, 7)
    with pytest.raises(ValueError):
        F.pad(x, [1, 1], mode="reflect")
    with pytest.raises(ValueError):
        F.pad(x, [1, 1, 1, 1], mode="mirror")


def _interp1d(v, out, align_corners, align_mode, scale=None):
    n = len(v)
    res = np.zeros(out)
    for d in range(out):
        if align_corners:
            src = d * (n - 1) / (out - 1) if out > 1 else 0.0
        else:
            ratio = (1.0 / scale) if scale else n / out
            src = (d + 0.5) * ratio - 0.5 if align_mode == 0 else d * ratio
            src = max(src, 0.0)
        i0 = min(int(math.floor(src)), n - 1)
        i1 = min(i0 + 1, n - 1)
        lam = src - i0
        res[d] = v[i0] * (1 - lam) + v[i1] * lam
    return res


def test_interpolate_modes():
    v = np.array([1.0, 4.0, 2.0, 8.0, 5.0])
    x = torch.tensor(v, dtype=torch.float64).reshape(1, 1, 5)
    for ac, am in ((True, 0), (False, 0), (False, 1)):
        for out in (3, 5, 8, 11):
            got = F.interpolate(x, size=[out], mode="linear", align_corners=ac, align_mode=am)
            np.testing.assert_allclose(got.numpy()[0, 0], _interp1d(v, out, ac, am), atol=1e-12, err_msg=f"{ac} {am} {out}")
    # a given scale_factor (not out / in) drives the coordinates
    got = F.interpolate(x, scale_factor=1.7, mode="linear", align_mode=1)
    np.testing.assert_allclose(got.numpy()[0, 0], _interp1d(v, 8, False, 1, scale=1.7), atol=1e-12)
    got = F.interpolate(x, scale_factor=1.7, mode="linear")
    np.testing.assert_allclose(got.numpy()[0, 0], _interp1d(v, 8, False, 0, scale=1.7), atol=1e-12)
    # bilinear is separable; align_mode=1 differs from align_mode=0
    img = torch.randn(2, 3, 4, 5, dtype=torch.float64)
    a0 = F.interpolate(img, size=[7, 9], mode="bilinear", align_mode=0)
    a1 = F.interpolate(img, size=[7, 9], mode="bilinear", align_mode=1)
    assert not torch.allclose(a0, a1)
    row = _interp1d(img[0, 0, :, 2].numpy(), 7, False, 1)
    col_then = np.stack([_interp1d(img[0, 0, r].numpy(), 9, False, 1) for r in range(4)])
    want = np.stack([_interp1d(col_then[:, c], 7, False, 1) for c in range(9)], 1)
    np.testing.assert_allclose(a1[0, 0].numpy(), want, atol=1e-12)
    assert row.shape == (7,)
    torch.testing.assert_close(a0, TF.interpolate(img, size=(7, 9), mode="bilinear", align_corners=False))
    # nearest: floor(dst * in / out); area = adaptive average; channel-last; trilinear; bicubic
    got = F.interpolate(x, size=[8], mode="nearest")
    np.testing.assert_array_equal(got.numpy()[0, 0], v[[int(math.floor(d * 5 / 8)) for d in range(8)]])
    torch.testing.assert_close(F.interpolate(img, size=[2, 2], mode="area"), TF.adaptive_avg_pool2d(img, 2))
    torch.testing.assert_close(F.interpolate(img.permute(0, 2, 3, 1), scale_factor=2, mode="bilinear", data_format="NHWC"),
                               TF.interpolate(img, scale_factor=2, mode="bilinear").permute(0, 2, 3, 1))
    vol = torch.randn(1, 2, 3, 4, 5)
    assert F.interpolate(vol, scale_factor=[2, 1.5, 1], mode="trilinear").shape == (1, 2, 6, 6, 5)
    assert F.interpolate(vol, size=[4, 4, 4], mode="nearest", data_format="NCDHW").shape == (1, 2, 4, 4, 4)
    torch.testing.assert_close(F.interpolate(img, size=[8, 8], mode="bicubic", align_corners=True),
                               TF.interpolate(img, size=(8, 8), mode="bicubic", align_corners=True))
    assert F.interpolate(img, size=torch.tensor([6, 7]), mode="nearest").shape == (2, 3, 6, 7)
    assert F.upsample(img, scale_factor=2).shape == (2, 3, 8, 10)
    # layers
    up = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True)
    torch.testing.assert_close(up(img), TF.interpolate(img, scale_factor=2, mode="bilinear", align_corners=True))
    torch.testing.assert_close(nn.UpsamplingBilinear2D(size=[6, 6])(img), TF.interpolate(img, size=(6, 6), mode="bilinear", align_corners=True))
    torch.testing.assert_close(nn.UpsamplingNearest2D(scale_factor=2)(img), TF.interpolate(img, scale_factor=2, mode="nearest"))
    assert nn.Upsample(size=[3, 3], mode="bilinear", align_mode=1, data_format="NHWC")(img.permute(0, 2, 3, 1)).shape == (2, 3, 3, 3)
    for bad in (dict(size=[4, 4], scale_factor=2), dict(), dict(size=[4, 4], mode="linear"), dict(size=[4, 4], mode="nearest", align_corners=True),
                dict(size=[4], mode="bilinear"), dict(scale_factor=-1.0), dict(size=[4, 4], mode="cubic")):
        with pytest.raises(ValueError):
            F.interpolate(img, **bad)


def test_norm_layers_statistics_and_formats():
    x = torch.randn(4, 3, 5, 6) * 2 + 1
    bn = nn.BatchNorm2D(3, momentum=0.8)
    bn.train()
    y = bn(x)
    mean, var = x.mean((0, 2, 3)), x.var((0, 2, 3), unbiased=False)
    torch.testing.assert_close(y, (x - mean[None, :, None, None]) / torch.sqrt(var[None, :, None, None] + 1e-5), rtol=1e-4, atol=1e-5)
    # running = momentum * running + (1 - momentum) * batch, with the biased batch variance
    torch.testing.assert_close(bn._mean, 0.2 * mean)
    torch.testing.assert_close(bn._variance, 0.8 + 0.2 * var)
    bn.eval()
    y = bn(x)
    torch.testing.assert_close(y, (x - bn._mean[None, :, None, None]) / torch.sqrt(bn._variance[None, :, None, None] + 1e-5), rtol=1e-4, atol=1e-5)
    assert sorted(bn.state_dict()) == ["_mean", "_variance", "bias", "weight"]
    # use_global_stats=True normalises with the running statistics even in training, and leaves them alone
    g = nn.BatchNorm2D(3, use_global_stats=True)
    g.train()
    torch.testing.assert_close(g(x), x / math.sqrt(1 + 1e-5))
    assert float(g._mean.abs().max()) == 0.0
    # channel-last and other ranks
    cl = nn.BatchNorm2D(3, data_format="NHWC")
    cl.train()
    torch.testing.assert_close(cl(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2), nn.BatchNorm2D(3).train()(x))
    torch.testing.assert_close(cl._variance, 0.9 + 0.1 * var)
    b1 = nn.BatchNorm1D(3)
    assert b1(torch.randn(8, 3)).shape == (8, 3) and b1(torch.randn(8, 3, 7)).shape == (8, 3, 7)
    assert nn.BatchNorm1D(3, data_format="NLC")(torch.randn(8, 7, 3)).shape == (8, 7, 3)
    assert nn.BatchNorm3D(3, data_format="NDHWC")(torch.randn(2, 2, 3, 4, 3)).shape == (2, 2, 3, 4, 3)
    # instance norm: per sample and channel, parameters named scale / bias, no running statistics
    inn = nn.InstanceNorm2D(3)
    assert sorted(inn.state_dict()) == ["bias", "scale"]
    m, v = x.mean((2, 3), keepdim=True), x.var((2, 3), unbiased=False, keepdim=True)
    torch.testing.assert_close(inn(x), (x - m) / torch.sqrt(v + 1e-5), rtol=1e-4, atol=1e-5)
    plain = nn.InstanceNorm1D(3, weight_attr=False, bias_attr=False)
    assert plain.scale is None and len(plain.state_dict()) == 0 and plain(torch.randn(2, 3, 9)).shape == (2, 3, 9)
    assert nn.InstanceNorm3D(2)(torch.randn(2, 2, 3, 4, 5)).shape == (2, 2, 3, 4, 5)
    with pytest.raises(ValueError):
        inn(torch.randn(2, 3, 4))
    # group norm: statistics over (channels of the group, spatial); channel-last
    gn = nn.GroupNorm(3, 6)
    xg = torch.randn(2, 6, 4, 5)
    r = xg.reshape(2, 3, -1)
    want = ((r - r.mean(-1, keepdim=True)) / torch.sqrt(r.var(-1, unbiased=False, keepdim=True) + 1e-5)).reshape(xg.shape)
    torch.testing.assert_close(gn(xg), want, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(nn.GroupNorm(3, 6, data_format="NHWC")(xg.permute(0, 2, 3, 1)).permute(0, 3, 1, 2), want, rtol=1e-4, atol=1e-5)
    with pytest.raises(ValueError):
        nn.GroupNorm(4, 6)
    # local response norm over neighbouring channels
    xl = torch.rand(1, 5, 2, 2)
    got = nn.LocalResponseNorm(3, alpha=1e-2, beta=0.5, k=2.0)(xl)
    sq = TF.pad(xl ** 2, (0, 0, 0, 0, 1, 1))
    div = (2.0 + 1e-2 / 3 * (sq[:, 0:5] + sq[:, 1:6] + sq[:, 2:7])) ** 0.5
    torch.testing.assert_close(got, xl / div)
    # shuffles in both layouts
    xs = torch.randn(2, 8, 3, 4)
    torch.testing.assert_close(nn.PixelShuffle(2, "NHWC")(xs.permute(0, 2, 3, 1)).permute(0, 3, 1, 2), TF.pixel_shuffle(xs, 2))
    torch.testing.assert_close(nn.PixelUnshuffle(2)(TF.pixel_shuffle(xs, 2)), xs)
    torch.testing.assert_close(nn.ChannelShuffle(4, "NHWC")(xs.permute(0, 2, 3, 1)).permute(0, 3, 1, 2), TF.channel_shuffle(xs, 4))
    # unfold / fold with (top, left, bottom, right) paddings
    xu = torch.randn(2, 3, 6, 7)
    torch.testing.assert_close(F.unfold(xu, [2, 3], 2, [1, 0, 0, 2], 1), TF.unfold(TF.pad(xu, (0, 2, 1, 0)), (2, 3), 1, 0, 2))
    cols = F.unfold(xu, 3, 1, 1, 1)
    torch.testing.assert_close(nn.Fold([6, 7], 3, paddings=1)(cols), TF.fold(cols, (6, 7), 3, 1, 1, 1))
    assert nn.Unfold(3, paddings=[1, 1])(xu).shape == (2, 27, 42)
