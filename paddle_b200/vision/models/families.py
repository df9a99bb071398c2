"""The remaining classification families of paddle.vision.models (reference: python/paddle/vision/models/
{mobilenetv1,mobilenetv3,densenet,squeezenet,googlenet,inceptionv3,shufflenetv2}.py).

Sub-layer names follow the reference so its ``.pdparams`` checkpoints load by key; the networks themselves are built
from small tables (one row per block / branch) instead of being spelled out layer by layer.  Convolutions and batch
norms run through the library kernels - these models are not on the hot path this framework is about."""
from __future__ import annotations

import math
from functools import partial

import torch

from paddle_b200 import ParamAttr, nn
from paddle_b200.nn import initializer as I
from paddle_b200.vision.ops import ConvNormActivation


def _no_pretrained(pretrained, arch):
    if pretrained:
        raise ValueError(f"{arch}: pretrained weights cannot be downloaded in this environment; load a local .pdparams with "
                         f"model.set_state_dict(paddle.load(path))")


def _make_divisible(v, divisor=8, min_value=None):
    """Round to a multiple of ``divisor`` without going more than 10 % below ``v``."""
    min_value = divisor if min_value is None else min_value
    new_v = max(min_value, int(v + divisor / 2) // divisor * divisor)
    return new_v + divisor if new_v < 0.9 * v else new_v


class _Head:
    """with_pool / num_classes handling shared by the table-built models."""

    def _classify(self, x, pool, fc):
        if self.with_pool:
            x = pool(x)
        if self.num_classes > 0:
            x = fc(torch.flatten(x, 1))
        return x


# --------------------------------------------------------------------------------------------------------- MobileNetV1
class DepthwiseSeparable(nn.Layer):
    def __init__(self, in_channels, out_channels1, out_channels2, num_groups, stride, scale):
        super().__init__()
        mid = int(out_channels1 * scale)
        self._depthwise_conv = ConvNormActivation(in_channels, mid, 3, stride=stride, padding=1, groups=int(num_groups * scale))
        self._pointwise_conv = ConvNormActivation(mid, int(out_channels2 * scale), 1, stride=1, padding=0)

    def forward(self, x):
        return self._pointwise_conv(self._depthwise_conv(x))


class MobileNetV1(nn.Layer, _Head):
    # name, input, depthwise width, pointwise width, stride
    _TABLE = ([("conv2_1", 32, 32, 64, 1), ("conv2_2", 64, 64, 128, 2), ("conv3_1", 128, 128, 128, 1), ("conv3_2", 128, 128, 256, 2),
               ("conv4_1", 256, 256, 256, 1), ("conv4_2", 256, 256, 512, 2)]
              + [(f"conv5_{i}", 512, 512, 512, 1) for i in range(1, 6)]
              + [("conv5_6", 512, 512, 1024, 2), ("conv6", 1024, 1024, 1024, 1)])

    def __init__(self, scale=1.0, num_classes=1000, with_pool=True):
        super().__init__()
        self.scale, self.num_classes, self.with_pool = scale, num_classes, with_pool
        self.conv1 = ConvNormActivation(3, int(32 * scale), 3, stride=2, padding=1)
        self.dwsl = []
        for name, cin, c1, c2, stride in self._TABLE:
            block = DepthwiseSeparable(int(cin * scale), c1, c2, c1, stride, scale)
            self.add_sublayer(name, block)
            self.dwsl.append(block)
        if with_pool:
            self.pool2d_avg = nn.AdaptiveAvgPool2D(1)
        if num_classes > 0:
            self.fc = nn.Linear(int(1024 * scale), num_classes)

    def forward(self, x):
        x = self.conv1(x)
        for block in self.dwsl:
            x = block(x)
        return self._classify(x, getattr(self, "pool2d_avg", None), getattr(self, "fc", None))


def mobilenet_v1(pretrained=False, scale=1.0, **kwargs):
    _no_pretrained(pretrained, "mobilenet_v1")
    return MobileNetV1(scale=scale, **kwargs)


# --------------------------------------------------------------------------------------------------------- MobileNetV3
class SqueezeExcitation(nn.Layer):
    def __init__(self, input_channels, squeeze_channels, activation=nn.ReLU, scale_activation=nn.Sigmoid):
        super().__init__()
        self.avgpool = nn.AdaptiveAvgPool2D(1)
        self.fc1 = nn.Conv2D(input_channels, squeeze_channels, 1)
        self.fc2 = nn.Conv2D(squeeze_channels, input_channels, 1)
        self.activation = activation()
        self.scale_activation = scale_activation()

    def forward(self, x):
        return self.scale_activation(self.fc2(self.activation(self.fc1(self.avgpool(x))))) * x


class InvertedResidualConfig:
    def __init__(self, in_channels, kernel, expanded_channels, out_channels, use_se, activation, stride, scale=1.0):
        self.in_channels = self.adjust_channels(in_channels, scale)
        self.kernel = kernel
        self.expanded_channels = self.adjust_channels(expanded_channels, scale)
        self.out_channels = self.adjust_channels(out_channels, scale)
        self.use_se = use_se
        if activation is None:
            self.activation_layer = None
        elif activation == "relu":
            self.activation_layer = nn.ReLU
        elif activation == "hardswish":
            self.activation_layer = nn.Hardswish
        else:
            raise RuntimeError(f"The activation function is not supported: {activation}")
        self.stride = stride

    @staticmethod
    def adjust_channels(channels, scale=1.0):
        return _make_divisible(channels * scale, 8)


class _MBv3Block(nn.Layer):
    def __init__(self, in_channels, expanded_channels, out_channels, filter_size, stride, use_se, activation_layer, norm_layer):
        super().__init__()
        self.use_res_connect = stride == 1 and in_channels == out_channels
        self.use_se = use_se
        self.expand = in_channels != expanded_channels
        if self.expand:
            self.expand_conv = ConvNormActivation(in_channels, expanded_channels, 1, stride=1, padding=0, norm_layer=norm_layer,
                                                  activation_layer=activation_layer)
        self.bottleneck_conv = ConvNormActivation(expanded_channels, expanded_channels, filter_size, stride=stride,
                                                  padding=(filter_size - 1) // 2, groups=expanded_channels, norm_layer=norm_layer,
                                                  activation_layer=activation_layer)
        if use_se:
            self.mid_se = SqueezeExcitation(expanded_channels, _make_divisible(expanded_channels // 4), scale_activation=nn.Hardsigmoid)
        self.linear_conv = ConvNormActivation(expanded_channels, out_channels, 1, stride=1, padding=0, norm_layer=norm_layer,
                                              activation_layer=None)

    def forward(self, x):
        identity = x
        if self.expand:
            x = self.expand_conv(x)
        x = self.bottleneck_conv(x)
        if self.use_se:
            x = self.mid_se(x)
        x = self.linear_conv(x)
        return x + identity if self.use_res_connect else x


class MobileNetV3(nn.Layer, _Head):
    def __init__(self, config, last_channel, scale=1.0, num_classes=1000, with_pool=True):
        super().__init__()
        self.config, self.scale, self.last_channel = config, scale, last_channel
        self.num_classes, self.with_pool = num_classes, with_pool
        self.firstconv_in_channels = config[0].in_channels
        self.lastconv_in_channels = config[-1].in_channels
        self.lastconv_out_channels = self.lastconv_in_channels * 6
        norm_layer = partial(nn.BatchNorm2D, epsilon=0.001, momentum=0.99)
        self.conv = ConvNormActivation(3, self.firstconv_in_channels, 3, stride=2, padding=1, activation_layer=nn.Hardswish,
                                       norm_layer=norm_layer)
        self.blocks = nn.Sequential(*[_MBv3Block(c.in_channels, c.expanded_channels, c.out_channels, c.kernel, c.stride, c.use_se,
                                                 c.activation_layer, norm_layer) for c in config])
        self.lastconv = ConvNormActivation(self.lastconv_in_channels, self.lastconv_out_channels, 1, stride=1, padding=0,
                                           norm_layer=norm_layer, activation_layer=nn.Hardswish)
        if with_pool:
            self.avgpool = nn.AdaptiveAvgPool2D(1)
        if num_classes > 0:
            self.classifier = nn.Sequential(nn.Linear(self.lastconv_out_channels, last_channel), nn.Hardswish(), nn.Dropout(p=0.2),
                                            nn.Linear(last_channel, num_classes))

    def forward(self, x):
        x = self.lastconv(self.blocks(self.conv(x)))
        return self._classify(x, getattr(self, "avgpool", None), getattr(self, "classifier", None))


# in, kernel, expanded, out, squeeze-excite, activation, stride
_MBV3_SMALL = [(16, 3, 16, 16, True, "relu", 2), (16, 3, 72, 24, False, "relu", 2), (24, 3, 88, 24, False, "relu", 1),
               (24, 5, 96, 40, True, "hardswish", 2), (40, 5, 240, 40, True, "hardswish", 1), (40, 5, 240, 40, True, "hardswish", 1),
               (40, 5, 120, 48, True, "hardswish", 1), (48, 5, 144, 48, True, "hardswish", 1), (48, 5, 288, 96, True, "hardswish", 2),
               (96, 5, 576, 96, True, "hardswish", 1), (96, 5, 576, 96, True, "hardswish", 1)]
_MBV3_LARGE = [(16, 3, 16, 16, False, "relu", 1), (16, 3, 64, 24, False, "relu", 2), (24, 3, 72, 24, False, "relu", 1),
               (24, 5, 72, 40, True, "relu", 2), (40, 5, 120, 40, True, "relu", 1), (40, 5, 120, 40, True, "relu", 1),
               (40, 3, 240, 80, False, "hardswish", 2), (80, 3, 200, 80, False, "hardswish", 1), (80, 3, 184, 80, False, "hardswish", 1),
               (80, 3, 184, 80, False, "hardswish", 1), (80, 3, 480, 112, True, "hardswish", 1), (112, 3, 672, 112, True, "hardswish", 1),
               (112, 5, 672, 160, True, "hardswish", 2), (160, 5, 960, 160, True, "hardswish", 1), (160, 5, 960, 160, True, "hardswish", 1)]


class MobileNetV3Small(MobileNetV3):
    def __init__(self, scale=1.0, num_classes=1000, with_pool=True):
        super().__init__([InvertedResidualConfig(*row, scale) for row in _MBV3_SMALL], _make_divisible(1024 * scale, 8), scale,
                         num_classes, with_pool)


class MobileNetV3Large(MobileNetV3):
    def __init__(self, scale=1.0, num_classes=1000, with_pool=True):
        super().__init__([InvertedResidualConfig(*row, scale) for row in _MBV3_LARGE], _make_divisible(1280 * scale, 8), scale,
                         num_classes, with_pool)


def mobilenet_v3_small(pretrained=False, scale=1.0, **kwargs):
    _no_pretrained(pretrained, "mobilenet_v3_small")
    return MobileNetV3Small(scale=scale, **kwargs)


def mobilenet_v3_large(pretrained=False, scale=1.0, **kwargs):
    _no_pretrained(pretrained, "mobilenet_v3_large")
    return MobileNetV3Large(scale=scale, **kwargs)


# ------------------------------------------------------------------------------------------------------------ DenseNet
class BNACConvLayer(nn.Layer):
    """BatchNorm + activation, then convolution (pre-activation order)."""

    def __init__(self, num_channels, num_filters, filter_size, stride=1, pad=0, groups=1, act="relu"):
        super().__init__()
        self._batch_norm = nn.BatchNorm(num_channels, act=act)
        self._conv = nn.Conv2D(num_channels, num_filters, filter_size, stride=stride, padding=pad, groups=groups, bias_attr=False)

    def forward(self, x):
        return self._conv(self._batch_norm(x))


class ConvBNLayer(nn.Layer):
    def __init__(self, num_channels, num_filters, filter_size, stride=1, pad=0, groups=1, act="relu"):
        super().__init__()
        self._conv = nn.Conv2D(num_channels, num_filters, filter_size, stride=stride, padding=pad, groups=groups, bias_attr=False)
        self._batch_norm = nn.BatchNorm(num_filters, act=act)

    def forward(self, x):
        return self._batch_norm(self._conv(x))


class DenseLayer(nn.Layer):
    def __init__(self, num_channels, growth_rate, bn_size, dropout):
        super().__init__()
        self.dropout = dropout
        self.bn_ac_func1 = BNACConvLayer(num_channels, bn_size * growth_rate, 1)
        self.bn_ac_func2 = BNACConvLayer(bn_size * growth_rate, growth_rate, 3, pad=1)
        if dropout:
            self.dropout_func = nn.Dropout(p=dropout, mode="downscale_in_infer")

    def forward(self, x):
        y = self.bn_ac_func2(self.bn_ac_func1(x))
        if self.dropout:
            y = self.dropout_func(y)
        return torch.cat([x, y], 1)


class DenseBlock(nn.Layer):
    def __init__(self, num_channels, num_layers, bn_size, growth_rate, dropout, name=None):
        super().__init__()
        self.dropout = dropout
        self.dense_layer_func = []
        for i in range(num_layers):
            layer = DenseLayer(num_channels + i * growth_rate, growth_rate, bn_size, dropout)
            self.add_sublayer(f"{name}_{i + 1}", layer)
            self.dense_layer_func.append(layer)

    def forward(self, x):
        for layer in self.dense_layer_func:
            x = layer(x)
        return x


class TransitionLayer(nn.Layer):
    def __init__(self, num_channels, num_output_features):
        super().__init__()
        self.conv_ac_func = BNACConvLayer(num_channels, num_output_features, 1)
        self.pool2d_avg = nn.AvgPool2D(kernel_size=2, stride=2, padding=0)

    def forward(self, x):
        return self.pool2d_avg(self.conv_ac_func(x))


class DenseNet(nn.Layer):
    # depth -> (stem width, growth rate, layers per dense block)
    _SPEC = {121: (64, 32, [6, 12, 24, 16]), 161: (96, 48, [6, 12, 36, 24]), 169: (64, 32, [6, 12, 32, 32]),
             201: (64, 32, [6, 12, 48, 32]), 264: (64, 32, [6, 12, 64, 48])}

    def __init__(self, layers=121, bn_size=4, dropout=0.0, num_classes=1000, with_pool=True):
        super().__init__()
        assert layers in self._SPEC, f"supported layers are {sorted(self._SPEC)} but input layer is {layers}"
        self.num_classes, self.with_pool = num_classes, with_pool
        width, growth_rate, self.block_config = self._SPEC[layers]
        self.conv1_func = ConvBNLayer(3, width, 7, stride=2, pad=3)
        self.pool2d_max = nn.MaxPool2D(kernel_size=3, stride=2, padding=1)
        self.dense_block_func_list, self.transition_func_list = [], []
        for i, num_layers in enumerate(self.block_config):
            block = DenseBlock(width, num_layers, bn_size, growth_rate, dropout, name=f"conv{i + 2}")
            self.add_sublayer(f"db_conv_{i + 2}", block)
            self.dense_block_func_list.append(block)
            width += num_layers * growth_rate
            if i != len(self.block_config) - 1:
                transition = TransitionLayer(width, width // 2)
                self.add_sublayer(f"tr_conv{i + 2}_blk", transition)
                self.transition_func_list.append(transition)
                width //= 2
        self.batch_norm = nn.BatchNorm(width, act="relu")
        if with_pool:
            self.pool2d_avg = nn.AdaptiveAvgPool2D(1)
        if num_classes > 0:
            bound = 1.0 / math.sqrt(width)
            self.out = nn.Linear(width, num_classes, weight_attr=ParamAttr(initializer=I.Uniform(-bound, bound)))

    def forward(self, x):
        x = self.pool2d_max(self.conv1_func(x))
        for i, block in enumerate(self.dense_block_func_list):
            x = block(x)
            if i < len(self.transition_func_list):
                x = self.transition_func_list[i](x)
        x = self.batch_norm(x)
        if self.with_pool:
            x = self.pool2d_avg(x)
        if self.num_classes > 0:
            x = self.out(torch.flatten(x, 1))
        return x


def _densenet(layers, pretrained, **kwargs):
    _no_pretrained(pretrained, f"densenet{layers}")
    return DenseNet(layers=layers, **kwargs)


def densenet121(pretrained=False, **kwargs):
    return _densenet(121, pretrained, **kwargs)


def densenet161(pretrained=False, **kwargs):
    return _densenet(161, pretrained, **kwargs)


def densenet169(pretrained=False, **kwargs):
    return _densenet(169, pretrained, **kwargs)


def densenet201(pretrained=False, **kwargs):
    return _densenet(201, pretrained, **kwargs)


def densenet264(pretrained=False, **kwargs):
    return _densenet(264, pretrained, **kwargs)


# ---------------------------------------------------------------------------------------------------------- SqueezeNet
class MakeFireConv(nn.Layer):
    def __init__(self, input_channels, output_channels, filter_size, padding=0):
        super().__init__()
        self._conv = nn.Conv2D(input_channels, output_channels, filter_size, padding=padding)

    def forward(self, x):
        return torch.relu(self._conv(x))


class MakeFire(nn.Layer):
    """Fire module: 1x1 squeeze, then parallel 1x1 and 3x3 expands, concatenated."""

    def __init__(self, input_channels, squeeze_channels, expand1x1_channels, expand3x3_channels):
        super().__init__()
        self._conv = MakeFireConv(input_channels, squeeze_channels, 1)
        self._conv_path1 = MakeFireConv(squeeze_channels, expand1x1_channels, 1)
        self._conv_path2 = MakeFireConv(squeeze_channels, expand3x3_channels, 3, padding=1)

    def forward(self, x):
        x = self._conv(x)
        return torch.cat([self._conv_path1(x), self._conv_path2(x)], 1)


class SqueezeNet(nn.Layer):
    # fire modules (input, squeeze, expand 1x1, expand 3x3); the two versions differ in the stem and in where they pool
    _FIRES = [(None, 16, 64, 64), (128, 16, 64, 64), (128, 32, 128, 128), (256, 32, 128, 128), (256, 48, 192, 192),
              (384, 48, 192, 192), (384, 64, 256, 256), (512, 64, 256, 256)]
    _POOL_AFTER = {"1.0": (3, 7), "1.1": (2, 4)}

    def __init__(self, version, num_classes=1000, with_pool=True):
        super().__init__()
        self.version, self.num_classes, self.with_pool = version, num_classes, with_pool
        assert version in self._POOL_AFTER, f"supported versions are {sorted(self._POOL_AFTER)} but input version is {version}"
        if version == "1.0":
            stem = 96
            self._conv = nn.Conv2D(3, stem, 7, stride=2)
        else:
            stem = 64
            self._conv = nn.Conv2D(3, stem, 3, stride=2, padding=1)
        self._pool = nn.MaxPool2D(kernel_size=3, stride=2, padding=0)
        for i, (cin, sq, e1, e3) in enumerate(self._FIRES, 1):
            setattr(self, f"_conv{i}", MakeFire(stem if cin is None else cin, sq, e1, e3))
        self._drop = nn.Dropout(p=0.5, mode="downscale_in_infer")
        self._conv9 = nn.Conv2D(512, num_classes, 1)
        self._avg_pool = nn.AdaptiveAvgPool2D(1)

    def forward(self, x):
        x = self._pool(torch.relu(self._conv(x)))
        for i in range(1, 9):
            x = getattr(self, f"_conv{i}")(x)
            if i in self._POOL_AFTER[self.version]:
                x = self._pool(x)
        if self.num_classes > 0:
            x = self._conv9(self._drop(x))
        if self.with_pool:
            x = self._avg_pool(torch.relu(x)).squeeze(-1).squeeze(-1)
        return x


def squeezenet1_0(pretrained=False, **kwargs):
    _no_pretrained(pretrained, "squeezenet1_0")
    return SqueezeNet("1.0", **kwargs)


def squeezenet1_1(pretrained=False, **kwargs):
    _no_pretrained(pretrained, "squeezenet1_1")
    return SqueezeNet("1.1", **kwargs)


# ----------------------------------------------------------------------------------------------------------- GoogLeNet
def _xavier(channels, filter_size):
    bound = (3.0 / (filter_size ** 2 * channels)) ** 0.5
    return ParamAttr(initializer=I.Uniform(-bound, bound))


class ConvLayer(nn.Layer):
    def __init__(self, num_channels, num_filters, filter_size, stride=1, groups=1):
        super().__init__()
        self._conv = nn.Conv2D(num_channels, num_filters, filter_size, stride=stride, padding=(filter_size - 1) // 2, groups=groups,
                               bias_attr=False)

    def forward(self, x):
        return self._conv(x)


class Inception(nn.Layer):
    """Four parallel branches (1x1, 1x1 -> 3x3, 1x1 -> 5x5, pool -> 1x1), concatenated, then ReLU."""

    def __init__(self, input_channels, output_channels, filter1, filter3R, filter3, filter5R, filter5, proj):
        super().__init__()
        self._conv1 = ConvLayer(input_channels, filter1, 1)
        self._conv3r = ConvLayer(input_channels, filter3R, 1)
        self._conv3 = ConvLayer(filter3R, filter3, 3)
        self._conv5r = ConvLayer(input_channels, filter5R, 1)
        self._conv5 = ConvLayer(filter5R, filter5, 5)
        self._pool = nn.MaxPool2D(kernel_size=3, stride=1, padding=1)
        self._convprj = ConvLayer(input_channels, proj, 1)

    def forward(self, x):
        return torch.relu(torch.cat([self._conv1(x), self._conv3(self._conv3r(x)), self._conv5(self._conv5r(x)),
                                     self._convprj(self._pool(x))], 1))


class GoogLeNet(nn.Layer):
    """Inception v1.  ``forward`` returns [main output, auxiliary output 1 (from 4a), auxiliary output 2 (from 4d)]."""
    _INCEPTIONS = [("3a", 192, 64, 96, 128, 16, 32, 32), ("3b", 256, 128, 128, 192, 32, 96, 64), ("4a", 480, 192, 96, 208, 16, 48, 64),
                   ("4b", 512, 160, 112, 224, 24, 64, 64), ("4c", 512, 128, 128, 256, 24, 64, 64), ("4d", 512, 112, 144, 288, 32, 64, 64),
                   ("4e", 528, 256, 160, 320, 32, 128, 128), ("5a", 832, 256, 160, 320, 32, 128, 128),
                   ("5b", 832, 384, 192, 384, 48, 128, 128)]

    def __init__(self, num_classes=1000, with_pool=True):
        super().__init__()
        self.num_classes, self.with_pool = num_classes, with_pool
        self._conv = ConvLayer(3, 64, 7, 2)
        self._pool = nn.MaxPool2D(kernel_size=3, stride=2)
        self._conv_1 = ConvLayer(64, 64, 1)
        self._conv_2 = ConvLayer(64, 192, 3)
        for name, cin, *widths in self._INCEPTIONS:
            setattr(self, f"_ince{name}", Inception(cin, cin, *widths))
        if with_pool:
            self._pool_5 = nn.AdaptiveAvgPool2D(1)
            self._pool_o1 = nn.AvgPool2D(kernel_size=5, stride=3)
            self._pool_o2 = nn.AvgPool2D(kernel_size=5, stride=3)
        if num_classes > 0:
            self._drop = nn.Dropout(p=0.4, mode="downscale_in_infer")
            self._fc_out = nn.Linear(1024, num_classes, weight_attr=_xavier(1024, 1))
            self._conv_o1 = ConvLayer(512, 128, 1)
            self._fc_o1 = nn.Linear(1152, 1024, weight_attr=_xavier(2048, 1))
            self._drop_o1 = nn.Dropout(p=0.7, mode="downscale_in_infer")
            self._out1 = nn.Linear(1024, num_classes, weight_attr=_xavier(1024, 1))
            self._conv_o2 = ConvLayer(528, 128, 1)
            self._fc_o2 = nn.Linear(1152, 1024, weight_attr=_xavier(2048, 1))
            self._drop_o2 = nn.Dropout(p=0.7, mode="downscale_in_infer")
            self._out2 = nn.Linear(1024, num_classes, weight_attr=_xavier(1024, 1))

    def forward(self, x):
        x = self._pool(self._conv(x))
        x = self._pool(self._conv_2(self._conv_1(x)))
        x = self._pool(self._ince3b(self._ince3a(x)))
        ince4a = self._ince4a(x)
        ince4d = self._ince4d(self._ince4c(self._ince4b(ince4a)))
        x = self._pool(self._ince4e(ince4d))
        out, out1, out2 = self._ince5b(self._ince5a(x)), ince4a, ince4d
        if self.with_pool:
            out, out1, out2 = self._pool_5(out), self._pool_o1(out1), self._pool_o2(out2)
        if self.num_classes > 0:
            out = self._fc_out(self._drop(out).squeeze(-1).squeeze(-1))
            out1 = self._out1(self._drop_o1(torch.relu(self._fc_o1(torch.flatten(self._conv_o1(out1), 1)))))
            out2 = self._out2(self._drop_o2(self._fc_o2(torch.flatten(self._conv_o2(out2), 1))))
        return [out, out1, out2]


def googlenet(pretrained=False, **kwargs):
    _no_pretrained(pretrained, "googlenet")
    return GoogLeNet(**kwargs)


# --------------------------------------------------------------------------------------------------------- InceptionV3
class _Branches(nn.Layer):
    """An Inception-v3 block from a table: every branch is a chain of (attribute name, out channels, kernel, stride,
    padding) convolutions; a chain may fork at its end (a tuple of two rows, both applied to the chain's value and
    concatenated); the pooling branch is ("avg" | "max") plus an optional 1x1 projection."""

    def _build(self, num_channels, branches, pool, pool_proj):
        self._plan = []
        for chain in branches:
            cin, steps = num_channels, []
            for row in chain:
                rows = row if isinstance(row[0], tuple) else (row,)
                names = []
                for name, cout, k, s, p in rows:
                    setattr(self, name, ConvNormActivation(cin, cout, k, stride=s, padding=p))
                    names.append(name)
                steps.append(names)
                cin = rows[0][1]
            self._plan.append(steps)
        if pool == "avg":
            self.branch_pool = nn.AvgPool2D(kernel_size=3, stride=1, padding=1, exclusive=False)
        else:
            self.branch_pool = nn.MaxPool2D(kernel_size=3, stride=2)
        if pool_proj:
            self.branch_pool_conv = ConvNormActivation(num_channels, pool_proj, 1, padding=0)
        self._has_proj = bool(pool_proj)

    def forward(self, x):
        outs = []
        for steps in self._plan:
            y = x
            for names in steps:
                y = getattr(self, names[0])(y) if len(names) == 1 else torch.cat([getattr(self, n)(y) for n in names], 1)
            outs.append(y)
        pooled = self.branch_pool(x)
        outs.append(self.branch_pool_conv(pooled) if self._has_proj else pooled)
        return torch.cat(outs, 1)


class InceptionStem(nn.Layer):
    def __init__(self):
        super().__init__()
        self.conv_1a_3x3 = ConvNormActivation(3, 32, 3, stride=2, padding=0)
        self.conv_2a_3x3 = ConvNormActivation(32, 32, 3, stride=1, padding=0)
        self.conv_2b_3x3 = ConvNormActivation(32, 64, 3, padding=1)
        self.max_pool = nn.MaxPool2D(kernel_size=3, stride=2, padding=0)
        self.conv_3b_1x1 = ConvNormActivation(64, 80, 1, padding=0)
        self.conv_4a_3x3 = ConvNormActivation(80, 192, 3, padding=0)

    def forward(self, x):
        x = self.max_pool(self.conv_2b_3x3(self.conv_2a_3x3(self.conv_1a_3x3(x))))
        return self.max_pool(self.conv_4a_3x3(self.conv_3b_1x1(x)))


class InceptionA(_Branches):
    def __init__(self, num_channels, pool_features):
        super().__init__()
        self._build(num_channels, [[("branch1x1", 64, 1, 1, 0)],
                                   [("branch5x5_1", 48, 1, 1, 0), ("branch5x5_2", 64, 5, 1, 2)],
                                   [("branch3x3dbl_1", 64, 1, 1, 0), ("branch3x3dbl_2", 96, 3, 1, 1), ("branch3x3dbl_3", 96, 3, 1, 1)]],
                    "avg", pool_features)


class InceptionB(_Branches):
    def __init__(self, num_channels):
        super().__init__()
        self._build(num_channels, [[("branch3x3", 384, 3, 2, 0)],
                                   [("branch3x3dbl_1", 64, 1, 1, 0), ("branch3x3dbl_2", 96, 3, 1, 1), ("branch3x3dbl_3", 96, 3, 2, 0)]],
                    "max", 0)


class InceptionC(_Branches):
    def __init__(self, num_channels, channels_7x7):
        super().__init__()
        c = channels_7x7
        self._build(num_channels, [[("branch1x1", 192, 1, 1, 0)],
                                   [("branch7x7_1", c, 1, 1, 0), ("branch7x7_2", c, (1, 7), 1, (0, 3)), ("branch7x7_3", 192, (7, 1), 1, (3, 0))],
                                   [("branch7x7dbl_1", c, 1, 1, 0), ("branch7x7dbl_2", c, (7, 1), 1, (3, 0)), ("branch7x7dbl_3", c, (1, 7), 1, (0, 3)),
                                    ("branch7x7dbl_4", c, (7, 1), 1, (3, 0)), ("branch7x7dbl_5", 192, (1, 7), 1, (0, 3))]],
                    "avg", 192)


class InceptionD(_Branches):
    def __init__(self, num_channels):
        super().__init__()
        self._build(num_channels, [[("branch3x3_1", 192, 1, 1, 0), ("branch3x3_2", 320, 3, 2, 0)],
                                   [("branch7x7x3_1", 192, 1, 1, 0), ("branch7x7x3_2", 192, (1, 7), 1, (0, 3)),
                                    ("branch7x7x3_3", 192, (7, 1), 1, (3, 0)), ("branch7x7x3_4", 192, 3, 2, 0)]],
                    "max", 0)


class InceptionE(_Branches):
    def __init__(self, num_channels):
        super().__init__()
        self._build(num_channels, [[("branch1x1", 320, 1, 1, 0)],
                                   [("branch3x3_1", 384, 1, 1, 0),
                                    (("branch3x3_2a", 384, (1, 3), 1, (0, 1)), ("branch3x3_2b", 384, (3, 1), 1, (1, 0)))],
                                   [("branch3x3dbl_1", 448, 1, 1, 0), ("branch3x3dbl_2", 384, 3, 1, 1),
                                    (("branch3x3dbl_3a", 384, (1, 3), 1, (0, 1)), ("branch3x3dbl_3b", 384, (3, 1), 1, (1, 0)))]],
                    "avg", 192)


class InceptionV3(nn.Layer):
    def __init__(self, num_classes=1000, with_pool=True):
        super().__init__()
        self.num_classes, self.with_pool = num_classes, with_pool
        self.layers_config = {"inception_a": [[192, 256, 288], [32, 64, 64]], "inception_b": [288],
                              "inception_c": [[768, 768, 768, 768], [128, 160, 160, 192]], "inception_d": [768],
                              "inception_e": [1280, 2048]}
        cfg = self.layers_config
        self.inception_stem = InceptionStem()
        blocks = [InceptionA(c, p) for c, p in zip(*cfg["inception_a"])]
        blocks += [InceptionB(c) for c in cfg["inception_b"]]
        blocks += [InceptionC(c, w) for c, w in zip(*cfg["inception_c"])]
        blocks += [InceptionD(c) for c in cfg["inception_d"]]
        blocks += [InceptionE(c) for c in cfg["inception_e"]]
        self.inception_block_list = nn.LayerList(blocks)
        if with_pool:
            self.avg_pool = nn.AdaptiveAvgPool2D(1)
        if num_classes > 0:
            self.dropout = nn.Dropout(p=0.2, mode="downscale_in_infer")
            bound = 1.0 / math.sqrt(2048.0)
            self.fc = nn.Linear(2048, num_classes, weight_attr=ParamAttr(initializer=I.Uniform(-bound, bound)))

    def forward(self, x):
        x = self.inception_stem(x)
        for block in self.inception_block_list:
            x = block(x)
        if self.with_pool:
            x = self.avg_pool(x)
        if self.num_classes > 0:
            x = self.fc(self.dropout(torch.flatten(x, 1)))
        return x


def inception_v3(pretrained=False, **kwargs):
    _no_pretrained(pretrained, "inception_v3")
    return InceptionV3(**kwargs)


# -------------------------------------------------------------------------------------------------------- ShuffleNetV2
def create_activation_layer(act):
    if act == "swish":
        return nn.Swish
    if act == "relu":
        return nn.ReLU
    if act is None:
        return None
    raise RuntimeError(f"The activation function is not supported: {act}")


def channel_shuffle(x, groups):
    n, c, h, w = x.shape
    return x.reshape(n, groups, c // groups, h, w).transpose(1, 2).reshape(n, c, h, w)


class InvertedResidual(nn.Layer):
    """Stride-1 unit: half of the channels pass through, the other half go 1x1 -> depthwise 3x3 -> 1x1; then shuffle."""

    def __init__(self, in_channels, out_channels, stride, activation_layer=nn.ReLU):
        super().__init__()
        half = out_channels // 2
        self._conv_pw = ConvNormActivation(in_channels // 2, half, 1, stride=1, padding=0, activation_layer=activation_layer)
        self._conv_dw = ConvNormActivation(half, half, 3, stride=stride, padding=1, groups=half, activation_layer=None)
        self._conv_linear = ConvNormActivation(half, half, 1, stride=1, padding=0, activation_layer=activation_layer)

    def forward(self, x):
        x1, x2 = x.chunk(2, dim=1)
        return channel_shuffle(torch.cat([x1, self._conv_linear(self._conv_dw(self._conv_pw(x2)))], 1), 2)


class InvertedResidualDS(nn.Layer):
    """Down-sampling unit: both halves are computed from the whole input (depthwise stride 2 in each)."""

    def __init__(self, in_channels, out_channels, stride, activation_layer=nn.ReLU):
        super().__init__()
        half = out_channels // 2
        self._conv_dw_1 = ConvNormActivation(in_channels, in_channels, 3, stride=stride, padding=1, groups=in_channels, activation_layer=None)
        self._conv_linear_1 = ConvNormActivation(in_channels, half, 1, stride=1, padding=0, activation_layer=activation_layer)
        self._conv_pw_2 = ConvNormActivation(in_channels, half, 1, stride=1, padding=0, activation_layer=activation_layer)
        self._conv_dw_2 = ConvNormActivation(half, half, 3, stride=stride, padding=1, groups=half, activation_layer=None)
        self._conv_linear_2 = ConvNormActivation(half, half, 1, stride=1, padding=0, activation_layer=activation_layer)

    def forward(self, x):
        x1 = self._conv_linear_1(self._conv_dw_1(x))
        x2 = self._conv_linear_2(self._conv_dw_2(self._conv_pw_2(x)))
        return channel_shuffle(torch.cat([x1, x2], 1), 2)


class ShuffleNetV2(nn.Layer, _Head):
    _WIDTHS = {0.25: [24, 24, 48, 96, 512], 0.33: [24, 32, 64, 128, 512], 0.5: [24, 48, 96, 192, 1024],
               1.0: [24, 116, 232, 464, 1024], 1.5: [24, 176, 352, 704, 1024], 2.0: [24, 224, 488, 976, 2048]}

    def __init__(self, scale=1.0, act="relu", num_classes=1000, with_pool=True):
        super().__init__()
        self.scale, self.num_classes, self.with_pool = scale, num_classes, with_pool
        if scale not in self._WIDTHS:
            raise NotImplementedError("This scale size:[" + str(scale) + "] is not implemented!")
        widths = self._WIDTHS[scale]
        activation_layer = create_activation_layer(act)
        self._conv1 = ConvNormActivation(3, widths[0], 3, stride=2, padding=1, activation_layer=activation_layer)
        self._max_pool = nn.MaxPool2D(kernel_size=3, stride=2, padding=1)
        self._block_list = []
        for stage, repeats in enumerate([4, 8, 4]):
            for i in range(repeats):
                if i == 0:
                    block = InvertedResidualDS(widths[stage], widths[stage + 1], 2, activation_layer)
                else:
                    block = InvertedResidual(widths[stage + 1], widths[stage + 1], 1, activation_layer)
                self.add_sublayer(f"{stage + 2}_{i + 1}", block)
                self._block_list.append(block)
        self._last_conv = ConvNormActivation(widths[-2], widths[-1], 1, stride=1, padding=0, activation_layer=activation_layer)
        if with_pool:
            self._pool2d_avg = nn.AdaptiveAvgPool2D(1)
        if num_classes > 0:
            self._out_c = widths[-1]
            self._fc = nn.Linear(widths[-1], num_classes)

    def forward(self, x):
        x = self._max_pool(self._conv1(x))
        for block in self._block_list:
            x = block(x)
        x = self._last_conv(x)
        return self._classify(x, getattr(self, "_pool2d_avg", None), getattr(self, "_fc", None))


def _shufflenet(arch, pretrained, **kwargs):
    _no_pretrained(pretrained, arch)
    return ShuffleNetV2(**kwargs)


def shufflenet_v2_x0_25(pretrained=False, **kwargs):
    return _shufflenet("shufflenet_v2_x0_25", pretrained, scale=0.25, **kwargs)


def shufflenet_v2_x0_33(pretrained=False, **kwargs):
    return _shufflenet("shufflenet_v2_x0_33", pretrained, scale=0.33, **kwargs)


def shufflenet_v2_x0_5(pretrained=False, **kwargs):
    return _shufflenet("shufflenet_v2_x0_5", pretrained, scale=0.5, **kwargs)


def shufflenet_v2_x1_0(pretrained=False, **kwargs):
    return _shufflenet("shufflenet_v2_x1_0", pretrained, scale=1.0, **kwargs)


def shufflenet_v2_x1_5(pretrained=False, **kwargs):
    return _shufflenet("shufflenet_v2_x1_5", pretrained, scale=1.5, **kwargs)


def shufflenet_v2_x2_0(pretrained=False, **kwargs):
    return _shufflenet("shufflenet_v2_x2_0", pretrained, scale=2.0, **kwargs)


def shufflenet_v2_swish(pretrained=False, **kwargs):
    return _shufflenet("shufflenet_v2_swish", pretrained, scale=1.0, act="swish", **kwargs)
