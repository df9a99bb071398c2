"""Convolution, pooling, padding, resampling and instance / group normalisation layers (reference:
python/paddle/nn/layer/{conv,pooling,common,norm,vision}.py).  Thin stateful shells over ``nn.functional.conv_pool``:
they own the parameters under the reference's names (``weight`` / ``bias``, InstanceNorm's ``scale`` / ``bias``), keep
the constructor signatures, and pass every option (``data_format``, ``padding_mode``, ``output_size``,
``return_mask``, ``exclusive`` ...) through."""
from __future__ import annotations

import math

import torch

from .. import functional as F
from .. import initializer as I
from .layers import Layer


def _tup(v, nd):
    return tuple(int(e) for e in v) if isinstance(v, (list, tuple)) else (int(v),) * nd


# --------------------------------------------------------------------------------------------------------------- conv
class _ConvNd(Layer):
    """Weight ``[out, in / groups, *k]`` (transposed: ``[in, out / groups, *k]``).  Default weight initializer of the
    forward convolutions: Normal(0, sqrt(2 / (prod(k) * in_channels))) (conv.py _ConvNd._get_default_param_initializer);
    transposed ones take the global default."""

    _nd = 2
    _transposed = False
    _formats = ("NCHW", "NHWC")

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 padding_mode="zeros", weight_attr=None, bias_attr=None, data_format=None, output_padding=0):
        super().__init__()
        nd = self._nd
        data_format = data_format or self._formats[0]
        if data_format not in self._formats and not (nd == 1 and data_format in ("NCW", "NWC")):
            raise ValueError(f"data_format must be one of {self._formats}, but got data_format='{data_format}'")
        if padding_mode not in ("zeros", "reflect", "replicate", "circular"):
            raise ValueError(f"padding_mode must be one of zeros, reflect, replicate, circular, but got '{padding_mode}'")
        if groups <= 0 or in_channels % groups != 0:
            raise ValueError("in_channels must be divisible by groups.")
        if out_channels % groups != 0:
            raise ValueError("out_channels must be divisible by groups.")
        if self._transposed and padding_mode != "zeros":
            raise ValueError("transposed convolutions only support padding_mode='zeros'")
        self._in_channels, self._out_channels, self._groups = in_channels, out_channels, groups
        self._kernel_size, self._stride, self._dilation = _tup(kernel_size, nd), _tup(stride, nd), _tup(dilation, nd)
        self._padding, self._padding_mode, self._data_format = padding, padding_mode, data_format
        self._output_padding = output_padding
        if padding_mode != "zeros":
            if isinstance(padding, str):
                raise ValueError(f"padding='{padding}' needs padding_mode='zeros'")
            p = _tup(padding, nd)
            if len(p) != nd:
                raise ValueError(f"padding_mode='{padding_mode}' takes an int or {nd} ints as padding")
            # F.pad order: innermost dim first, (before, after) each
            self._explicit_pad = [v for e in reversed(p) for v in (e, e)]
        if self._transposed:
            shape = [in_channels, out_channels // groups, *self._kernel_size]
            init = None
        else:
            shape = [out_channels, in_channels // groups, *self._kernel_size]
            init = I.Normal(0.0, math.sqrt(2.0 / (math.prod(self._kernel_size) * in_channels)))
        self.weight = self.create_parameter(shape, attr=weight_attr, default_initializer=init)
        self.bias = self.create_parameter([out_channels], attr=bias_attr, is_bias=True)

    def _fn(self):
        name = f"conv{self._nd}d" + ("_transpose" if self._transposed else "")
        return getattr(F, name)

    def forward(self, x, output_size=None):
        if self._transposed:
            opad = 0 if output_size is not None else self._output_padding
            return self._fn()(x, self.weight, self.bias, stride=self._stride, padding=self._padding,
                              output_padding=opad, groups=self._groups, dilation=self._dilation,
                              output_size=output_size, data_format=self._data_format)
        padding = self._padding
        if self._padding_mode != "zeros":
            x = F.pad(x, self._explicit_pad, mode=self._padding_mode, data_format=self._data_format)
            padding = 0
        return self._fn()(x, self.weight, self.bias, stride=self._stride, padding=padding, dilation=self._dilation,
                          groups=self._groups, data_format=self._data_format)

    def extra_repr(self):
        s = f"{self._in_channels}, {self._out_channels}, kernel_size={list(self._kernel_size)}"
        if self._stride != (1,) * self._nd:
            s += f", stride={list(self._stride)}"
        if self._padding != 0:
            s += f", padding={self._padding}"
        if self._padding_mode != "zeros":
            s += f", padding_mode={self._padding_mode}"
        if self._dilation != (1,) * self._nd:
            s += f", dilation={list(self._dilation)}"
        if self._groups != 1:
            s += f", groups={self._groups}"
        return s + f", data_format={self._data_format}"


class Conv1D(_ConvNd):
    _nd, _formats = 1, ("NCL", "NLC")

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 padding_mode="zeros", weight_attr=None, bias_attr=None, data_format="NCL"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, padding_mode,
                         weight_attr, bias_attr, data_format)


class Conv2D(_ConvNd):
    _nd, _formats = 2, ("NCHW", "NHWC")

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 padding_mode="zeros", weight_attr=None, bias_attr=None, data_format="NCHW"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, padding_mode,
                         weight_attr, bias_attr, data_format)


class Conv3D(_ConvNd):
    _nd, _formats = 3, ("NCDHW", "NDHWC")

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 padding_mode="zeros", weight_attr=None, bias_attr=None, data_format="NCDHW"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, padding_mode,
                         weight_attr, bias_attr, data_format)


class Conv1DTranspose(_ConvNd):
    _nd, _formats, _transposed = 1, ("NCL", "NLC"), True

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, output_padding=0, groups=1,
                 dilation=1, weight_attr=None, bias_attr=None, data_format="NCL"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, "zeros",
                         weight_attr, bias_attr, data_format, output_padding)


class Conv2DTranspose(_ConvNd):
    _nd, _formats, _transposed = 2, ("NCHW", "NHWC"), True

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, output_padding=0, dilation=1,
                 groups=1, weight_attr=None, bias_attr=None, data_format="NCHW"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, "zeros",
                         weight_attr, bias_attr, data_format, output_padding)


class Conv3DTranspose(_ConvNd):
    _nd, _formats, _transposed = 3, ("NCDHW", "NDHWC"), True

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, output_padding=0, dilation=1,
                 groups=1, weight_attr=None, bias_attr=None, data_format="NCDHW"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, "zeros",
                         weight_attr, bias_attr, data_format, output_padding)


# ------------------------------------------------------------------------------------------------------------ pooling
class _Pool(Layer):
    def __init__(self, fn, **kw):
        super().__init__()
        self._fn_name, self._kw = fn, kw
        for k, v in kw.items():
            setattr(self, k, v)

    def forward(self, x):
        return getattr(F, self._fn_name)(x, **self._kw)

    def extra_repr(self):
        return ", ".join(f"{k}={v}" for k, v in self._kw.items())


class AvgPool1D(_Pool):
    def __init__(self, kernel_size, stride=None, padding=0, exclusive=True, ceil_mode=False, name=None):
        super().__init__("avg_pool1d", kernel_size=kernel_size, stride=stride, padding=padding, exclusive=exclusive,
                         ceil_mode=ceil_mode)


class AvgPool2D(_Pool):
    def __init__(self, kernel_size, stride=None, padding=0, ceil_mode=False, exclusive=True, divisor_override=None,
                 data_format="NCHW", name=None):
        super().__init__("avg_pool2d", kernel_size=kernel_size, stride=stride, padding=padding, ceil_mode=ceil_mode,
                         exclusive=exclusive, divisor_override=divisor_override, data_format=data_format)


class AvgPool3D(_Pool):
    def __init__(self, kernel_size, stride=None, padding=0, ceil_mode=False, exclusive=True, divisor_override=None,
                 data_format="NCDHW", name=None):
        super().__init__("avg_pool3d", kernel_size=kernel_size, stride=stride, padding=padding, ceil_mode=ceil_mode,
                         exclusive=exclusive, divisor_override=divisor_override, data_format=data_format)


class MaxPool1D(_Pool):
    def __init__(self, kernel_size, stride=None, padding=0, return_mask=False, ceil_mode=False, name=None):
        super().__init__("max_pool1d", kernel_size=kernel_size, stride=stride, padding=padding, return_mask=return_mask,
                         ceil_mode=ceil_mode)


class MaxPool2D(_Pool):
    def __init__(self, kernel_size, stride=None, padding=0, return_mask=False, ceil_mode=False, data_format="NCHW",
                 name=None):
        super().__init__("max_pool2d", kernel_size=kernel_size, stride=stride, padding=padding, return_mask=return_mask,
                         ceil_mode=ceil_mode, data_format=data_format)


class MaxPool3D(_Pool):
    def __init__(self, kernel_size, stride=None, padding=0, return_mask=False, ceil_mode=False, data_format="NCDHW",
                 name=None):
        super().__init__("max_pool3d", kernel_size=kernel_size, stride=stride, padding=padding, return_mask=return_mask,
                         ceil_mode=ceil_mode, data_format=data_format)


class LPPool1D(_Pool):
    def __init__(self, norm_type, kernel_size, stride=None, padding=0, ceil_mode=False, data_format="NCL", name=None):
        super().__init__("lp_pool1d", norm_type=norm_type, kernel_size=kernel_size, stride=stride, padding=padding,
                         ceil_mode=ceil_mode, data_format=data_format)


class LPPool2D(_Pool):
    def __init__(self, norm_type, kernel_size, stride=None, padding=0, ceil_mode=False, data_format="NCHW", name=None):
        super().__init__("lp_pool2d", norm_type=norm_type, kernel_size=kernel_size, stride=stride, padding=padding,
                         ceil_mode=ceil_mode, data_format=data_format)


class AdaptiveAvgPool1D(_Pool):
    def __init__(self, output_size, name=None):
        super().__init__("adaptive_avg_pool1d", output_size=output_size)


class AdaptiveAvgPool2D(_Pool):
    def __init__(self, output_size, data_format="NCHW", name=None):
        super().__init__("adaptive_avg_pool2d", output_size=output_size, data_format=data_format)


class AdaptiveAvgPool3D(_Pool):
    def __init__(self, output_size, data_format="NCDHW", name=None):
        super().__init__("adaptive_avg_pool3d", output_size=output_size, data_format=data_format)


class AdaptiveMaxPool1D(_Pool):
    def __init__(self, output_size, return_mask=False, name=None):
        super().__init__("adaptive_max_pool1d", output_size=output_size, return_mask=return_mask)


class AdaptiveMaxPool2D(_Pool):
    def __init__(self, output_size, return_mask=False, name=None):
        super().__init__("adaptive_max_pool2d", output_size=output_size, return_mask=return_mask)


class AdaptiveMaxPool3D(_Pool):
    def __init__(self, output_size, return_mask=False, name=None):
        super().__init__("adaptive_max_pool3d", output_size=output_size, return_mask=return_mask)


class _MaxUnPool(Layer):
    _fn_name = "max_unpool2d"

    def __init__(self, kernel_size, stride=None, padding=0, data_format=None, output_size=None, name=None):
        super().__init__()
        self.ksize, self.stride, self.padding = kernel_size, stride, padding
        self.data_format, self.output_size = data_format, output_size

    def forward(self, x, indices):
        return getattr(F, self._fn_name)(x, indices, kernel_size=self.ksize, stride=self.stride, padding=self.padding,
                                         data_format=self.data_format, output_size=self.output_size)


class MaxUnPool1D(_MaxUnPool):
    _fn_name = "max_unpool1d"

    def __init__(self, kernel_size, stride=None, padding=0, data_format="NCL", output_size=None, name=None):
        super().__init__(kernel_size, stride, padding, data_format, output_size)


class MaxUnPool2D(_MaxUnPool):
    _fn_name = "max_unpool2d"

    def __init__(self, kernel_size, stride=None, padding=0, data_format="NCHW", output_size=None, name=None):
        super().__init__(kernel_size, stride, padding, data_format, output_size)


class MaxUnPool3D(_MaxUnPool):
    _fn_name = "max_unpool3d"

    def __init__(self, kernel_size, stride=None, padding=0, data_format="NCDHW", output_size=None, name=None):
        super().__init__(kernel_size, stride, padding, data_format, output_size)


# ------------------------------------------------------------------------------------------------- padding / resample
class _PadNd(Layer):
    """``padding``: an int (every side) or (left, right[, top, bottom[, front, back]]) - innermost dim first."""

    _nd, _default_format = 2, "NCHW"

    def __init__(self, padding, mode="constant", value=0.0, data_format=None, name=None):
        super().__init__()
        if isinstance(padding, int):
            padding = [padding] * (2 * self._nd)
        padding = [int(p) for p in (padding.tolist() if isinstance(padding, torch.Tensor) else padding)]
        if len(padding) != 2 * self._nd:
            raise ValueError(f"padding should be an int or {2 * self._nd} ints, got {padding}")
        self._pad, self._mode, self._value = padding, mode, value
        self._data_format = data_format or self._default_format

    def forward(self, x):
        return F.pad(x, self._pad, mode=self._mode, value=self._value, data_format=self._data_format)

    def extra_repr(self):
        return f"padding={self._pad}, mode={self._mode}, value={self._value}, data_format={self._data_format}"


class Pad1D(_PadNd):
    _nd, _default_format = 1, "NCL"


class Pad2D(_PadNd):
    _nd, _default_format = 2, "NCHW"


class Pad3D(_PadNd):
    _nd, _default_format = 3, "NCDHW"


class ZeroPad1D(_PadNd):
    _nd, _default_format = 1, "NCL"

    def __init__(self, padding, data_format="NCL", name=None):
        super().__init__(padding, "constant", 0.0, data_format)


class ZeroPad2D(_PadNd):
    _nd, _default_format = 2, "NCHW"

    def __init__(self, padding, data_format="NCHW", name=None):
        super().__init__(padding, "constant", 0.0, data_format)


class ZeroPad3D(_PadNd):
    _nd, _default_format = 3, "NCDHW"

    def __init__(self, padding, data_format="NCDHW", name=None):
        super().__init__(padding, "constant", 0.0, data_format)


class Upsample(Layer):
    def __init__(self, size=None, scale_factor=None, mode="nearest", align_corners=False, align_mode=0,
                 data_format=None, name=None):
        super().__init__()
        self.size, self.scale_factor, self.mode = size, scale_factor, mode.lower()
        self.align_corners, self.align_mode, self.data_format = align_corners, align_mode, data_format

    def forward(self, x):
        return F.interpolate(x, size=self.size, scale_factor=self.scale_factor, mode=self.mode,
                             align_corners=self.align_corners, align_mode=self.align_mode, data_format=self.data_format)

    def extra_repr(self):
        main = f"scale_factor={self.scale_factor}" if self.scale_factor is not None else f"size={self.size}"
        return f"{main}, mode={self.mode}, align_corners={self.align_corners}, align_mode={self.align_mode}"


class UpsamplingNearest2D(Upsample):
    def __init__(self, size=None, scale_factor=None, data_format="NCHW", name=None):
        super().__init__(size, scale_factor, "nearest", False, 0, data_format)


class UpsamplingBilinear2D(Upsample):
    def __init__(self, size=None, scale_factor=None, data_format="NCHW", name=None):
        super().__init__(size, scale_factor, "bilinear", True, 0, data_format)


class PixelShuffle(Layer):
    def __init__(self, upscale_factor, data_format="NCHW", name=None):
        super().__init__()
        if not isinstance(upscale_factor, int):
            raise TypeError("upscale factor must be int type")
        self._upscale_factor, self._data_format = upscale_factor, data_format

    def forward(self, x):
        return F.pixel_shuffle(x, self._upscale_factor, self._data_format)


class PixelUnshuffle(Layer):
    def __init__(self, downscale_factor, data_format="NCHW", name=None):
        super().__init__()
        if not isinstance(downscale_factor, int):
            raise TypeError("Downscale factor must be int type")
        self._downscale_factor, self._data_format = downscale_factor, data_format

    def forward(self, x):
        return F.pixel_unshuffle(x, self._downscale_factor, self._data_format)


class ChannelShuffle(Layer):
    def __init__(self, groups, data_format="NCHW", name=None):
        super().__init__()
        if not isinstance(groups, int):
            raise TypeError("groups must be int type")
        self._groups, self._data_format = groups, data_format

    def forward(self, x):
        return F.channel_shuffle(x, self._groups, self._data_format)


class Unfold(Layer):
    def __init__(self, kernel_sizes, dilations=1, paddings=0, strides=1, name=None):
        super().__init__()
        self.kernel_sizes, self.dilations, self.paddings, self.strides = kernel_sizes, dilations, paddings, strides

    def forward(self, x):
        return F.unfold(x, self.kernel_sizes, self.strides, self.paddings, self.dilations)


class Fold(Layer):
    def __init__(self, output_sizes, kernel_sizes, dilations=1, paddings=0, strides=1, name=None):
        super().__init__()
        self.output_sizes, self.kernel_sizes = output_sizes, kernel_sizes
        self.dilations, self.paddings, self.strides = dilations, paddings, strides

    def forward(self, x):
        return F.fold(x, self.output_sizes, self.kernel_sizes, self.strides, self.paddings, self.dilations)


# -------------------------------------------------------------------------------------------------------------- norms
class _InstanceNormBase(Layer):
    """Per-sample, per-channel normalisation; learnable ``scale`` / ``bias`` unless ``weight_attr`` / ``bias_attr`` is
    False.  No running statistics are kept (``momentum`` is accepted and unused, as in the reference, norm.py:66)."""

    _rank, _default_format = 4, "NCHW"

    def __init__(self, num_features, epsilon=1e-05, momentum=0.9, weight_attr=None, bias_attr=None, data_format=None,
                 name=None):
        super().__init__()
        if weight_attr is False or bias_attr is False:
            assert weight_attr == bias_attr, "weight_attr and bias_attr must be set to False at the same time in InstanceNorm"
        self._num_features, self._epsilon = num_features, epsilon
        self._data_format = data_format or self._default_format
        if weight_attr is not False and bias_attr is not False:
            self.scale = self.create_parameter([num_features], attr=weight_attr, default_initializer=I.Constant(1.0))
            self.bias = self.create_parameter([num_features], attr=bias_attr, is_bias=True,
                                              default_initializer=I.Constant(0.0))
        else:
            self.scale = None
            self.bias = None

    def forward(self, input):
        if input.dim() != self._rank and not (self._rank == 3 and input.dim() == 2):
            raise ValueError(f"expected {self._rank}D input (got {input.dim()}D input)")
        return F.instance_norm(input, weight=self.scale, bias=self.bias, eps=self._epsilon,
                               data_format=self._data_format)

    def extra_repr(self):
        return f"num_features={self._num_features}, epsilon={self._epsilon}"


class InstanceNorm1D(_InstanceNormBase):
    _rank, _default_format = 3, "NCL"


class InstanceNorm2D(_InstanceNormBase):
    _rank, _default_format = 4, "NCHW"


class InstanceNorm3D(_InstanceNormBase):
    _rank, _default_format = 5, "NCDHW"


class GroupNorm(Layer):
    def __init__(self, num_groups, num_channels, epsilon=1e-05, weight_attr=None, bias_attr=None, data_format="NCHW",
                 name=None):
        super().__init__()
        if num_channels % num_groups != 0:
            raise ValueError("num_channels must be divisible by num_groups")
        self._num_groups, self._num_channels, self._epsilon = num_groups, num_channels, epsilon
        self._data_format = data_format
        self.weight = None if weight_attr is False else self.create_parameter(
            [num_channels], attr=weight_attr, default_initializer=I.Constant(1.0))
        self.bias = None if bias_attr is False else self.create_parameter([num_channels], attr=bias_attr, is_bias=True)

    def forward(self, input):
        return F.group_norm(input, self._num_groups, self._epsilon, self.weight, self.bias, self._data_format)

    def extra_repr(self):
        return f"num_groups={self._num_groups}, num_channels={self._num_channels}, epsilon={self._epsilon}"


class LocalResponseNorm(Layer):
    def __init__(self, size, alpha=0.0001, beta=0.75, k=1.0, data_format="NCHW", name=None):
        super().__init__()
        self.size, self.alpha, self.beta, self.k, self.data_format = size, alpha, beta, k, data_format

    def forward(self, input):
        return F.local_response_norm(input, self.size, self.alpha, self.beta, self.k, self.data_format)


__all__ = [n for n, v in list(globals().items()) if isinstance(v, type) and issubclass(v, Layer) and not n.startswith("_")
           and n != "Layer"]
