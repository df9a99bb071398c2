"""paddle.sparse.nn (reference: python/paddle/sparse/nn/layer/{activation,conv,norm,pooling}.py)."""
from __future__ import annotations

import math

import torch

from ...nn import initializer as I
from ...nn.layer.layers import Layer
from . import functional  # noqa: F401
from . import functional as F

__all__ = ["ReLU", "ReLU6", "LeakyReLU", "Softmax", "BatchNorm", "SyncBatchNorm", "Conv2D", "Conv3D", "SubmConv2D", "SubmConv3D",
           "MaxPool3D"]


class ReLU(Layer):
    def forward(self, x):
        return F.relu(x)


class ReLU6(Layer):
    def forward(self, x):
        return F.relu6(x)


class LeakyReLU(Layer):
    def __init__(self, negative_slope=0.01, name=None):
        super().__init__()
        self._negative_slope = negative_slope

    def forward(self, x):
        return F.leaky_relu(x, self._negative_slope)


class Softmax(Layer):
    def __init__(self, axis=-1, name=None):
        super().__init__()
        self._axis = axis

    def forward(self, x):
        return F.softmax(x, self._axis)


class BatchNorm(Layer):
    """Batch normalisation of the stored values of a channel-last COO tensor: statistics per channel over the active
    sites.  Same running-statistics convention as the dense layer (running = m * running + (1 - m) * batch, biased
    variance)."""

    def __init__(self, num_features, momentum=0.9, epsilon=1e-05, weight_attr=None, bias_attr=None, data_format="NDHWC",
                 use_global_stats=None, name=None):
        super().__init__()
        if data_format not in ("NDHWC", "NHWC"):
            raise ValueError("sparse BatchNorm only support layout of 'NDHWC' and 'NHWC'")
        self._num_features, self._momentum, self._epsilon = num_features, momentum, epsilon
        self._data_format, self._use_global_stats = data_format, use_global_stats
        self.weight = None if weight_attr is False else self.create_parameter([num_features], attr=weight_attr,
                                                                              default_initializer=I.Constant(1.0))
        self.bias = None if bias_attr is False else self.create_parameter([num_features], attr=bias_attr, is_bias=True)
        self.register_buffer("_mean", torch.zeros(num_features))
        self.register_buffer("_variance", torch.ones(num_features))

    def _stats(self, values):
        mean = values.mean(0)
        var = values.var(0, unbiased=False)
        return mean, var

    def forward(self, input):
        x = input.coalesce()
        v = x.values()
        if v.dim() != 2 or v.shape[1] != self._num_features:
            raise ValueError(f"expected values of shape [nnz, {self._num_features}], got {tuple(v.shape)}")
        use_batch = self.training and not self._use_global_stats
        if use_batch:
            mean, var = self._stats(v.float())
            with torch.no_grad():
                self._mean.mul_(self._momentum).add_(mean.detach(), alpha=1 - self._momentum)
                self._variance.mul_(self._momentum).add_(var.detach(), alpha=1 - self._momentum)
        else:
            mean, var = self._mean, self._variance
        out = (v - mean.to(v.dtype)) * torch.rsqrt(var.to(v.dtype) + self._epsilon)
        if self.weight is not None:
            out = out * self.weight
        if self.bias is not None:
            out = out + self.bias
        return torch.sparse_coo_tensor(x.indices(), out, x.shape, is_coalesced=True)


class SyncBatchNorm(BatchNorm):
    """Statistics over the active sites of every rank of the default group."""

    def _stats(self, values):
        from ...distributed import collective as C
        from ...distributed import communication as comm
        if not C.is_initialized() or C.get_world_size() == 1:
            return super()._stats(values)
        packed = torch.cat([values.sum(0), (values * values).sum(0), values.new_tensor([float(values.shape[0])])])
        total = packed.detach().clone()
        comm.all_reduce(total)
        # gradient of the local share flows through `packed`; the other ranks' shares are constants here
        packed = packed + (total - packed.detach())
        c = values.shape[1]
        n = packed[-1]
        mean = packed[:c] / n
        return mean, packed[c:2 * c] / n - mean * mean

    @classmethod
    def convert_sync_batchnorm(cls, layer):
        out = layer
        if isinstance(layer, BatchNorm) and not isinstance(layer, SyncBatchNorm):
            out = SyncBatchNorm(layer._num_features, layer._momentum, layer._epsilon, data_format=layer._data_format)
            out.weight, out.bias = layer.weight, layer.bias
            out._mean.copy_(layer._mean)
            out._variance.copy_(layer._variance)
        for name, child in list(layer.named_children()):
            setattr(out, name, cls.convert_sync_batchnorm(child))
        return out


class _Conv(Layer):
    _nd, _subm, _format = 3, False, "NDHWC"

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, padding_mode="zeros",
                 key=None, weight_attr=None, bias_attr=None, data_format=None):
        super().__init__()
        assert weight_attr is not False, "weight_attr should not be False in Conv."
        nd = self._nd
        data_format = data_format or self._format
        if data_format != self._format:
            raise ValueError(f"data_format must be '{self._format}', but got data_format='{data_format}'")
        if padding_mode != "zeros":
            raise ValueError("Currently, only support padding_mode='zeros'")
        if groups != 1:
            raise ValueError("Currently, only support groups=1")
        tup = lambda v: tuple(v) if isinstance(v, (list, tuple)) else (v,) * nd  # noqa: E731
        self._in_channels, self._out_channels, self._groups, self._key = in_channels, out_channels, groups, key
        self._kernel_size, self._stride, self._padding, self._dilation = tup(kernel_size), tup(stride), tup(padding), tup(dilation)
        self._data_format = data_format
        std = math.sqrt(2.0 / (math.prod(self._kernel_size) * in_channels))
        self.weight = self.create_parameter([*self._kernel_size, in_channels // groups, out_channels], attr=weight_attr,
                                            default_initializer=I.Normal(0.0, std))
        self.bias = self.create_parameter([out_channels], attr=bias_attr, is_bias=True)

    def forward(self, x):
        return F._conv(self._nd, x, self.weight, self.bias, self._stride, self._padding, self._dilation, self._groups, self._subm,
                       self._data_format)

    def extra_repr(self):
        return (f"{self._in_channels}, {self._out_channels}, kernel_size={list(self._kernel_size)}, stride={list(self._stride)}, "
                f"padding={list(self._padding)}, data_format={self._data_format}")


class Conv3D(_Conv):
    _nd, _subm, _format = 3, False, "NDHWC"

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, padding_mode="zeros",
                 weight_attr=None, bias_attr=None, data_format="NDHWC"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, padding_mode, None, weight_attr,
                         bias_attr, data_format)


class Conv2D(_Conv):
    _nd, _subm, _format = 2, False, "NHWC"

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, padding_mode="zeros",
                 weight_attr=None, bias_attr=None, data_format="NHWC"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, padding_mode, None, weight_attr,
                         bias_attr, data_format)


class SubmConv3D(_Conv):
    _nd, _subm, _format = 3, True, "NDHWC"

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, padding_mode="zeros",
                 key=None, weight_attr=None, bias_attr=None, data_format="NDHWC"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, padding_mode, key, weight_attr,
                         bias_attr, data_format)


class SubmConv2D(_Conv):
    _nd, _subm, _format = 2, True, "NHWC"

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, padding_mode="zeros",
                 key=None, weight_attr=None, bias_attr=None, data_format="NHWC"):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, padding_mode, key, weight_attr,
                         bias_attr, data_format)


class MaxPool3D(Layer):
    def __init__(self, kernel_size, stride=None, padding=0, return_mask=False, ceil_mode=False, data_format="NDHWC", name=None):
        super().__init__()
        if return_mask:
            raise NotImplementedError("sparse MaxPool3D has no return_mask (neither has the reference)")
        self.ksize, self.stride, self.padding, self.ceil_mode, self.data_format = kernel_size, stride, padding, ceil_mode, data_format

    def forward(self, x):
        return F.max_pool3d(x, self.ksize, self.stride, self.padding, self.ceil_mode, self.data_format)
