"""Initializers (reference: python/paddle/nn/initializer/*.py).  Each is a callable ``init(param)`` that fills the
tensor in place; fan computation follows the reference (weight layout [in, out] for Linear, [out, in, kh, kw] conv)."""
from __future__ import annotations

import math

import numpy as np
import torch


def _fans(shape):
    shape = list(shape)
    if len(shape) == 0:
        return 1, 1
    if len(shape) == 1:
        return shape[0], shape[0]
    if len(shape) == 2:
        return shape[0], shape[1]
    rf = int(np.prod(shape[2:]))
    return shape[1] * rf, shape[0] * rf


class Initializer:
    def __call__(self, param, block=None):
        with torch.no_grad():
            self._fill(param)
        return param

    def _fill(self, p):
        raise NotImplementedError


class Constant(Initializer):
    def __init__(self, value=0.0):
        self.value = value

    def _fill(self, p):
        p.fill_(self.value)


class Uniform(Initializer):
    def __init__(self, low=-1.0, high=1.0, name=None):
        self.low, self.high = low, high

    def _fill(self, p):
        p.uniform_(self.low, self.high)


class Normal(Initializer):
    def __init__(self, mean=0.0, std=1.0, name=None):
        self.mean, self.std = mean, std

    def _fill(self, p):
        p.normal_(self.mean, self.std)


class TruncatedNormal(Initializer):
    def __init__(self, mean=0.0, std=1.0, a=-2.0, b=2.0, name=None):
        self.mean, self.std, self.a, self.b = mean, std, a, b

    def _fill(self, p):
        torch.nn.init.trunc_normal_(p, self.mean, self.std, self.mean + self.a * self.std, self.mean + self.b * self.std)


class XavierUniform(Initializer):
    def __init__(self, fan_in=None, fan_out=None, gain=1.0, name=None):
        self.fan_in, self.fan_out, self.gain = fan_in, fan_out, gain

    def _fill(self, p):
        fi, fo = _fans(p.shape)
        fi, fo = self.fan_in or fi, self.fan_out or fo
        limit = self.gain * math.sqrt(6.0 / float(fi + fo))
        p.uniform_(-limit, limit)


class XavierNormal(Initializer):
    def __init__(self, fan_in=None, fan_out=None, gain=1.0, name=None):
        self.fan_in, self.fan_out, self.gain = fan_in, fan_out, gain

    def _fill(self, p):
        fi, fo = _fans(p.shape)
        fi, fo = self.fan_in or fi, self.fan_out or fo
        p.normal_(0.0, self.gain * math.sqrt(2.0 / float(fi + fo)))


class KaimingUniform(Initializer):
    def __init__(self, fan_in=None, negative_slope=0.0, nonlinearity="relu"):
        self.fan_in, self.negative_slope, self.nonlinearity = fan_in, negative_slope, nonlinearity

    def _fill(self, p):
        fi = self.fan_in or _fans(p.shape)[0]
        gain = torch.nn.init.calculate_gain(self.nonlinearity, self.negative_slope) if self.nonlinearity != "relu" else math.sqrt(2.0)
        limit = gain * math.sqrt(3.0 / float(fi))
        p.uniform_(-limit, limit)


class KaimingNormal(Initializer):
    def __init__(self, fan_in=None, negative_slope=0.0, nonlinearity="relu"):
        self.fan_in, self.negative_slope, self.nonlinearity = fan_in, negative_slope, nonlinearity

    def _fill(self, p):
        fi = self.fan_in or _fans(p.shape)[0]
        gain = torch.nn.init.calculate_gain(self.nonlinearity, self.negative_slope) if self.nonlinearity != "relu" else math.sqrt(2.0)
        p.normal_(0.0, gain / math.sqrt(float(fi)))


class Assign(Initializer):
    def __init__(self, value, name=None):
        self.value = value

    def _fill(self, p):
        v = self.value
        if not isinstance(v, torch.Tensor):
            v = torch.as_tensor(np.asarray(v))
        p.copy_(v.to(device=p.device, dtype=p.dtype).reshape(p.shape))


class Orthogonal(Initializer):
    def __init__(self, gain=1.0, name=None):
        self.gain = gain

    def _fill(self, p):
        torch.nn.init.orthogonal_(p, self.gain)


class Dirac(Initializer):
    def __init__(self, groups=1, name=None):
        self.groups = groups

    def _fill(self, p):
        torch.nn.init.dirac_(p, self.groups)


NumpyArrayInitializer = Assign
ConstantInitializer = Constant
NormalInitializer = Normal
UniformInitializer = Uniform
XavierInitializer = XavierUniform
MSRAInitializer = KaimingNormal
TruncatedNormalInitializer = TruncatedNormal

_global_weight_init = None
_global_bias_init = None


def set_global_initializer(weight_init, bias_init=None):
    global _global_weight_init, _global_bias_init
    _global_weight_init, _global_bias_init = weight_init, bias_init


def calculate_gain(nonlinearity, param=None):
    return torch.nn.init.calculate_gain(nonlinearity, param)


class Bilinear(Initializer):
    """Bilinear-upsampling kernel for transposed convolutions (reference: nn/initializer/bilinear.py)."""

    def _fill(self, p):
        if p.dim() != 4 or p.shape[2] != p.shape[3]:
            raise ValueError("Bilinear initializer needs a 4-D weight with square kernels")
        k = p.shape[3]
        f = (k + 1) // 2
        c = (2 * f - 1 - f % 2) / (2.0 * f)
        og = torch.arange(k, dtype=torch.float32)
        filt = (1 - (og / f - c).abs())
        w = filt[:, None] * filt[None, :]
        p.copy_(w.to(p.dtype).expand_as(p))
