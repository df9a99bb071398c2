"""paddle.distribution.transform (reference: python/paddle/distribution/transform.py): invertible maps with
log-det-Jacobians, used by ``TransformedDistribution``.

A transform declares how many rightmost dims of its input / output form one event (``_domain_event_rank`` /
``_codomain_event_rank``): ``forward_log_det_jacobian`` returns one number per event, i.e. it has the input's shape minus
the event dims.  ``inverse_log_det_jacobian(y) = -forward_log_det_jacobian(inverse(y))``.
"""
from __future__ import annotations

import enum
import math

import torch
import torch.nn.functional as TF

__all__ = ["Transform", "AbsTransform", "AffineTransform", "ChainTransform", "ExpTransform", "IndependentTransform",
           "PowerTransform", "ReshapeTransform", "SigmoidTransform", "SoftmaxTransform", "StackTransform",
           "StickBreakingTransform", "TanhTransform"]


class Type(enum.Enum):
    BIJECTION = "bijection"
    INJECTION = "injection"
    SURJECTION = "surjection"
    OTHER = "other"

    @classmethod
    def is_injective(cls, _type):
        return _type in (cls.BIJECTION, cls.INJECTION)


def _t(x):
    return x if isinstance(x, torch.Tensor) else torch.as_tensor(x, dtype=torch.get_default_dtype())


def _sum_rightmost(x, n):
    return x.sum(list(range(-n, 0))) if n > 0 else x


class Transform:
    _type = Type.INJECTION
    _domain_event_rank = 0
    _codomain_event_rank = 0

    @classmethod
    def _is_injective(cls):
        return Type.is_injective(cls._type)

    def __call__(self, input):
        """A transform applied to a distribution gives the transformed distribution; to a transform, their chain."""
        from . import Distribution, TransformedDistribution
        if isinstance(input, Distribution):
            return TransformedDistribution(input, [self])
        if isinstance(input, Transform):
            return ChainTransform([self, input])
        return self.forward(input)

    def forward(self, x):
        if not isinstance(x, torch.Tensor):
            raise TypeError(f"Expected 'x' is a Tensor or Real, but got {type(x)}.")
        if x.dim() < self._domain_event_rank:
            raise ValueError(f"The dimensions of x({x.dim()}) should be grater than or equal to {self._domain_event_rank}")
        return self._forward(x)

    def inverse(self, y):
        if not isinstance(y, torch.Tensor):
            raise TypeError(f"Expected 'y' is a Tensor or Real, but got {type(y)}.")
        if y.dim() < self._codomain_event_rank:
            raise ValueError(f"The dimensions of y({y.dim()}) should be grater than or equal to {self._codomain_event_rank}")
        return self._inverse(y)

    def forward_log_det_jacobian(self, x):
        if not isinstance(x, torch.Tensor):
            raise TypeError(f"Expected 'x' is a Tensor or Real, but got {type(x)}.")
        if not self._is_injective():
            raise NotImplementedError("forward_log_det_jacobian can't be implemented for non-injective transforms.")
        return self._call_forward_log_det_jacobian(x)

    def inverse_log_det_jacobian(self, y):
        if not isinstance(y, torch.Tensor):
            raise TypeError(f"Expected 'y' is a Tensor or Real, but got {type(y)}.")
        return self._call_inverse_log_det_jacobian(y)

    def forward_shape(self, shape):
        return self._forward_shape(tuple(shape))

    def inverse_shape(self, shape):
        return self._inverse_shape(tuple(shape))

    # --- to override
    def _forward(self, x):
        raise NotImplementedError("Forward not implemented")

    def _inverse(self, y):
        raise NotImplementedError("Inverse not implemented")

    def _call_forward_log_det_jacobian(self, x):
        if type(self)._forward_log_det_jacobian is not Transform._forward_log_det_jacobian:
            return self._forward_log_det_jacobian(x)
        if type(self)._inverse_log_det_jacobian is not Transform._inverse_log_det_jacobian:
            return -self._inverse_log_det_jacobian(self.forward(x))
        raise NotImplementedError("Neither _forward_log_det_jacobian nor _inverse_log_det_jacobian is implemented.")

    def _call_inverse_log_det_jacobian(self, y):
        if type(self)._inverse_log_det_jacobian is not Transform._inverse_log_det_jacobian:
            return self._inverse_log_det_jacobian(y)
        if type(self)._forward_log_det_jacobian is not Transform._forward_log_det_jacobian:
            return -self._forward_log_det_jacobian(self._inverse(y))
        raise NotImplementedError("Neither _forward_log_det_jacobian nor _inverse_log_det_jacobian is implemented")

    def _forward_log_det_jacobian(self, x):
        raise NotImplementedError

    def _inverse_log_det_jacobian(self, y):
        raise NotImplementedError

    def _forward_shape(self, shape):
        return shape

    def _inverse_shape(self, shape):
        return shape


class AbsTransform(Transform):
    """y = |x|.  Not injective: ``inverse`` returns both pre-images (-y, y); its log-det is (0, 0)."""
    _type = Type.SURJECTION

    def _forward(self, x):
        return x.abs()

    def _inverse(self, y):
        return -y, y

    def _inverse_log_det_jacobian(self, y):
        zero = torch.zeros([], dtype=y.dtype, device=y.device)
        return zero, zero


class AffineTransform(Transform):
    """y = loc + scale * x."""
    _type = Type.BIJECTION

    def __init__(self, loc, scale):
        if not isinstance(loc, torch.Tensor):
            raise TypeError(f"Expected 'loc' is a Tensor, but got {type(loc)}")
        if not isinstance(scale, torch.Tensor):
            raise TypeError(f"Expected scale is a Tensor, but got {type(scale)}")
        self._loc, self._scale = loc, scale

    @property
    def loc(self):
        return self._loc

    @property
    def scale(self):
        return self._scale

    def _forward(self, x):
        return self._loc + self._scale * x

    def _inverse(self, y):
        return (y - self._loc) / self._scale

    def _forward_log_det_jacobian(self, x):
        return torch.log(torch.abs(self._scale)).expand(torch.broadcast_shapes(x.shape, self._scale.shape))

    def _forward_shape(self, shape):
        return tuple(torch.broadcast_shapes(shape, self._loc.shape, self._scale.shape))

    _inverse_shape = _forward_shape


class ExpTransform(Transform):
    _type = Type.BIJECTION

    def _forward(self, x):
        return x.exp()

    def _inverse(self, y):
        return y.log()

    def _forward_log_det_jacobian(self, x):
        return x


class PowerTransform(Transform):
    """y = x ** power on the positive reals."""
    _type = Type.BIJECTION

    def __init__(self, power):
        if not isinstance(power, torch.Tensor):
            raise TypeError(f"Expected 'power' is a tensor, but got {type(power)}")
        self._power = power

    @property
    def power(self):
        return self._power

    def _forward(self, x):
        return x.pow(self._power)

    def _inverse(self, y):
        return y.pow(1.0 / self._power)

    def _forward_log_det_jacobian(self, x):
        return torch.log(torch.abs(self._power * x.pow(self._power - 1)))

    def _forward_shape(self, shape):
        return tuple(torch.broadcast_shapes(shape, self._power.shape))

    _inverse_shape = _forward_shape


class SigmoidTransform(Transform):
    _type = Type.BIJECTION

    def _forward(self, x):
        return torch.sigmoid(x)

    def _inverse(self, y):
        return y.log() - (-y).log1p()

    def _forward_log_det_jacobian(self, x):
        return -TF.softplus(-x) - TF.softplus(x)


class TanhTransform(Transform):
    _type = Type.BIJECTION

    def _forward(self, x):
        return torch.tanh(x)

    def _inverse(self, y):
        return torch.atanh(y)

    def _forward_log_det_jacobian(self, x):
        # log(1 - tanh(x)^2) in the overflow-free form
        return 2.0 * (math.log(2.0) - x - TF.softplus(-2.0 * x))


class SoftmaxTransform(Transform):
    """y = softmax(x) over the last dim; the "inverse" is log(y) (softmax is invariant to shifts, so not injective)."""
    _type = Type.OTHER
    _domain_event_rank = 1
    _codomain_event_rank = 1

    def _forward(self, x):
        x = (x - x.max(-1, keepdim=True).values).exp()
        return x / x.sum(-1, keepdim=True)

    def _inverse(self, y):
        return y.log()

    def _forward_shape(self, shape):
        if len(shape) < 1:
            raise ValueError(f"Expected length of shape is grater than 1, but got {len(shape)}")
        return shape

    _inverse_shape = _forward_shape


class StickBreakingTransform(Transform):
    """R^{k} -> the open simplex in R^{k+1} by stick breaking."""
    _type = Type.BIJECTION
    _domain_event_rank = 1
    _codomain_event_rank = 1

    def _forward(self, x):
        offset = x.shape[-1] + 1 - torch.ones([x.shape[-1]], dtype=x.dtype, device=x.device).cumsum(-1)
        z = torch.sigmoid(x - offset.log())
        z_cumprod = (1 - z).cumprod(-1)
        return TF.pad(z, (0, 1), value=1.0) * TF.pad(z_cumprod, (1, 0), value=1.0)

    def _inverse(self, y):
        y_crop = y[..., :-1]
        offset = y.shape[-1] - torch.ones([y_crop.shape[-1]], dtype=y.dtype, device=y.device).cumsum(-1)
        sf = 1 - y_crop.cumsum(-1)
        return y_crop.log() - sf.log() + offset.log()

    def _forward_log_det_jacobian(self, x):
        y = self.forward(x)
        offset = x.shape[-1] + 1 - torch.ones([x.shape[-1]], dtype=x.dtype, device=x.device).cumsum(-1)
        x = x - offset.log()
        return (-x + TF.logsigmoid(x) + y[..., :-1].log()).sum(-1)

    def _forward_shape(self, shape):
        if not shape:
            raise ValueError(f"Expected 'shape' is not empty, but got {shape}")
        return shape[:-1] + (shape[-1] + 1,)

    def _inverse_shape(self, shape):
        if not shape:
            raise ValueError(f"Expected 'shape' is not empty, but got {shape}")
        return shape[:-1] + (shape[-1] - 1,)


class ChainTransform(Transform):
    """transforms[-1] ∘ ... ∘ transforms[0]; the log-dets add, each summed down to the chain's event rank."""

    def __init__(self, transforms):
        if not isinstance(transforms, (list, tuple)):
            raise TypeError(f"Expected type of 'transforms' is Sequence, but got {type(transforms)}")
        if not all(isinstance(t, Transform) for t in transforms):
            raise TypeError("All elements of transforms should be Transform type.")
        self.transforms = list(transforms)
        # event ranks of the whole chain: the largest rank any member needs, carried to either end
        self._domain_event_rank = self._rank_at_start()
        self._codomain_event_rank = self._rank_at_end()

    def _rank_at_start(self):
        need = 0
        for t in reversed(self.transforms):
            need = max(need - (t._codomain_event_rank - t._domain_event_rank), t._domain_event_rank)
        return max(need, 0)

    def _rank_at_end(self):
        need = 0
        for t in self.transforms:
            need = max(need + (t._codomain_event_rank - t._domain_event_rank), t._codomain_event_rank)
        return max(need, 0)

    def _is_injective(self):
        return all(t._is_injective() for t in self.transforms)

    def _forward(self, x):
        for t in self.transforms:
            x = t.forward(x)
        return x

    def _inverse(self, y):
        for t in reversed(self.transforms):
            y = t.inverse(y)
        return y

    def _forward_log_det_jacobian(self, x):
        value = 0.0
        event_rank = self._domain_event_rank
        for t in self.transforms:
            value = value + _sum_rightmost(t.forward_log_det_jacobian(x), event_rank - t._domain_event_rank)
            x = t.forward(x)
            event_rank += t._codomain_event_rank - t._domain_event_rank
        return value

    def _forward_shape(self, shape):
        for t in self.transforms:
            shape = t.forward_shape(shape)
        return shape

    def _inverse_shape(self, shape):
        for t in reversed(self.transforms):
            shape = t.inverse_shape(shape)
        return shape


class IndependentTransform(Transform):
    """Treat ``reinterpreted_batch_rank`` extra rightmost dims as part of the event: their log-dets are summed."""

    def __init__(self, base, reinterpreted_batch_rank):
        if not isinstance(base, Transform):
            raise TypeError(f"Expected 'base' is Transform type, but get {type(base)}")
        if reinterpreted_batch_rank <= 0:
            raise ValueError(f"Expected 'reinterpreted_batch_rank' is grater than zero, but got {reinterpreted_batch_rank}")
        self._base, self._reinterpreted_batch_rank = base, reinterpreted_batch_rank
        self._domain_event_rank = base._domain_event_rank + reinterpreted_batch_rank
        self._codomain_event_rank = base._codomain_event_rank + reinterpreted_batch_rank

    def _is_injective(self):
        return self._base._is_injective()

    def _forward(self, x):
        if x.dim() < self._domain_event_rank:
            raise ValueError("Input dimensions is less than event dimensions.")
        return self._base.forward(x)

    def _inverse(self, y):
        if y.dim() < self._codomain_event_rank:
            raise ValueError("Input dimensions is less than event dimensions.")
        return self._base.inverse(y)

    def _forward_log_det_jacobian(self, x):
        return _sum_rightmost(self._base.forward_log_det_jacobian(x), self._reinterpreted_batch_rank)

    def _forward_shape(self, shape):
        return self._base.forward_shape(shape)

    def _inverse_shape(self, shape):
        return self._base.inverse_shape(shape)


class ReshapeTransform(Transform):
    """Reshape the event part: ``in_event_shape`` -> ``out_event_shape`` (same number of elements)."""
    _type = Type.BIJECTION

    def __init__(self, in_event_shape, out_event_shape):
        if not isinstance(in_event_shape, (list, tuple)) or not isinstance(out_event_shape, (list, tuple)):
            raise TypeError(f"Expected type of 'in_event_shape' and 'out_event_shape' is Sequence[int], but got "
                            f"{type(in_event_shape)} and {type(out_event_shape)}")
        if math.prod(in_event_shape) != math.prod(out_event_shape):
            raise ValueError(f"The numel of 'in_event_shape' should be 'out_event_shape', but got "
                             f"{math.prod(in_event_shape)}!={math.prod(out_event_shape)}")
        self._in_event_shape, self._out_event_shape = tuple(in_event_shape), tuple(out_event_shape)
        self._domain_event_rank, self._codomain_event_rank = len(in_event_shape), len(out_event_shape)

    @property
    def in_event_shape(self):
        return self._in_event_shape

    @property
    def out_event_shape(self):
        return self._out_event_shape

    def _forward(self, x):
        return x.reshape(tuple(x.shape[:x.dim() - len(self._in_event_shape)]) + self._out_event_shape)

    def _inverse(self, y):
        return y.reshape(tuple(y.shape[:y.dim() - len(self._out_event_shape)]) + self._in_event_shape)

    def _forward_shape(self, shape):
        n = len(self._in_event_shape)
        if len(shape) < n:
            raise ValueError(f"Expected length of 'shape' is not less than {n}, but got {len(shape)}")
        if tuple(shape[len(shape) - n:]) != self._in_event_shape:
            raise ValueError(f"Event shape mismatch, expected: {self._in_event_shape}, but got {shape[len(shape) - n:]}")
        return tuple(shape[:len(shape) - n]) + self._out_event_shape

    def _inverse_shape(self, shape):
        n = len(self._out_event_shape)
        if len(shape) < n:
            raise ValueError(f"Expected 'shape' length is not less than {n}, but got {len(shape)}")
        if tuple(shape[len(shape) - n:]) != self._out_event_shape:
            raise ValueError(f"Event shape mismatch, expected: {self._out_event_shape}, but got {shape[len(shape) - n:]}")
        return tuple(shape[:len(shape) - n]) + self._in_event_shape

    def _forward_log_det_jacobian(self, x):
        return torch.zeros(tuple(x.shape[:x.dim() - len(self._in_event_shape)]), dtype=x.dtype, device=x.device)


class StackTransform(Transform):
    """Apply ``transforms[i]`` to slice ``i`` of ``axis``."""

    def __init__(self, transforms, axis=0):
        if not transforms or not isinstance(transforms, (list, tuple)):
            raise TypeError(f"Expected 'transforms' is Sequence[Transform], but got {type(transforms)}.")
        if not all(isinstance(t, Transform) for t in transforms):
            raise TypeError("Expected all element in transforms is Transform Type.")
        if not isinstance(axis, int):
            raise TypeError(f"Expected 'axis' is int, but got{type(axis)}.")
        self._transforms, self._axis = list(transforms), axis

    def _is_injective(self):
        return all(t._is_injective() for t in self._transforms)

    @property
    def transforms(self):
        return self._transforms

    @property
    def axis(self):
        return self._axis

    def _check(self, v):
        if not (-v.dim() <= self._axis < v.dim()):
            raise ValueError(f"Input dimensions {v.dim()} should be grater than stack transform axis {self._axis}.")
        if v.shape[self._axis] != len(self._transforms):
            raise ValueError(f"Input size along {self._axis} should be equal to the length of transforms.")

    def _map(self, v, name):
        self._check(v)
        return torch.stack([getattr(t, name)(s) for s, t in zip(v.unbind(self._axis), self._transforms)], self._axis)

    def _forward(self, x):
        return self._map(x, "forward")

    def _inverse(self, y):
        return self._map(y, "inverse")

    def _forward_log_det_jacobian(self, x):
        return self._map(x, "forward_log_det_jacobian")
