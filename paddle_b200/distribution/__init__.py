"""paddle.distribution (reference: python/paddle/distribution/*.py).

Every class keeps the reference's constructor arguments, property / method set and quirks (for example ``Categorical``
normalises its ``logits`` by their SUM for ``probs`` / ``log_prob`` but by softmax for ``sample`` / ``entropy`` /
``kl_divergence``, categorical.py:90-208; ``Geometric`` counts failures before the first success, so its mean is
``1 / p - 1``).  Sampling and special functions come from torch; ``kl_divergence`` dispatches through a registry that
user code can extend with ``register_kl``.
"""
from __future__ import annotations

import math
import numbers

import numpy as np
import torch
import torch.distributions as TD

from . import transform
from .transform import (  # noqa: F401
    AbsTransform, AffineTransform, ChainTransform, ExpTransform, IndependentTransform, PowerTransform, ReshapeTransform,
    SigmoidTransform, SoftmaxTransform, StackTransform, StickBreakingTransform, TanhTransform, Transform,
)

__all__ = [
    "Bernoulli", "Beta", "Categorical", "Cauchy", "ContinuousBernoulli", "Dirichlet", "Distribution", "Exponential",
    "ExponentialFamily", "Multinomial", "MultivariateNormal", "Normal", "Uniform", "kl_divergence", "register_kl",
    "Independent", "TransformedDistribution", "Laplace", "LogNormal", "Gamma", "Gumbel", "Geometric", "Binomial",
    "Poisson", "StudentT",
]


def _t(x, dtype=None):
    """Python number / list / ndarray / Tensor -> floating tensor."""
    if isinstance(x, torch.Tensor):
        return x if x.dtype.is_floating_point else x.to(torch.get_default_dtype())
    if isinstance(x, np.ndarray) and x.dtype in (np.float32, np.float64):
        return torch.from_numpy(x)
    return torch.as_tensor(x, dtype=dtype or torch.get_default_dtype())


def _broadcast(*xs):
    return torch.broadcast_tensors(*[_t(x) for x in xs])


def _shape(shape):
    if isinstance(shape, torch.Tensor):
        shape = shape.tolist()
    if isinstance(shape, numbers.Integral):
        shape = (shape,)
    return tuple(int(s) for s in shape)


class Distribution:
    """Base class: ``batch_shape`` (independent parameterisations) and ``event_shape`` (shape of one draw)."""

    def __init__(self, batch_shape=(), event_shape=()):
        self._batch_shape = tuple(batch_shape)
        self._event_shape = tuple(event_shape)

    @property
    def batch_shape(self):
        return self._batch_shape

    @property
    def event_shape(self):
        return self._event_shape

    @property
    def mean(self):
        raise NotImplementedError

    @property
    def variance(self):
        raise NotImplementedError

    def sample(self, shape=()):
        raise NotImplementedError

    def rsample(self, shape=()):
        raise NotImplementedError

    def entropy(self):
        raise NotImplementedError

    def kl_divergence(self, other):
        return kl_divergence(self, other)

    def prob(self, value):
        return self.log_prob(value).exp()

    def log_prob(self, value):
        raise NotImplementedError

    def probs(self, value):
        """Old name of ``prob`` (the reference keeps it with a deprecation warning)."""
        return self.prob(value)

    def _extend_shape(self, sample_shape):
        return _shape(sample_shape) + self._batch_shape + self._event_shape


class ExponentialFamily(Distribution):
    """p(x) = h(x) exp(<eta, T(x)> - A(eta)).  Subclasses give ``_natural_parameters``, ``_log_normalizer`` and
    ``_mean_carrier_measure``; entropy and KL then follow from derivatives of A (Bregman divergence)."""

    @property
    def _natural_parameters(self):
        raise NotImplementedError

    def _log_normalizer(self, *natural_parameters):
        raise NotImplementedError

    @property
    def _mean_carrier_measure(self):
        raise NotImplementedError

    def entropy(self):
        nat = [p.detach().clone().requires_grad_(True) for p in self._natural_parameters]
        with torch.enable_grad():
            log_norm = self._log_normalizer(*nat)
            grads = torch.autograd.grad(log_norm.sum(), nat, create_graph=False)
        ent = -self._mean_carrier_measure + log_norm.detach()
        for p, g in zip(nat, grads):
            ent = ent - (p.detach() * g).reshape(self._batch_shape + (-1,)).sum(-1)
        return ent


class _Wrapped(Distribution):
    """A distribution whose math is a ``torch.distributions`` object (``self._base``)."""

    def _set_base(self, base):
        self._base = base
        super().__init__(tuple(base.batch_shape), tuple(base.event_shape))

    @property
    def mean(self):
        return self._base.mean

    @property
    def variance(self):
        return self._base.variance

    @property
    def stddev(self):
        return self._base.variance.sqrt()

    def sample(self, shape=()):
        with torch.no_grad():
            return self._base.sample(_shape(shape))

    def rsample(self, shape=()):
        return self._base.rsample(_shape(shape))

    def log_prob(self, value):
        return self._base.log_prob(_t(value))

    def entropy(self):
        return self._base.entropy()

    def cdf(self, value):
        return self._base.cdf(_t(value))

    def icdf(self, value):
        return self._base.icdf(_t(value))


# ---------------------------------------------------------------------------------------------------------- continuous
class Normal(_Wrapped):
    def __init__(self, loc, scale, name=None):
        self.all_arg_is_float = isinstance(loc, (int, float)) and isinstance(scale, (int, float))
        self.loc, self.scale = _broadcast(loc, scale)
        self.name = name or "Normal"
        self._set_base(TD.Normal(self.loc, self.scale, validate_args=False))

    def sample(self, shape=(), seed=0):
        """Draws of shape ``shape + batch_shape``; ``seed`` != 0 makes the draw reproducible."""
        if seed:
            gen = torch.Generator(device=self.loc.device).manual_seed(int(seed))
            eps = torch.randn(_shape(shape) + tuple(self.loc.shape), generator=gen, dtype=self.loc.dtype, device=self.loc.device)
            return (self.loc + self.scale * eps).detach()
        return super().sample(shape)

    def probs(self, value):
        return self.prob(value)


class LogNormal(_Wrapped):
    """exp(Normal(loc, scale))."""

    def __init__(self, loc, scale, name=None):
        self.loc, self.scale = _broadcast(loc, scale)
        self._normal = Normal(self.loc, self.scale)
        self._set_base(TD.LogNormal(self.loc, self.scale, validate_args=False))

    def probs(self, value):
        return self.prob(value)


class Uniform(_Wrapped):
    def __init__(self, low, high, name=None):
        self.low, self.high = _broadcast(low, high)
        self.name = name or "Uniform"
        self._set_base(TD.Uniform(self.low, self.high, validate_args=False))

    def sample(self, shape=(), seed=0):
        if seed:
            gen = torch.Generator(device=self.low.device).manual_seed(int(seed))
            u = torch.rand(_shape(shape) + tuple(self.low.shape), generator=gen, dtype=self.low.dtype, device=self.low.device)
            return self.low + (self.high - self.low) * u
        return super().sample(shape)

    def log_prob(self, value):
        """log(1[low < value < high]) - log(high - low): -inf outside the OPEN interval."""
        value = _t(value)
        inside = ((self.low < value) & (value < self.high)).to(value.dtype)
        return torch.log(inside) - torch.log(self.high - self.low)

    def probs(self, value):
        value = _t(value)
        inside = ((self.low < value) & (value < self.high)).to(value.dtype)
        return inside / (self.high - self.low)

    prob = probs

    def entropy(self):
        return torch.log(self.high - self.low)


class Laplace(_Wrapped):
    def __init__(self, loc, scale):
        self.loc, self.scale = _broadcast(loc, scale)
        self._set_base(TD.Laplace(self.loc, self.scale, validate_args=False))


class Cauchy(_Wrapped):
    def __init__(self, loc, scale, name=None):
        self.loc, self.scale = _broadcast(loc, scale)
        self.name = name or "Cauchy"
        self._set_base(TD.Cauchy(self.loc, self.scale, validate_args=False))

    @property
    def mean(self):
        raise ValueError("Cauchy distribution has no mean.")

    @property
    def variance(self):
        raise ValueError("Cauchy distribution has no variance.")

    @property
    def stddev(self):
        raise ValueError("Cauchy distribution has no stddev.")


class Gumbel(_Wrapped):
    def __init__(self, loc, scale):
        self.loc, self.scale = _broadcast(loc, scale)
        self._set_base(TD.Gumbel(self.loc, self.scale, validate_args=False))

    def cdf(self, value):
        return torch.exp(-torch.exp(-(_t(value) - self.loc) / self.scale))


class Exponential(_Wrapped, ExponentialFamily):
    def __init__(self, rate):
        self.rate = _t(rate)
        self._set_base(TD.Exponential(self.rate, validate_args=False))

    def entropy(self):
        return 1.0 - torch.log(self.rate)

    @property
    def _natural_parameters(self):
        return (-self.rate,)

    def _log_normalizer(self, x):
        return -torch.log(-x)

    @property
    def _mean_carrier_measure(self):
        return 0.0


class Gamma(_Wrapped, ExponentialFamily):
    def __init__(self, concentration, rate):
        self.concentration, self.rate = _broadcast(concentration, rate)
        self._set_base(TD.Gamma(self.concentration, self.rate, validate_args=False))

    def entropy(self):
        return self._base.entropy()

    @property
    def _natural_parameters(self):
        return (self.concentration - 1, -self.rate)

    def _log_normalizer(self, x, y):
        return torch.lgamma(x + 1) + (x + 1) * torch.log(-y.reciprocal())

    @property
    def _mean_carrier_measure(self):
        return 0.0


class Beta(_Wrapped, ExponentialFamily):
    def __init__(self, alpha, beta):
        self.alpha, self.beta = _broadcast(alpha, beta)
        self._set_base(TD.Beta(self.alpha, self.beta, validate_args=False))

    def entropy(self):
        return self._base.entropy()

    @property
    def _natural_parameters(self):
        return (self.alpha, self.beta)

    def _log_normalizer(self, x, y):
        return torch.lgamma(x) + torch.lgamma(y) - torch.lgamma(x + y)

    @property
    def _mean_carrier_measure(self):
        # T(x) = (log x, log(1 - x)) with natural parameters (alpha, beta) leaves h(x) = 1 / (x (1 - x))
        a, b = self.alpha, self.beta
        return -(torch.digamma(a) + torch.digamma(b) - 2 * torch.digamma(a + b))


class Dirichlet(_Wrapped, ExponentialFamily):
    def __init__(self, concentration):
        self.concentration = _t(concentration)
        if self.concentration.dim() < 1:
            raise ValueError("`concentration` parameter must be at least one dimensional")
        self._set_base(TD.Dirichlet(self.concentration, validate_args=False))

    def entropy(self):
        return self._base.entropy()

    @property
    def _natural_parameters(self):
        return (self.concentration,)

    def _log_normalizer(self, x):
        return x.lgamma().sum(-1) - torch.lgamma(x.sum(-1))

    @property
    def _mean_carrier_measure(self):
        # T(x) = log x with natural parameter alpha leaves h(x) = 1 / prod(x)
        c = self.concentration
        return -(torch.digamma(c) - torch.digamma(c.sum(-1, keepdim=True))).sum(-1)


class StudentT(_Wrapped):
    def __init__(self, df, loc, scale, name=None):
        self.df, self.loc, self.scale = _broadcast(df, loc, scale)
        self.name = name or "StudentT"
        if not bool((self.df > 0).all()) or not bool((self.scale > 0).all()):
            raise ValueError("df and scale must be positive")
        self._set_base(TD.StudentT(self.df, self.loc, self.scale, validate_args=False))


class MultivariateNormal(_Wrapped):
    """Exactly one of ``covariance_matrix`` / ``precision_matrix`` / ``scale_tril`` (lower Cholesky factor)."""

    def __init__(self, loc, covariance_matrix=None, precision_matrix=None, scale_tril=None):
        given = [m is not None for m in (covariance_matrix, precision_matrix, scale_tril)]
        if sum(given) != 1:
            raise ValueError("Exactly one of covariance_matrix or precision_matrix or scale_tril may be specified.")
        self.loc = _t(loc)
        if self.loc.dim() < 1:
            raise ValueError("loc must be at least one-dimensional.")
        kw = {}
        if covariance_matrix is not None:
            kw["covariance_matrix"] = _t(covariance_matrix)
        elif precision_matrix is not None:
            kw["precision_matrix"] = _t(precision_matrix)
        else:
            kw["scale_tril"] = _t(scale_tril)
        if next(iter(kw.values())).dim() < 2:
            raise ValueError("the matrix argument must be at least two-dimensional, with optional leading batch dimensions")
        self._set_base(TD.MultivariateNormal(self.loc, validate_args=False, **kw))

    @property
    def covariance_matrix(self):
        return self._base.covariance_matrix

    @property
    def precision_matrix(self):
        return self._base.precision_matrix

    @property
    def scale_tril(self):
        return self._base.scale_tril


# ------------------------------------------------------------------------------------------------------------ discrete
class Bernoulli(_Wrapped, ExponentialFamily):
    def __init__(self, probs, name=None):
        self.probs_ = _t(probs)
        self.name = name or "Bernoulli"
        eps = torch.finfo(self.probs_.dtype).eps
        self._clipped = self.probs_.clamp(eps, 1 - eps)
        self.logits = torch.log(self._clipped) - torch.log1p(-self._clipped)
        self._set_base(TD.Bernoulli(probs=self.probs_, validate_args=False))

    # the reference exposes the parameter as ``.probs`` AND (through the base class) a ``probs(value)`` method; the
    # attribute wins there, so it does here
    @property
    def probs(self):
        return self.probs_

    def rsample(self, shape=(), temperature=1.0):
        """Relaxed (Gumbel-sigmoid) draws in (0, 1), differentiable w.r.t. ``probs``; they sharpen to {0, 1} as
        ``temperature`` -> 0."""
        shape = self._extend_shape(shape)
        u = torch.rand(shape, dtype=self.probs_.dtype, device=self.probs_.device).clamp(1e-6, 1 - 1e-6)
        noise = torch.log(u) - torch.log1p(-u)
        return torch.sigmoid((self.logits + noise) / temperature)

    def cdf(self, value):
        value = _t(value)
        zeros, ones = torch.zeros_like(self.probs_ + value), torch.ones_like(self.probs_ + value)
        return torch.where(value < 0, zeros, torch.where(value < 1, (1 - self.probs_) + zeros, ones))

    def entropy(self):
        p = self._clipped
        return -(p * torch.log(p) + (1 - p) * torch.log1p(-p))

    @property
    def _natural_parameters(self):
        return (self.logits,)

    def _log_normalizer(self, x):
        return torch.log1p(torch.exp(x))

    @property
    def _mean_carrier_measure(self):
        return 0.0


class ContinuousBernoulli(Distribution):
    """p(x) = C(lambda) lambda^x (1 - lambda)^(1 - x) on [0, 1].  Near lambda = 1/2 (inside ``lims``) the normaliser and
    the moments use their Taylor expansions."""

    def __init__(self, probs, lims=(0.499, 0.501)):
        self.probs = _t(probs)
        eps = torch.finfo(self.probs.dtype).eps
        self.probs = self.probs.clamp(eps, 1 - eps)
        self.lims = lims
        self._base = TD.ContinuousBernoulli(probs=self.probs, lims=tuple(lims), validate_args=False)
        super().__init__(tuple(self.probs.shape), ())

    mean = property(lambda self: self._base.mean)
    variance = property(lambda self: self._base.variance)

    def sample(self, shape=()):
        with torch.no_grad():
            return self._base.sample(_shape(shape))

    def rsample(self, shape=()):
        return self._base.rsample(_shape(shape))

    def log_prob(self, value):
        return self._base.log_prob(_t(value))

    def cdf(self, value):
        return self._base.cdf(_t(value))

    def icdf(self, value):
        return self._base.icdf(_t(value))

    def entropy(self):
        return self._base.entropy()


class Categorical(Distribution):
    """``logits``: non-negative weights.  ``probs`` / ``log_prob`` use weights / sum(weights); ``sample``, ``entropy``
    and ``kl_divergence`` use softmax(logits) - the reference's behaviour (categorical.py:90, 130, 170, 204)."""

    def __init__(self, logits, name=None):
        self.logits = _t(logits)
        self.name = name or "Categorical"
        self._prob = self.logits / self.logits.sum(-1, keepdim=True)
        super().__init__(tuple(self.logits.shape[:-1]), ())

    def sample(self, shape=()):
        shape = _shape(shape)
        n = int(np.prod(shape)) if shape else 1
        p = torch.softmax(self.logits, -1)
        flat = p.reshape(-1, p.shape[-1])
        with torch.no_grad():
            draws = torch.multinomial(flat, n, replacement=True)            # [batch, n]
        return draws.t().reshape(shape + tuple(self.logits.shape[:-1]))

    def entropy(self):
        logits = self.logits - self.logits.max(-1, keepdim=True).values
        z = logits.exp().sum(-1, keepdim=True)
        p = logits.exp() / z
        return -(p * (logits - torch.log(z))).sum(-1)

    def probs(self, value):
        value = torch.as_tensor(value).long()
        if self._prob.dim() == 1:
            return self._prob[value.reshape(-1)].reshape(value.shape)
        if value.dim() == 1:                       # the same categories for every batch entry
            return self._prob[..., value]
        return torch.gather(self._prob, -1, value)

    prob = probs

    def log_prob(self, value):
        return torch.log(self.probs(value))


class Multinomial(_Wrapped):
    def __init__(self, total_count, probs):
        if not isinstance(total_count, int) or total_count < 1:
            raise ValueError("input parameter total_count must be int type and grater than zero.")
        self.probs = _t(probs)
        if self.probs.dim() < 1:
            raise ValueError("probs parameter should not be none and over one dimension")
        self.probs = self.probs / self.probs.sum(-1, keepdim=True)
        self.total_count = total_count
        self._categorical = Categorical(self.probs)
        self._set_base(TD.Multinomial(total_count, probs=self.probs, validate_args=False))

    def entropy(self):
        """n * H(categorical) - lgamma(n + 1) + sum_k E[lgamma(x_k + 1)], the expectation taken under Binomial(n, p_k)."""
        n = self.total_count
        support = torch.arange(1, n + 1, dtype=self.probs.dtype).reshape((-1,) + (1,) * self.probs.dim())
        binom = TD.Binomial(n, probs=self.probs, validate_args=False).log_prob(support).exp()
        cat_entropy = -(self.probs * torch.log(self.probs.clamp_min(torch.finfo(self.probs.dtype).tiny))).sum(-1)
        return n * cat_entropy - math.lgamma(n + 1) + (binom * torch.lgamma(support + 1)).sum((0, -1))


class Binomial(Distribution):
    def __init__(self, total_count, probs):
        self.total_count, self.probs = torch.broadcast_tensors(torch.as_tensor(total_count), _t(probs))
        if not bool((self.total_count >= 0).all()):
            raise ValueError("total_count must be non-negative")
        self._base = TD.Binomial(self.total_count, probs=self.probs, validate_args=False)
        super().__init__(tuple(self.probs.shape), ())

    @property
    def mean(self):
        return self.total_count * self.probs

    @property
    def variance(self):
        return self.total_count * self.probs * (1 - self.probs)

    def sample(self, shape=()):
        with torch.no_grad():
            return self._base.sample(_shape(shape))

    def log_prob(self, value):
        return self._base.log_prob(_t(value))

    def entropy(self):
        """-sum_k p(k) log p(k) over the whole support 0..max(n)."""
        n = int(self.total_count.max())
        support = torch.arange(0, n + 1, dtype=self.probs.dtype).reshape((-1,) + (1,) * self.probs.dim())
        logp = self._base.log_prob(support.clamp(max=self.total_count.to(self.probs.dtype)))
        valid = support <= self.total_count
        return -torch.where(valid, logp.exp() * logp, torch.zeros_like(logp)).sum(0)


class Geometric(Distribution):
    """Number of failures before the first success: P(X = k) = (1 - p)^k p, k = 0, 1, ..."""

    def __init__(self, probs):
        if isinstance(probs, (int, float)):
            probs = float(probs)
        self.probs = _t(probs)
        if not bool(((self.probs > 0) & (self.probs <= 1)).all()):
            raise ValueError("Expected probs in (0, 1]")
        super().__init__(tuple(self.probs.shape), ())

    @property
    def mean(self):
        return 1.0 / self.probs - 1.0

    @property
    def variance(self):
        return (1.0 / self.probs - 1.0) / self.probs

    @property
    def stddev(self):
        return self.variance.sqrt()

    def pmf(self, k):
        if not isinstance(k, (numbers.Integral, torch.Tensor)):
            raise TypeError(f"Expected type of k is number.Real|framework.Variable, but got {type(k)}")
        return torch.pow(1.0 - self.probs, _t(k)) * self.probs

    def log_pmf(self, k):
        if not isinstance(k, (numbers.Integral, torch.Tensor)):
            raise TypeError(f"Expected type of k is number.Real|framework.Variable, but got {type(k)}")
        return torch.log(self.pmf(k))

    prob, log_prob = pmf, log_pmf

    def sample(self, shape=()):
        with torch.no_grad():
            return self.rsample(shape)

    def rsample(self, shape=()):
        shape = self._extend_shape(shape)
        u = torch.rand(shape, dtype=self.probs.dtype, device=self.probs.device).clamp_min(torch.finfo(self.probs.dtype).tiny)
        return torch.floor(torch.log(u) / torch.log1p(-self.probs))

    def entropy(self):
        x = (1.0 - self.probs) * torch.log(1.0 - self.probs)
        y = self.probs * torch.log(self.probs)
        return -(x + y) / self.probs

    def cdf(self, k):
        if not isinstance(k, (numbers.Integral, torch.Tensor)):
            raise TypeError(f"Expected type of k is number.Real|framework.Variable, but got {type(k)}")
        return 1.0 - torch.pow(1.0 - self.probs, _t(k) + 1)


class Poisson(_Wrapped):
    def __init__(self, rate):
        self.rate = _t(rate)
        if not bool((self.rate >= 0).all()):
            raise ValueError("rate must be non-negative")
        self._set_base(TD.Poisson(self.rate, validate_args=False))

    def entropy(self):
        """-sum_k p(k) log p(k) over [0, rate + 30 sqrt(rate)] - the mass beyond is below float precision."""
        upper = int(torch.ceil(self.rate.max() + 30 * self.rate.max().sqrt()).item()) + 2
        support = torch.arange(0, upper, dtype=self.rate.dtype).reshape((-1,) + (1,) * self.rate.dim())
        logp = self._base.log_prob(support)
        return -(logp.exp() * logp).sum(0)


# -------------------------------------------------------------------------------------------------------- combinators
def _sum_rightmost(x, n):
    return x.sum(list(range(-n, 0))) if n > 0 else x


class Independent(Distribution):
    """Reinterpret the ``reinterpreted_batch_rank`` rightmost batch dims of ``base`` as event dims: ``log_prob`` and
    ``entropy`` are summed over them."""

    def __init__(self, base, reinterpreted_batch_rank):
        if not isinstance(base, Distribution):
            raise TypeError(f"Expected type of 'base' is Distribution, but got {type(base)}")
        if not 0 < reinterpreted_batch_rank <= len(base.batch_shape):
            raise ValueError(f"Expected 0 < reinterpreted_batch_rank <= {len(base.batch_shape)}, but got {reinterpreted_batch_rank}")
        self._base, self._reinterpreted_batch_rank = base, reinterpreted_batch_rank
        shape = base.batch_shape + base.event_shape
        split = len(base.batch_shape) - reinterpreted_batch_rank
        super().__init__(shape[:split], shape[split:])

    @property
    def mean(self):
        return self._base.mean

    @property
    def variance(self):
        return self._base.variance

    def sample(self, shape=()):
        return self._base.sample(shape)

    def rsample(self, shape=()):
        return self._base.rsample(shape)

    def log_prob(self, value):
        return _sum_rightmost(self._base.log_prob(value), self._reinterpreted_batch_rank)

    def prob(self, value):
        return self.log_prob(value).exp()

    def entropy(self):
        return _sum_rightmost(self._base.entropy(), self._reinterpreted_batch_rank)


class TransformedDistribution(Distribution):
    """The law of ``t_n(...t_1(x))`` for ``x ~ base``: log p(y) = log p_base(x) - sum_i log|det J_{t_i}|."""

    def __init__(self, base, transforms):
        if not isinstance(base, Distribution):
            raise TypeError(f"Expected type of 'base' is Distribution, but got {type(base)}.")
        if not isinstance(transforms, (list, tuple)):
            raise TypeError(f"Expected type of 'transforms' is Sequence[Transform] or Chain, but got {type(transforms)}.")
        if not all(isinstance(t, Transform) for t in transforms):
            raise TypeError("All element of transforms must be Transform type.")
        chain = ChainTransform(transforms)
        base_shape = base.batch_shape + base.event_shape
        if len(base_shape) < chain._domain_event_rank:
            raise ValueError(f"'base' needs to have shape with size at least {chain._domain_event_rank}, but got {len(base_shape)}.")
        if chain._domain_event_rank > len(base.event_shape):
            base = Independent(base, chain._domain_event_rank - len(base.event_shape))
        self._base, self._transforms = base, list(transforms)
        out_shape = chain.forward_shape(base.batch_shape + base.event_shape)
        event_rank = chain._codomain_event_rank + max(len(base.event_shape) - chain._domain_event_rank, 0)
        super().__init__(out_shape[:len(out_shape) - event_rank], out_shape[len(out_shape) - event_rank:])

    def sample(self, shape=()):
        x = self._base.sample(shape)
        for t in self._transforms:
            x = t.forward(x)
        return x

    def rsample(self, shape=()):
        x = self._base.rsample(shape)
        for t in self._transforms:
            x = t.forward(x)
        return x

    def log_prob(self, value):
        log_prob = 0.0
        y = _t(value)
        event_rank = len(self.event_shape)
        for t in reversed(self._transforms):
            x = t.inverse(y)
            event_rank += t._domain_event_rank - t._codomain_event_rank
            log_prob = log_prob - _sum_rightmost(t.forward_log_det_jacobian(x), event_rank - t._domain_event_rank)
            y = x
        return log_prob + _sum_rightmost(self._base.log_prob(y), event_rank - len(self._base.event_shape))


# ------------------------------------------------------------------------------------------------------------------ KL
_REGISTER_TABLE = {}


def register_kl(cls_p, cls_q):
    """Decorator: ``@register_kl(P, Q) def kl_p_q(p, q): ...``.  ``kl_divergence`` picks the most specific registered
    pair along the two classes' MROs and refuses an ambiguous choice."""
    if not (isinstance(cls_p, type) and issubclass(cls_p, Distribution) and isinstance(cls_q, type) and issubclass(cls_q, Distribution)):
        raise TypeError("cls_p and cls_q must be subclass of Distribution")

    def decorator(f):
        _REGISTER_TABLE[cls_p, cls_q] = f
        _dispatch_cache.clear()
        return f
    return decorator


_dispatch_cache = {}


def _dispatch(cls_p, cls_q):
    key = (cls_p, cls_q)
    if key in _dispatch_cache:
        return _dispatch_cache[key]
    matches = [(sp, sq) for sp, sq in _REGISTER_TABLE if issubclass(cls_p, sp) and issubclass(cls_q, sq)]
    if not matches:
        raise NotImplementedError(f"No KL(p || q) is registered for p type {cls_p.__name__} and q type {cls_q.__name__}")

    def rank(pair):
        return (cls_p.__mro__.index(pair[0]), cls_q.__mro__.index(pair[1]))
    left = min(matches, key=lambda m: (rank(m)[0], rank(m)[1]))
    right = min(matches, key=lambda m: (rank(m)[1], rank(m)[0]))
    if _REGISTER_TABLE[left] is not _REGISTER_TABLE[right]:
        raise RuntimeWarning(f"Ambiguous kl_divergence({cls_p.__name__}, {cls_q.__name__}). Please register_kl("
                             f"{left[0].__name__}, {right[1].__name__})")
    _dispatch_cache[key] = _REGISTER_TABLE[left]
    return _REGISTER_TABLE[left]


def kl_divergence(p, q):
    """KL(p || q) through the registry."""
    return _dispatch(type(p), type(q))(p, q)


def _torch_kl(p, q):
    return TD.kl_divergence(p._base, q._base)


for _cls in (Normal, Uniform, Laplace, Cauchy, Beta, Dirichlet, Gamma, Exponential, Bernoulli, MultivariateNormal, Poisson,
             LogNormal, Gumbel):
    register_kl(_cls, _cls)(_torch_kl)


@register_kl(ContinuousBernoulli, ContinuousBernoulli)
def _kl_cb_cb(p, q):
    return TD.kl_divergence(p._base, q._base)


@register_kl(Binomial, Binomial)
def _kl_binomial_binomial(p, q):
    if not bool((p.total_count == q.total_count).all()):
        raise NotImplementedError("KL between Binomials with different total_count is not supported")
    n = p.total_count.to(p.probs.dtype)
    return n * (p.probs * (torch.log(p.probs) - torch.log(q.probs)) + (1 - p.probs) * (torch.log1p(-p.probs) - torch.log1p(-q.probs)))


@register_kl(Categorical, Categorical)
def _kl_categorical_categorical(p, q):
    lp = p.logits - p.logits.max(-1, keepdim=True).values
    lq = q.logits - q.logits.max(-1, keepdim=True).values
    zp, zq = lp.exp().sum(-1, keepdim=True), lq.exp().sum(-1, keepdim=True)
    prob = lp.exp() / zp
    return (prob * (lp - torch.log(zp) - lq + torch.log(zq))).sum(-1)


@register_kl(Geometric, Geometric)
def _kl_geometric_geometric(p, q):
    return p.probs * torch.log(p.probs / q.probs) + (1.0 - p.probs) * torch.log((1.0 - p.probs) / (1.0 - q.probs))


@register_kl(ExponentialFamily, ExponentialFamily)
def _kl_expfamily_expfamily(p, q):
    """Bregman divergence of the log normaliser: A(eta_q) - A(eta_p) - <eta_q - eta_p, grad A(eta_p)>."""
    if type(p) is not type(q):
        raise NotImplementedError
    p_nat = [t.detach().clone().requires_grad_(True) for t in p._natural_parameters]
    q_nat = [t.detach() for t in q._natural_parameters]
    with torch.enable_grad():
        p_log_norm = p._log_normalizer(*p_nat)
        grads = torch.autograd.grad(p_log_norm.sum(), p_nat)
    kl = q._log_normalizer(*q_nat) - p_log_norm.detach()
    for pn, qn, g in zip(p_nat, q_nat, grads):
        term = (qn - pn.detach()) * g
        kl = kl - _sum_rightmost(term, len(q.event_shape))
    return kl


for _cls in (Normal, Uniform, Laplace, Cauchy, Bernoulli, Categorical, Beta, Dirichlet, Gamma, Exponential, MultivariateNormal,
             Poisson, LogNormal, Gumbel, Geometric, Binomial, ContinuousBernoulli):
    _cls.kl_divergence = lambda self, other: kl_divergence(self, other)
