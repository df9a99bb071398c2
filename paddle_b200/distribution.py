"""paddle.distribution (reference: python/paddle/distribution/*.py): the common distributions on torch.distributions
with Paddle's constructor argument names and method set (sample / rsample / log_prob / prob / entropy / kl_divergence)."""
from __future__ import annotations

import torch
import torch.distributions as D


class Distribution:
    def __init__(self, base):
        self._base = base

    @property
    def mean(self):
        return self._base.mean

    @property
    def variance(self):
        return self._base.variance

    def sample(self, shape=()):
        return self._base.sample(tuple(shape))

    def rsample(self, shape=()):
        return self._base.rsample(tuple(shape))

    def log_prob(self, value):
        return self._base.log_prob(torch.as_tensor(value))

    def prob(self, value):
        return self.log_prob(value).exp()

    def probs(self, value):
        return self.prob(value)

    def entropy(self):
        return self._base.entropy()

    def kl_divergence(self, other):
        return D.kl_divergence(self._base, other._base)


def _t(x):
    return x if isinstance(x, torch.Tensor) else torch.as_tensor(x, dtype=torch.get_default_dtype())


class Normal(Distribution):
    def __init__(self, loc, scale, name=None):
        self.loc, self.scale = _t(loc), _t(scale)
        super().__init__(D.Normal(self.loc, self.scale))


class LogNormal(Distribution):
    def __init__(self, loc, scale, name=None):
        super().__init__(D.LogNormal(_t(loc), _t(scale)))


class Uniform(Distribution):
    def __init__(self, low, high, name=None):
        self.low, self.high = _t(low), _t(high)
        super().__init__(D.Uniform(self.low, self.high))


class Bernoulli(Distribution):
    def __init__(self, probs, name=None):
        super().__init__(D.Bernoulli(probs=_t(probs)))


class Categorical(Distribution):
    def __init__(self, logits, name=None):
        # Paddle's Categorical takes unnormalised probabilities called ``logits`` and normalises them by their sum
        self.logits = _t(logits)
        super().__init__(D.Categorical(probs=self.logits / self.logits.sum(-1, keepdim=True)))


class Multinomial(Distribution):
    def __init__(self, total_count, probs, name=None):
        super().__init__(D.Multinomial(total_count, probs=_t(probs)))


class Beta(Distribution):
    def __init__(self, alpha, beta, name=None):
        super().__init__(D.Beta(_t(alpha), _t(beta)))


class Dirichlet(Distribution):
    def __init__(self, concentration, name=None):
        super().__init__(D.Dirichlet(_t(concentration)))


class Gamma(Distribution):
    def __init__(self, concentration, rate, name=None):
        super().__init__(D.Gamma(_t(concentration), _t(rate)))


class Exponential(Distribution):
    def __init__(self, rate, name=None):
        super().__init__(D.Exponential(_t(rate)))


class Laplace(Distribution):
    def __init__(self, loc, scale, name=None):
        super().__init__(D.Laplace(_t(loc), _t(scale)))


class Gumbel(Distribution):
    def __init__(self, loc, scale, name=None):
        super().__init__(D.Gumbel(_t(loc), _t(scale)))


class Geometric(Distribution):
    def __init__(self, probs, name=None):
        super().__init__(D.Geometric(probs=_t(probs)))


class Poisson(Distribution):
    def __init__(self, rate, name=None):
        super().__init__(D.Poisson(_t(rate)))


class Cauchy(Distribution):
    def __init__(self, loc, scale, name=None):
        super().__init__(D.Cauchy(_t(loc), _t(scale)))


class StudentT(Distribution):
    def __init__(self, df, loc, scale, name=None):
        super().__init__(D.StudentT(_t(df), _t(loc), _t(scale)))


class MultivariateNormal(Distribution):
    def __init__(self, loc, covariance_matrix=None, precision_matrix=None, scale_tril=None, name=None):
        super().__init__(D.MultivariateNormal(_t(loc), covariance_matrix=covariance_matrix, precision_matrix=precision_matrix,
                                              scale_tril=scale_tril))


def kl_divergence(p, q):
    return p.kl_divergence(q)
