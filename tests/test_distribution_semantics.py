"""paddle.distribution against scipy.stats and the reference's documented quirks (python/paddle/distribution/*.py)."""
import math

import numpy as np
import pytest
import scipy.stats as st
import torch

import paddle  # noqa: F401
import paddle.distribution as D
from paddle.distribution import transform as T


def _np(t):
    return t.detach().numpy()


def test_continuous_against_scipy():
    x = torch.tensor([0.3, 1.2, 2.5])
    cases = [
        (D.Normal(0.5, 2.0), st.norm(0.5, 2.0)),
        (D.Laplace(0.5, 2.0), st.laplace(0.5, 2.0)),
        (D.Cauchy(0.5, 2.0), st.cauchy(0.5, 2.0)),
        (D.Gumbel(0.5, 2.0), st.gumbel_r(0.5, 2.0)),
        (D.LogNormal(0.5, 0.7), st.lognorm(0.7, scale=math.exp(0.5))),
        (D.Exponential(1.5), st.expon(scale=1 / 1.5)),
        (D.Gamma(2.0, 3.0), st.gamma(2.0, scale=1 / 3.0)),
        (D.StudentT(4.0, 0.5, 2.0), st.t(4.0, 0.5, 2.0)),
        (D.Uniform(0.0, 3.0), st.uniform(0.0, 3.0)),
    ]
    for d, ref in cases:
        name = type(d).__name__
        np.testing.assert_allclose(_np(d.log_prob(x)), ref.logpdf(_np(x)), rtol=1e-5, atol=1e-6, err_msg=name)
        np.testing.assert_allclose(_np(d.prob(x)), ref.pdf(_np(x)), rtol=1e-5, atol=1e-6, err_msg=name)
        np.testing.assert_allclose(float(d.entropy()), ref.entropy(), rtol=1e-5, err_msg=name)
        if name != "Cauchy":
            np.testing.assert_allclose(float(d.mean), ref.mean(), rtol=1e-5, err_msg=name)
            np.testing.assert_allclose(float(d.variance), ref.var(), rtol=1e-5, err_msg=name)
        if hasattr(d, "cdf") and name not in ("StudentT",):
            np.testing.assert_allclose(_np(d.cdf(x)), ref.cdf(_np(x)), rtol=1e-5, atol=1e-6, err_msg=name)
        s = d.sample([2000])
        assert s.shape == (2000,) and not s.requires_grad
    with pytest.raises(ValueError):
        D.Cauchy(0.0, 1.0).mean
    b = D.Beta(torch.tensor([1.0, 3.0]), torch.tensor([2.0, 0.5]))
    xb = torch.tensor([0.2, 0.7])
    np.testing.assert_allclose(_np(b.log_prob(xb)), st.beta([1.0, 3.0], [2.0, 0.5]).logpdf([0.2, 0.7]), rtol=1e-5)
    np.testing.assert_allclose(_np(b.entropy()), st.beta([1.0, 3.0], [2.0, 0.5]).entropy(), rtol=1e-5)
    assert b.batch_shape == (2,) and b.event_shape == ()
    dr = D.Dirichlet(torch.tensor([1.0, 2.0, 3.0]))
    xs = torch.tensor([0.2, 0.3, 0.5])
    np.testing.assert_allclose(float(dr.log_prob(xs)), st.dirichlet([1.0, 2.0, 3.0]).logpdf(_np(xs)), rtol=1e-5)
    np.testing.assert_allclose(float(dr.entropy()), st.dirichlet([1.0, 2.0, 3.0]).entropy(), rtol=1e-5)
    assert dr.batch_shape == () and dr.event_shape == (3,) and dr.sample([4]).shape == (4, 3)
    cov = np.array([[2.0, 0.3], [0.3, 1.0]], dtype="float32")
    mv = D.MultivariateNormal(torch.tensor([0.5, -1.0]), covariance_matrix=torch.from_numpy(cov))
    ref = st.multivariate_normal([0.5, -1.0], cov)
    np.testing.assert_allclose(float(mv.log_prob(torch.tensor([0.1, 0.2]))), ref.logpdf([0.1, 0.2]), rtol=1e-5)
    np.testing.assert_allclose(float(mv.entropy()), ref.entropy(), rtol=1e-5)
    mv2 = D.MultivariateNormal(torch.tensor([0.5, -1.0]), scale_tril=torch.linalg.cholesky(torch.from_numpy(cov)))
    torch.testing.assert_close(mv2.covariance_matrix, torch.from_numpy(cov), rtol=1e-5, atol=1e-6)
    assert mv.event_shape == (2,) and mv.rsample([5]).shape == (5, 2)
    with pytest.raises(ValueError):
        D.MultivariateNormal(torch.zeros(2))
    # Uniform: open interval, -inf outside; Normal: seeded draws repeat
    u = D.Uniform(0.0, 2.0)
    assert float(u.log_prob(torch.tensor(2.5))) == float("-inf") and float(u.probs(torch.tensor(1.0))) == 0.5
    n = D.Normal(torch.zeros(3), torch.ones(3))
    torch.testing.assert_close(n.sample([2], seed=7), n.sample([2], seed=7))
    assert n.sample([2]).shape == (2, 3) and n.rsample([2]).shape == (2, 3)


def test_discrete_against_scipy_and_quirks():
    b = D.Bernoulli(torch.tensor([0.3, 0.8]))
    torch.testing.assert_close(b.probs, torch.tensor([0.3, 0.8]))
    np.testing.assert_allclose(_np(b.log_prob(torch.tensor([1.0, 0.0]))), [math.log(0.3), math.log(0.2)], rtol=1e-5)
    np.testing.assert_allclose(_np(b.entropy()), st.bernoulli([0.3, 0.8]).entropy(), rtol=1e-5)
    np.testing.assert_allclose(_np(b.cdf(torch.tensor([0.5, -1.0]))), [0.7, 0.0], rtol=1e-6)
    r = b.rsample([1000], temperature=0.1)
    assert r.shape == (1000, 2) and float(r.min()) >= 0 and float(r.max()) <= 1
    assert abs(float((r[:, 0] > 0.5).float().mean()) - 0.3) < 0.06
    # Categorical: probs by SUM normalisation, entropy / kl / sampling by softmax
    w = torch.tensor([1.0, 2.0, 3.0])
    c = D.Categorical(w)
    torch.testing.assert_close(c.probs(torch.tensor([0, 2])), torch.tensor([1 / 6, 0.5]))
    torch.testing.assert_close(c.log_prob(torch.tensor([1])), torch.log(torch.tensor([2 / 6])))
    sm = torch.softmax(w, -1)
    torch.testing.assert_close(c.entropy(), -(sm * sm.log()).sum())
    other = D.Categorical(torch.tensor([1.0, 1.0, 1.0]))
    torch.testing.assert_close(c.kl_divergence(other), (sm * (sm.log() - math.log(1 / 3))).sum())
    draws = c.sample([4000])
    assert draws.shape == (4000,) and abs(float((draws == 2).float().mean()) - float(sm[2])) < 0.04
    cb = D.Categorical(torch.tensor([[1.0, 1.0], [1.0, 3.0]]))
    assert cb.batch_shape == (2,) and cb.sample([5]).shape == (5, 2)
    torch.testing.assert_close(cb.probs(torch.tensor([1])), torch.tensor([[0.5], [0.75]]))
    # Geometric counts failures before the first success
    g = D.Geometric(0.3)
    ref = st.geom(0.3, loc=-1)
    np.testing.assert_allclose(float(g.mean), ref.mean(), rtol=1e-6)
    np.testing.assert_allclose(float(g.variance), ref.var(), rtol=1e-6)
    np.testing.assert_allclose(float(g.pmf(2)), ref.pmf(2), rtol=1e-6)
    np.testing.assert_allclose(float(g.cdf(2)), ref.cdf(2), rtol=1e-6)
    np.testing.assert_allclose(float(g.entropy()), ref.entropy(), rtol=1e-5)
    assert float(g.sample([100]).min()) >= 0
    with pytest.raises(TypeError):
        g.pmf(1.5)
    # Binomial / Poisson / Multinomial
    bi = D.Binomial(10, torch.tensor([0.3, 0.6]))
    np.testing.assert_allclose(_np(bi.log_prob(torch.tensor([3.0, 7.0]))), st.binom(10, [0.3, 0.6]).logpmf([3, 7]), rtol=1e-5)
    np.testing.assert_allclose(_np(bi.entropy()), st.binom(10, [0.3, 0.6]).entropy(), rtol=1e-4)
    np.testing.assert_allclose(_np(bi.mean), [3.0, 6.0], rtol=1e-6)
    po = D.Poisson(torch.tensor(3.0))
    np.testing.assert_allclose(float(po.log_prob(torch.tensor(4.0))), st.poisson(3.0).logpmf(4), rtol=1e-5)
    np.testing.assert_allclose(float(po.entropy()), st.poisson(3.0).entropy(), rtol=1e-4)
    mu = D.Multinomial(5, torch.tensor([2.0, 3.0, 5.0]))                  # probs are normalised
    ref = st.multinomial(5, [0.2, 0.3, 0.5])
    np.testing.assert_allclose(float(mu.log_prob(torch.tensor([1.0, 1.0, 3.0]))), ref.logpmf([1, 1, 3]), rtol=1e-5)
    np.testing.assert_allclose(float(mu.entropy()), ref.entropy(), rtol=1e-4)
    assert mu.sample([6]).shape == (6, 3) and float(mu.sample([6]).sum(-1).min()) == 5
    with pytest.raises(ValueError):
        D.Multinomial(0, torch.tensor([0.5, 0.5]))
    # ContinuousBernoulli integrates to one and has a closed-form mean away from 1/2
    cbern = D.ContinuousBernoulli(torch.tensor(0.3))
    grid = torch.linspace(1e-4, 1 - 1e-4, 20001)
    np.testing.assert_allclose(float(torch.trapezoid(cbern.prob(grid), grid)), 1.0, rtol=1e-3)
    lam = 0.3
    np.testing.assert_allclose(float(cbern.mean), lam / (2 * lam - 1) + 1 / (2 * math.atanh(1 - 2 * lam)), rtol=1e-4)
    torch.testing.assert_close(cbern.icdf(cbern.cdf(torch.tensor(0.4))), torch.tensor(0.4), rtol=1e-4, atol=1e-5)


def test_kl_registry_and_exponential_family():
    pairs = [
        (D.Normal(0.0, 1.0), D.Normal(1.0, 2.0), math.log(2.0) + (1 + 1) / 8 - 0.5),
        (D.Bernoulli(torch.tensor(0.3)), D.Bernoulli(torch.tensor(0.6)), 0.3 * math.log(0.3 / 0.6) + 0.7 * math.log(0.7 / 0.4)),
        (D.Exponential(2.0), D.Exponential(0.5), math.log(2.0 / 0.5) + 0.5 / 2.0 - 1),
        (D.Geometric(0.3), D.Geometric(0.6), 0.3 * math.log(0.3 / 0.6) + 0.7 * math.log(0.7 / 0.4)),
        (D.Uniform(0.0, 1.0), D.Uniform(-1.0, 2.0), math.log(3.0)),
        (D.Laplace(0.0, 1.0), D.Laplace(0.0, 1.0), 0.0),
    ]
    for p, q, want in pairs:
        np.testing.assert_allclose(float(D.kl_divergence(p, q)), want, rtol=1e-5, atol=1e-7, err_msg=type(p).__name__)
        np.testing.assert_allclose(float(p.kl_divergence(q)), want, rtol=1e-5, atol=1e-7)
    # numerically: KL = E_p[log p - log q]
    for p, q in [(D.Beta(2.0, 3.0), D.Beta(1.5, 1.0)), (D.Gamma(2.0, 3.0), D.Gamma(1.5, 1.0)),
                 (D.Cauchy(0.0, 1.0), D.Cauchy(1.0, 2.0)), (D.Gumbel(0.0, 1.0), D.Gumbel(0.5, 2.0))]:
        torch.manual_seed(0)
        x = p.sample([200000])
        mc = float((p.log_prob(x) - q.log_prob(x)).mean())
        np.testing.assert_allclose(float(D.kl_divergence(p, q)), mc, rtol=0.05, atol=5e-3, err_msg=type(p).__name__)
    bp, bq = D.Binomial(8, torch.tensor(0.3)), D.Binomial(8, torch.tensor(0.5))
    ks = torch.arange(0, 9.0)
    want = float((bp.prob(ks) * (bp.log_prob(ks) - bq.log_prob(ks))).sum())
    np.testing.assert_allclose(float(D.kl_divergence(bp, bq)), want, rtol=1e-5)
    with pytest.raises(NotImplementedError):
        D.kl_divergence(D.Normal(0.0, 1.0), D.Uniform(0.0, 1.0))
    # exponential-family machinery: entropy and Bregman KL from the log normaliser equal the closed forms
    from paddle.distribution import _kl_expfamily_expfamily
    for d in (D.Beta(torch.tensor([1.0, 3.0]), torch.tensor([2.0, 0.5])), D.Dirichlet(torch.tensor([[1.0, 2.0, 3.0], [0.5, 0.5, 2.0]])),
              D.Gamma(2.0, 3.0), D.Exponential(torch.tensor([0.5, 2.0])), D.Bernoulli(torch.tensor([0.2, 0.7]))):
        torch.testing.assert_close(D.ExponentialFamily.entropy(d), d.entropy(), rtol=1e-4, atol=1e-5)
    p, q = D.Dirichlet(torch.tensor([1.0, 2.0, 3.0])), D.Dirichlet(torch.tensor([2.0, 2.0, 2.0]))
    torch.testing.assert_close(_kl_expfamily_expfamily(p, q), D.kl_divergence(p, q), rtol=1e-4, atol=1e-5)

    # user registration: the most specific pair wins; a subclass falls back to its parent's rule
    class MyNormal(D.Normal):
        pass

    np.testing.assert_allclose(float(D.kl_divergence(MyNormal(0.0, 1.0), D.Normal(1.0, 2.0))), pairs[0][2], rtol=1e-5)

    @D.register_kl(MyNormal, D.Normal)
    def _mine(p, q):
        return torch.tensor(-1.0)

    assert float(D.kl_divergence(MyNormal(0.0, 1.0), D.Normal(1.0, 2.0))) == -1.0
    with pytest.raises(TypeError):
        D.register_kl(int, D.Normal)


def test_transforms_independent_and_transformed_distribution():
    x = torch.randn(4, 3, dtype=torch.float64)
    loc, scale = torch.tensor(0.5, dtype=torch.float64), torch.tensor(-2.0, dtype=torch.float64)
    bijections = [T.AffineTransform(loc, scale), T.ExpTransform(), T.SigmoidTransform(), T.TanhTransform(),
                  T.ChainTransform([T.AffineTransform(loc, scale), T.TanhTransform()])]
    for t in bijections:
        y = t.forward(x)
        torch.testing.assert_close(t.inverse(y), x, rtol=1e-6, atol=1e-8)
        # log|dy/dx| by autograd (elementwise maps: the Jacobian is diagonal)
        xr = x.clone().requires_grad_(True)
        (g,) = torch.autograd.grad(t.forward(xr).sum(), xr)
        torch.testing.assert_close(t.forward_log_det_jacobian(x), g.abs().log(), rtol=1e-6, atol=1e-8)
        torch.testing.assert_close(t.inverse_log_det_jacobian(y), -g.abs().log(), rtol=1e-6, atol=1e-8)
    pw = T.PowerTransform(torch.tensor(3.0, dtype=torch.float64))
    xp = x.abs() + 0.1
    torch.testing.assert_close(pw.inverse(pw.forward(xp)), xp)
    torch.testing.assert_close(pw.forward_log_det_jacobian(xp), torch.log(3 * xp ** 2))
    ab = T.AbsTransform()
    assert ab.inverse(torch.tensor(2.0))[0] == -2.0 and not ab._is_injective()
    with pytest.raises(NotImplementedError):
        ab.forward_log_det_jacobian(x)
    sm = T.SoftmaxTransform()
    torch.testing.assert_close(sm.forward(x).sum(-1), torch.ones(4, dtype=torch.float64))
    sb = T.StickBreakingTransform()
    y = sb.forward(x)
    assert y.shape == (4, 4) and sb.forward_shape((4, 3)) == (4, 4) and sb.inverse_shape((4, 4)) == (4, 3)
    torch.testing.assert_close(y.sum(-1), torch.ones(4, dtype=torch.float64))
    torch.testing.assert_close(sb.inverse(y), x, rtol=1e-6, atol=1e-8)
    jac = torch.autograd.functional.jacobian(lambda v: sb.forward(v)[:-1], x[0])      # 3 free outputs
    torch.testing.assert_close(sb.forward_log_det_jacobian(x)[0], torch.linalg.slogdet(jac)[1], rtol=1e-6, atol=1e-8)
    rs = T.ReshapeTransform((3,), (1, 3))
    assert rs.forward(x).shape == (4, 1, 3) and rs.inverse(rs.forward(x)).shape == (4, 3) and rs.forward_shape((4, 3)) == (4, 1, 3)
    assert rs.forward_log_det_jacobian(x).shape == (4,)
    with pytest.raises(ValueError):
        T.ReshapeTransform((3,), (2, 2))
    it = T.IndependentTransform(T.ExpTransform(), 1)
    torch.testing.assert_close(it.forward_log_det_jacobian(x), x.sum(-1))
    stk = T.StackTransform([T.ExpTransform(), T.AffineTransform(loc, scale), T.TanhTransform()], axis=1)
    torch.testing.assert_close(stk.forward(x)[:, 0], x[:, 0].exp())
    torch.testing.assert_close(stk.inverse(stk.forward(x)), x, rtol=1e-6, atol=1e-8)
    torch.testing.assert_close(stk.forward_log_det_jacobian(x)[:, 1], torch.full((4,), math.log(2.0), dtype=torch.float64))
    # a transform called on a distribution / transform
    ln = T.ExpTransform()(D.Normal(0.0, 1.0))
    assert isinstance(ln, D.TransformedDistribution)
    assert isinstance(T.ExpTransform()(T.TanhTransform()), T.ChainTransform)
    v = torch.tensor([0.5, 2.0])
    torch.testing.assert_close(ln.log_prob(v), D.LogNormal(0.0, 1.0).log_prob(v))
    assert ln.sample([7]).shape == (7,) and float(ln.sample([100]).min()) > 0
    # affine of a normal is a normal
    td = D.TransformedDistribution(D.Normal(0.0, 1.0), [T.AffineTransform(torch.tensor(1.0), torch.tensor(3.0))])
    torch.testing.assert_close(td.log_prob(v), D.Normal(1.0, 3.0).log_prob(v))
    # event-rank bookkeeping: stick breaking over a batch of normals -> a distribution over the simplex
    base = D.Normal(torch.zeros(5, 3), torch.ones(5, 3))
    simplex = D.TransformedDistribution(base, [T.StickBreakingTransform()])
    assert simplex.batch_shape == (5,) and simplex.event_shape == (4,)
    s = simplex.sample([2])
    assert s.shape == (2, 5, 4) and simplex.log_prob(s).shape == (2, 5)
    # Independent
    ind = D.Independent(D.Normal(torch.zeros(3, 2), torch.ones(3, 2)), 1)
    assert ind.batch_shape == (3,) and ind.event_shape == (2,)
    z = torch.randn(3, 2)
    torch.testing.assert_close(ind.log_prob(z), D.Normal(0.0, 1.0).log_prob(z).sum(-1))
    torch.testing.assert_close(ind.entropy(), D.Normal(0.0, 1.0).entropy() * 2 * torch.ones(3))
    with pytest.raises(ValueError):
        D.Independent(D.Normal(0.0, 1.0), 1)
