"""Update rules, regularisers and skipped steps against float64 restatements of the reference's formulas
(python/paddle/optimizer/*.py docstrings, paddle/phi/kernels/*/{adagrad,rmsprop,adadelta,adamax,lamb,rprop,asgd}_kernel*)."""
import numpy as np
import pytest
import torch

import paddle
import paddle.optimizer as O
from paddle.regularizer import L1Decay, L2Decay

P0 = np.array([1.0, -2.0, 3.0, -4.0])


def _param(reg=None):
    p = paddle.create_parameter([4], "float32")
    with torch.no_grad():
        p.copy_(torch.tensor(P0, dtype=torch.float32))
    if reg is not None:
        p.regularizer = reg
    return p


def _run(opt, p, n=3):
    for i in range(n):
        ((p ** 2).sum() * 0.5 + p.sum() * 0.1 * i).backward()      # grad = p + 0.1 i
        opt.step()
        opt.clear_grad()
    return p.detach().numpy().astype(np.float64)


def _reg(reg, q):
    if reg is None:
        return 0 * q
    if isinstance(reg, float):
        return reg * q
    return reg._coeff * np.sign(q) if isinstance(reg, L1Decay) else reg._coeff * q


def _sim(rule, reg=None, n=3):
    q, st = P0.copy(), {}
    for i in range(n):
        q = rule(q, q + 0.1 * i + _reg(reg, q), st, i)
    return q


def _sgd(q, g, st, i):
    return q - 0.1 * g


def _momentum(q, g, st, i):
    st["v"] = 0.9 * st.get("v", 0) + g
    return q - 0.1 * st["v"]


def _nesterov(q, g, st, i):
    st["v"] = 0.9 * st.get("v", 0) + g
    return q - 0.1 * (g + 0.9 * st["v"])


def _adagrad(q, g, st, i):
    st["m"] = st.get("m", 0.5) + g * g
    return q - 0.1 * g / (np.sqrt(st["m"]) + 1e-6)


def _adam(q, g, st, i):
    st["m"] = 0.9 * st.get("m", 0) + 0.1 * g
    st["v"] = 0.999 * st.get("v", 0) + 0.001 * g * g
    t = i + 1
    return q - 0.1 * (st["m"] / (1 - 0.9 ** t)) / (np.sqrt(st["v"] / (1 - 0.999 ** t)) + 1e-8)


def _rmsprop(centered):
    def rule(q, g, st, i):
        st["ms"] = 0.9 * st.get("ms", 0) + 0.1 * g * g
        if centered:
            st["mg"] = 0.9 * st.get("mg", 0) + 0.1 * g
            d = np.sqrt(st["ms"] - st["mg"] ** 2 + 1e-6)
        else:
            d = np.sqrt(st["ms"] + 1e-6)
        st["mom"] = 0.5 * st.get("mom", 0) + 0.1 * g / d
        return q - st["mom"]
    return rule


def _adadelta(q, g, st, i):
    st["eg"] = 0.95 * st.get("eg", 0) + 0.05 * g * g
    u = -np.sqrt((st.get("eu", 0) + 1e-6) / (st["eg"] + 1e-6)) * g
    st["eu"] = 0.95 * st.get("eu", 0) + 0.05 * u * u
    return q + u


def _adamax(q, g, st, i):
    st["m"] = 0.9 * st.get("m", 0) + 0.1 * g
    st["u"] = np.maximum(0.999 * st.get("u", 0) + 1e-8, np.abs(g))
    return q - 0.1 / (1 - 0.9 ** (i + 1)) * st["m"] / st["u"]


def _lamb(q, g, st, i):
    st["m"] = 0.9 * st.get("m", 0) + 0.1 * g
    st["v"] = 0.999 * st.get("v", 0) + 0.001 * g * g
    t = i + 1
    r = (st["m"] / (1 - 0.9 ** t)) / (np.sqrt(st["v"] / (1 - 0.999 ** t)) + 1e-6) + 0.01 * q
    return q - 0.1 * np.linalg.norm(q) / np.linalg.norm(r) * r


def _rprop(q, g, st, i):
    prev, lrs = st.get("prev", np.zeros(4)), st.get("lr", np.full(4, 0.01))
    s = g * prev
    lrs = np.where(s > 0, np.minimum(lrs * 1.2, 50), np.where(s < 0, np.maximum(lrs * 0.5, 1e-5), lrs))
    g = np.where(s < 0, 0, g)
    st["prev"], st["lr"] = g, lrs
    return q - np.sign(g) * lrs


def _asgd(q, g, st, i):
    ys, d = st.setdefault("ys", np.zeros((2, 4))), st.get("d", np.zeros(4))
    d = d - ys[i % 2] + g
    ys[i % 2] = g
    st["d"] = d
    return q - 0.1 * d / min(i + 1, 2)


@pytest.mark.parametrize("name,make,rule,steps", [
    ("sgd", lambda p: O.SGD(0.1, parameters=[p]), _sgd, 3),
    ("momentum", lambda p: O.Momentum(0.1, 0.9, parameters=[p]), _momentum, 3),
    ("nesterov", lambda p: O.Momentum(0.1, 0.9, parameters=[p], use_nesterov=True), _nesterov, 3),
    ("adagrad", lambda p: O.Adagrad(0.1, parameters=[p], initial_accumulator_value=0.5), _adagrad, 3),
    ("adam", lambda p: O.Adam(0.1, parameters=[p]), _adam, 3),
    ("rmsprop", lambda p: O.RMSProp(0.1, rho=0.9, momentum=0.5, epsilon=1e-6, parameters=[p]), _rmsprop(False), 3),
    ("rmsprop_centered", lambda p: O.RMSProp(0.1, rho=0.9, momentum=0.5, epsilon=1e-6, centered=True, parameters=[p]), _rmsprop(True), 3),
    ("adadelta", lambda p: O.Adadelta(1.0, rho=0.95, epsilon=1e-6, parameters=[p]), _adadelta, 3),
    ("adamax", lambda p: O.Adamax(0.1, parameters=[p]), _adamax, 3),
    ("lamb", lambda p: O.Lamb(0.1, lamb_weight_decay=0.01, parameters=[p]), _lamb, 3),
    ("rprop", lambda p: O.Rprop(0.01, parameters=[p]), _rprop, 4),
    ("asgd", lambda p: O.ASGD(0.1, batch_num=2, parameters=[p]), _asgd, 5),
])
def test_update_rules(name, make, rule, steps):
    p = _param()
    got = _run(make(p), p, steps)
    np.testing.assert_allclose(got, _sim(rule, n=steps), rtol=2e-5, atol=2e-5)


@pytest.mark.parametrize("cls,rule", [(O.SGD, _sgd), (O.Momentum, _momentum), (O.Adam, _adam), (O.Adamax, _adamax),
                                      (O.ASGD, None)])
@pytest.mark.parametrize("wd,own", [(0.05, None), (L1Decay(0.05), None), (L2Decay(0.05), None), (0.05, L1Decay(0.2)),
                                    (None, L2Decay(0.2)), (L1Decay(0.05), L2Decay(0.3))])
def test_regulariser_reaches_the_gradient(cls, rule, wd, own):
    """python/paddle/optimizer/optimizer.py append_regularization_ops: a float is L2, L1Decay adds coeff * sign(p),
    and a parameter's own regulariser replaces the optimizer's."""
    p = _param(own)
    if cls is O.ASGD:
        opt, rule = cls(0.1, batch_num=2, parameters=[p], weight_decay=wd), _asgd
    else:
        opt = cls(0.1, parameters=[p], weight_decay=wd)
    got = _run(opt, p)
    np.testing.assert_allclose(got, _sim(rule, own if own is not None else wd), rtol=2e-5, atol=2e-5)


def test_adamw_keeps_decay_decoupled_and_adds_a_parameters_own_regulariser():
    p = _param(L2Decay(0.2))
    got = _run(O.AdamW(0.1, parameters=[p], weight_decay=0.05), p)
    q, st = P0.copy(), {}
    for i in range(3):
        g = q + 0.1 * i + 0.2 * q
        q = _adam(q * (1 - 0.1 * 0.05), g, st, i)
    np.testing.assert_allclose(got, q, rtol=2e-5, atol=2e-5)
    for kw, ratio, decay in [(dict(lr_ratio=lambda t: 0.5), 0.5, True), (dict(apply_decay_param_fun=lambda n: False), 1.0, False)]:
        p = _param()
        got = _run(O.AdamW(0.1, parameters=[p], weight_decay=0.05, **kw), p)
        q, st = P0.copy(), {}
        for i in range(3):
            g = q + 0.1 * i
            if decay:
                q = q * (1 - 0.1 * ratio * 0.05)
            st["m"] = 0.9 * st.get("m", 0) + 0.1 * g
            st["v"] = 0.999 * st.get("v", 0) + 0.001 * g * g
            q = q - 0.1 * ratio * (st["m"] / (1 - 0.9 ** (i + 1))) / (np.sqrt(st["v"] / (1 - 0.999 ** (i + 1))) + 1e-8)
        np.testing.assert_allclose(got, q, rtol=2e-5, atol=2e-5)


@pytest.mark.parametrize("cls", [O.Adagrad, O.RMSProp, O.Adadelta, O.Adamax, O.NAdam, O.RAdam, O.ASGD, O.Rprop, O.Lamb,
                                 O.SGD, O.Momentum, O.Adam, O.AdamW])
def test_step_is_skipped_when_the_scaler_found_inf(cls):
    p = _param()
    opt = cls(learning_rate=0.1, parameters=[p])
    (p ** 2).sum().backward()
    opt._found_inf = torch.ones(1)
    opt.step()
    np.testing.assert_array_equal(p.detach().numpy(), P0.astype(np.float32))
    opt._found_inf = torch.zeros(1)
    opt.step()
    assert np.abs(p.detach().numpy() - P0).max() > 0


def test_asgd_state_round_trip():
    p = _param()
    opt = O.ASGD(0.1, batch_num=2, parameters=[p])
    _run(opt, p, 3)
    state = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in opt.state_dict().items()}   # a snapshot
    assert {k.rsplit("_", 2)[-2] for k in state if k.endswith("_0")} >= {"d", "y", "m"}
    q = _param()
    with torch.no_grad():
        q.copy_(p.detach())
    q.name = p.name
    twin = O.ASGD(0.1, batch_num=2, parameters=[q])
    twin.set_state_dict(state)
    for o, t in ((opt, p), (twin, q)):
        (t ** 2).sum().backward()
        o.step()
    torch.testing.assert_close(q.detach(), p.detach())
    with pytest.raises(ValueError):
        O.ASGD(0.1, batch_num=0, parameters=[p])
