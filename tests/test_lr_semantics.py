"""Learning-rate schedules against the closed forms and state machines in python/paddle/optimizer/lr.py."""
import math

import numpy as np
import pytest

import paddle
from paddle.optimizer import lr


def _run(s, n):
    out = []
    for _ in range(n):
        out.append(float(s()))
        s.step()
    return out


def test_closed_form_schedules():
    np.testing.assert_allclose(_run(lr.NoamDecay(d_model=16, warmup_steps=4, learning_rate=1.0), 6),
                               [16 ** -0.5 * min(1.0 if e == 0 else e ** -0.5, 4 ** -1.5 * e) for e in range(6)], rtol=1e-6)
    np.testing.assert_allclose(_run(lr.NaturalExpDecay(0.5, 0.1), 4), [0.5 * math.exp(-0.1 * e) for e in range(4)], rtol=1e-6)
    np.testing.assert_allclose(_run(lr.InverseTimeDecay(0.5, 0.1), 4), [0.5 / (1 + 0.1 * e) for e in range(4)], rtol=1e-6)
    np.testing.assert_allclose(_run(lr.PolynomialDecay(0.5, 5, end_lr=0.1, power=2.0), 8),
                               [0.4 * (1 - min(e, 5) / 5) ** 2 + 0.1 for e in range(8)], rtol=1e-6)
    # cycle=True stretches the horizon to the next multiple of decay_steps
    want = []
    for e in range(12):
        horizon = 5 * (1 if e == 0 else math.ceil(e / 5))
        want.append(0.4 * (1 - e / horizon) + 0.1)
    np.testing.assert_allclose(_run(lr.PolynomialDecay(0.5, 5, end_lr=0.1, power=1.0, cycle=True), 12), want, rtol=1e-6)
    assert _run(lr.PiecewiseDecay([3, 6], [1.0, 0.5, 0.1]), 8) == [1.0, 1.0, 1.0, 0.5, 0.5, 0.5, 0.1, 0.1]
    np.testing.assert_allclose(_run(lr.StepDecay(0.5, 3, 0.1), 7), [0.5 * 0.1 ** (e // 3) for e in range(7)], rtol=1e-6)
    np.testing.assert_allclose(_run(lr.MultiStepDecay(0.5, [2, 4], 0.1), 6), [0.5, 0.5, 0.05, 0.05, 0.005, 0.005], rtol=1e-6)
    np.testing.assert_allclose(_run(lr.LinearLR(0.5, 4, 0.25, 1.0), 7), [0.5 * (0.25 + 0.75 * min(e, 4) / 4) for e in range(7)], rtol=1e-6)
    np.testing.assert_allclose(_run(lr.CosineAnnealingDecay(0.5, 5, 0.0), 12),
                               [0.25 * (1 + math.cos(math.pi * e / 5)) for e in range(12)], atol=1e-6)
    # SGDR: periods 3, 6, 12, ...
    want, e, period = [], 0, 3
    for _ in range(12):
        want.append(0.25 * (1 + math.cos(math.pi * e / period)))
        e += 1
        if e == period:
            e, period = 0, period * 2
    np.testing.assert_allclose(_run(lr.CosineAnnealingWarmRestarts(0.5, 3, 2), 12), want, atol=1e-6)


def test_warmup_wraps_a_schedule_and_restarts_its_clock():
    got = _run(lr.LinearWarmup(lr.ExponentialDecay(0.5, 0.9), 3, 0.0, 0.5), 7)
    np.testing.assert_allclose(got, [0.0, 0.5 / 3, 1.0 / 3, 0.5, 0.45, 0.405, 0.3645], rtol=1e-6)


def test_cyclic_and_one_cycle():
    np.testing.assert_allclose(_run(lr.CyclicLR(0.1, 1.0, 3, 3), 10), [0.1, 0.4, 0.7, 1.0, 0.7, 0.4, 0.1, 0.4, 0.7, 1.0], rtol=1e-6)
    np.testing.assert_allclose(_run(lr.CyclicLR(0.1, 1.0, 2, 2, mode="triangular2"), 9),
                               [0.1, 0.55, 1.0, 0.55, 0.1, 0.325, 0.55, 0.325, 0.1], rtol=1e-6)
    one = _run(lr.OneCycleLR(1.0, 10), 10)                   # divide_factor 25, end 1e-4, 30 % warm-up, cosine
    assert one[0] == pytest.approx(0.04) and one[2] == pytest.approx(1.0) and one[-1] == pytest.approx(1e-4)
    np.testing.assert_allclose(one[3], 1e-4 + (1.0 - 1e-4) * (1 + math.cos(math.pi / 7)) / 2, rtol=1e-6)
    s = lr.OneCycleLR(1.0, 3)
    for _ in range(3):
        s.step()
    with pytest.raises(ValueError):
        s.step()


def test_reduce_on_plateau_state_machine_and_resume():
    """python/paddle/optimizer/lr.py:1501-1554 -- nothing is evaluated while cooling down."""
    s = lr.ReduceOnPlateau(1.0, patience=1, factor=0.5, cooldown=1)
    seen = []
    for m in [1, 1, 1, 1, 1, 1, 0.5, 0.5, 0.5, 0.5]:
        s.step(m)
        seen.append(float(s()))
    assert seen == [1.0, 1.0, 0.5, 0.5, 0.5, 0.25, 0.25, 0.25, 0.25, 0.125]
    state = s.state_dict()
    assert set(state) == {"cooldown_counter", "best", "num_bad_epochs", "last_epoch", "last_lr"}
    twin = lr.ReduceOnPlateau(1.0, patience=1, factor=0.5, cooldown=1)
    twin.set_state_dict(state)
    for m in [0.5, 0.5, 0.5]:
        s.step(m)
        twin.step(m)
        assert float(twin()) == float(s())
    # relative margin keeps the sign of best: with a negative best, "min" needs current < best * (1 - t)
    neg = lr.ReduceOnPlateau(1.0, patience=0, factor=0.5, threshold=0.1)
    neg.step(-1.0)
    neg.step(-1.0)                       # -1.0 < -1.0 + 0.1 = -0.9 : better by the reference's rule
    assert float(neg()) == 1.0
    s.step(np.array([0.3]))
    s.step(paddle.to_tensor(0.3))
    with pytest.raises(TypeError):
        s.step("0.3")
    with pytest.raises(RuntimeError):
        twin.set_state_dict({"last_epoch": 3})
