"""Learning-rate schedulers against the closed forms documented in the reference (python/paddle/optimizer/lr.py)."""
import math

import numpy as np
import pytest

import paddle_b200 as paddle
from paddle_b200.optimizer import lr as L


def run(s, n):
    out = []
    for _ in range(n):
        out.append(s())
        s.step()
    return out


def test_closed_forms():
    np.testing.assert_allclose(run(L.NoamDecay(d_model=64, warmup_steps=4, learning_rate=1.0), 8)[1:],
                               [64 ** -0.5 * min(e ** -0.5, e * 4 ** -1.5) for e in range(1, 8)], rtol=1e-6)
    assert run(L.PiecewiseDecay([2, 4], [1.0, 0.5, 0.1]), 6) == [1.0, 1.0, 0.5, 0.5, 0.1, 0.1]
    np.testing.assert_allclose(run(L.NaturalExpDecay(0.5, gamma=0.1), 4), [0.5 * math.exp(-0.1 * e) for e in range(4)], rtol=1e-6)
    np.testing.assert_allclose(run(L.InverseTimeDecay(0.5, gamma=0.1), 4), [0.5 / (1 + 0.1 * e) for e in range(4)], rtol=1e-6)
    np.testing.assert_allclose(run(L.ExponentialDecay(0.5, gamma=0.9), 4), [0.5 * 0.9 ** e for e in range(4)], rtol=1e-6)
    np.testing.assert_allclose(run(L.PolynomialDecay(1.0, decay_steps=4, end_lr=0.1, power=2.0), 7),
                               [(1.0 - 0.1) * (1 - min(e, 4) / 4) ** 2 + 0.1 for e in range(7)], rtol=1e-6)
    cyc = run(L.PolynomialDecay(1.0, decay_steps=4, end_lr=0.1, power=1.0, cycle=True), 7)
    assert abs(cyc[5] - ((1.0 - 0.1) * (1 - 5 / 8) + 0.1)) < 1e-6          # decay_steps stretches to the next multiple
    np.testing.assert_allclose(run(L.LinearWarmup(0.5, warmup_steps=4, start_lr=0.0, end_lr=0.5), 6), [0, 0.125, 0.25, 0.375, 0.5, 0.5], rtol=1e-6)
    warm = run(L.LinearWarmup(L.ExponentialDecay(0.5, 0.5), warmup_steps=2, start_lr=0.0, end_lr=0.5), 5)
    np.testing.assert_allclose(warm, [0, 0.25, 0.5, 0.25, 0.125], rtol=1e-6)   # the wrapped scheduler starts after warm-up
    np.testing.assert_allclose(run(L.StepDecay(1.0, step_size=2, gamma=0.1), 5), [1, 1, 0.1, 0.1, 0.01], rtol=1e-6)
    np.testing.assert_allclose(run(L.MultiStepDecay(1.0, milestones=[1, 3], gamma=0.1), 5), [1, 0.1, 0.1, 0.01, 0.01], rtol=1e-6)
    np.testing.assert_allclose(run(L.LambdaDecay(0.5, lambda e: 0.9 ** e), 3), [0.5, 0.45, 0.405], rtol=1e-6)
    np.testing.assert_allclose(run(L.MultiplicativeDecay(0.5, lambda e: 0.9), 3), [0.5, 0.45, 0.405], rtol=1e-6)
    np.testing.assert_allclose(run(L.CosineAnnealingDecay(1.0, T_max=4, eta_min=0.1), 5),
                               [0.1 + 0.9 * (1 + math.cos(math.pi * e / 4)) / 2 for e in range(5)], rtol=1e-5)
    np.testing.assert_allclose(run(L.LinearLR(1.0, total_steps=4, start_factor=0.5, end_factor=1.0), 6),
                               [0.5, 0.625, 0.75, 0.875, 1.0, 1.0], rtol=1e-6)
    wr = run(L.CosineAnnealingWarmRestarts(1.0, T_0=2, T_mult=2, eta_min=0.0), 7)
    np.testing.assert_allclose(wr, [1.0, 0.5, 1.0, (1 + math.cos(math.pi / 4)) / 2, 0.5, (1 + math.cos(3 * math.pi / 4)) / 2, 1.0], rtol=1e-5, atol=1e-7)
    oc = run(L.OneCycleLR(max_learning_rate=1.0, total_steps=10, divide_factor=10.0, end_learning_rate=0.001, phase_pct=0.3), 10)
    assert abs(oc[0] - 0.1) < 1e-6 and abs(max(oc) - 1.0) < 1e-6 and oc.index(max(oc)) == 2 and abs(oc[-1] - 0.001) < 1e-6
    tp = run(L.OneCycleLR(1.0, total_steps=20, divide_factor=10.0, end_learning_rate=0.001, phase_pct=0.25, three_phase=True, anneal_strategy="linear"), 20)
    assert tp.index(max(tp)) == 4 and abs(tp[8] - 0.1) < 1e-6 and abs(tp[-1] - 0.001) < 1e-6
    cy = run(L.CyclicLR(base_learning_rate=0.1, max_learning_rate=1.0, step_size_up=2, step_size_down=2), 9)
    np.testing.assert_allclose(cy, [0.1, 0.55, 1.0, 0.55, 0.1, 0.55, 1.0, 0.55, 0.1], rtol=1e-6)


def test_reduce_on_plateau_and_state():
    s = L.ReduceOnPlateau(1.0, mode="min", factor=0.5, patience=1, threshold=1e-4, cooldown=0, min_lr=0.2)
    lrs = []
    for m in [1.0, 1.0, 1.0, 1.0, 1.0, 1.0]:
        s.step(m)
        lrs.append(s())
    assert lrs == [1.0, 1.0, 0.5, 0.5, 0.25, 0.25] or lrs == [1.0, 0.5, 0.5, 0.25, 0.25, 0.2], lrs
    a = L.StepDecay(1.0, 2, 0.1)
    for _ in range(3):
        a.step()
    b = L.StepDecay(1.0, 2, 0.1)
    b.set_state_dict(a.state_dict())
    assert b() == a() and b.last_epoch == a.last_epoch
    opt = paddle.optimizer.SGD(learning_rate=L.StepDecay(1.0, 1, 0.5), parameters=[paddle.to_tensor([1.0], stop_gradient=False)])
    assert opt.get_lr() == 1.0
    opt._learning_rate.step()
    assert opt.get_lr() == 0.5
