"""paddle.nn.quant and paddle.quantization (reference: python/paddle/nn/quant/quantized_linear.py,
python/paddle/quantization/*.py, test/quantization/test_{qat,ptq,quant_config}.py)."""
import pytest
import torch

import paddle
from paddle.nn.quant import Stub, llm_int8_linear, weight_dequantize, weight_only_linear, weight_quantize
from paddle.quantization import PTQ, QAT, BaseQuanter, QuantConfig, quanter
from paddle.quantization.observers import AbsmaxObserver, GroupWiseWeightObserver
from paddle.quantization.quanters import FakeQuanterWithAbsMaxObserver


def test_weight_only_and_llm_int8():
    torch.manual_seed(0)
    w, x = torch.randn(128, 48), torch.randn(2, 5, 128)
    ref = x @ w
    for algo, dt, bound in (("weight_only_int8", "int8", 127), ("weight_only_int4", "int4", 7)):
        for g in (-1, 64):
            q, s = weight_quantize(w, algo, group_size=g)
            assert q.dtype == torch.int8 and s.dtype == torch.float32
            assert q.shape == ((48, 128) if dt == "int8" else (48, 64)) and s.shape == ((48,) if g == -1 else (2, 48))
            full = s[None] if g == -1 else s.repeat_interleave(64, 0)
            want_scale = w.abs().amax(0) / bound if g == -1 else w.reshape(2, 64, 48).abs().amax(1) / bound
            torch.testing.assert_close(s, want_scale)
            back = weight_dequantize(q, s, algo, "float32", g)
            torch.testing.assert_close(back, torch.round(w / full).clamp(-bound, bound) * full)     # on the grid
            assert float((back - w).abs().max()) <= float(full.max()) / 2 + 1e-6                     # half a step
            y = weight_only_linear(x, q, torch.ones(48), s, dt, group_size=g)
            torch.testing.assert_close(y, x @ back + 1, rtol=1e-4, atol=1e-4)
    q, s = weight_quantize(w.to(torch.bfloat16), "weight_only_int8")
    assert weight_dequantize(q, s, out_dtype="bfloat16").dtype == torch.bfloat16
    assert weight_only_linear(x.to(torch.bfloat16), q, None, s).dtype == torch.bfloat16
    # LLM.int8: an outlier feature is kept in floating point, the rest is exact integer arithmetic
    q, s = weight_quantize(w, "llm.int8")
    xo = x.clone()
    xo[0, 0, 5] = 40.0
    y = llm_int8_linear(xo, q, None, s, threshold=6.0)
    wd = weight_dequantize(q, s, "llm.int8", "float32")
    x2 = xo.reshape(-1, 128)
    keep = torch.ones(128, dtype=torch.bool)
    keep[5] = False
    rs = x2[:, keep].abs().amax(1, keepdim=True) / 127
    want = (torch.round(x2[:, keep] / rs) * rs) @ wd[keep] + x2[:, ~keep] @ wd[~keep]
    torch.testing.assert_close(y.reshape(-1, 48), want, rtol=1e-4, atol=1e-4)
    no_outlier = llm_int8_linear(x.clamp(-5, 5), q, torch.zeros(48), s, threshold=6.0)
    assert float((no_outlier - x.clamp(-5, 5) @ w).abs().max() / ref.abs().max()) < 0.03
    for bad in (dict(algo="int3"), dict(group_size=32)):
        with pytest.raises((ValueError, AssertionError)):
            weight_quantize(w, **bad)
    with pytest.raises(ValueError):
        weight_only_linear(x, q, None, None)


class Net(paddle.nn.Layer):
    def __init__(self):
        super().__init__()
        self.conv = paddle.nn.Conv2D(1, 4, 3, padding=1)
        self.relu = paddle.nn.ReLU()
        self.stub = Stub()
        self.fc = paddle.nn.Linear(4 * 8 * 8, 10)

    def forward(self, x):
        return self.fc(self.stub(self.relu(self.conv(x)).flatten(1)))


def test_qat_quantize_train_convert():
    torch.manual_seed(0)
    q = FakeQuanterWithAbsMaxObserver(moving_rate=0.9)
    assert "moving_rate=0.9" in str(q)
    cfg = QuantConfig(activation=q, weight=q)
    model = Net()
    model.train()
    qm = QAT(cfg).quantize(model)
    assert type(qm.conv).__name__ == "QuantedConv2D" and type(qm.fc).__name__ == "QuantedLinear"
    assert type(qm.relu).__name__ == "ObserveWrapper" and type(qm.stub).__name__ == "QuanterStub"
    assert type(model.conv).__name__ == "Conv2D"                                  # not in place
    assert qm.fc.weight_quanter is not qm.fc.activation_quanter                   # one quanter per use
    x = torch.randn(4, 1, 8, 8)
    opt = paddle.optimizer.SGD(learning_rate=0.05, parameters=qm.parameters())
    losses = []
    for _ in range(5):
        loss = qm(x).square().mean()
        loss.backward()
        opt.step()
        opt.clear_grad()
        losses.append(float(loss.detach()))
    assert losses[-1] < losses[0]                                                 # straight-through gradients train
    aq = qm.fc.activation_quanter
    # moving average after 5 steps: state = sum 0.9^i + 0.9^5, scale = accum / state
    state = 0.9 ** 5 + sum(0.9 ** i for i in range(5))
    assert float(aq._state) == pytest.approx(state, rel=1e-5) and float(aq.scales()) == pytest.approx(float(aq._accum) / state, rel=1e-5)
    assert aq.bit_length() == 8 and aq.quant_axis() == -1 and aq.zero_points() is None
    # fake quant puts values on a 255-level grid inside [-scale, scale]
    out = aq.eval()(torch.linspace(-3, 3, 1001))
    s = float(aq.scales())
    assert float(out.abs().max()) <= s + 1e-6 and len(torch.unique(out)) <= 255
    torch.testing.assert_close(out / s * 127, torch.round(out / s * 127), atol=1e-3, rtol=0)
    qm.eval()
    with torch.no_grad():
        before = qm(x)
        converted = QAT(cfg).convert(qm)
        after = converted(x)
    torch.testing.assert_close(after, before, rtol=1e-4, atol=1e-5)
    assert type(converted.fc.weight_quanter).__name__ == "LinearDequanter" and converted.fc.converted
    w = converted.fc.weight
    torch.testing.assert_close(w, torch.round(w))                                 # weights stored on the integer grid
    assert float(w.abs().max()) <= 127
    kept = QAT(cfg).convert(qm, remain_weight=True)
    assert type(kept.fc.weight_quanter).__name__ == "LinearQuanterDequanter"
    with torch.no_grad():
        torch.testing.assert_close(kept(x), before, rtol=1e-4, atol=1e-5)
    with pytest.raises(AssertionError):
        QAT(cfg).quantize(model.eval())


def test_ptq_and_config_priorities():
    torch.manual_seed(0)
    model = Net()
    model.eval()
    cfg = QuantConfig(activation=None, weight=None)
    cfg.add_type_config(paddle.nn.Linear, activation=AbsmaxObserver(quant_bits=8), weight=AbsmaxObserver(quant_bits=8))
    pm = PTQ(cfg).quantize(model)
    assert type(pm.fc).__name__ == "QuantedLinear" and type(pm.conv).__name__ == "Conv2D" and type(pm.relu).__name__ == "ReLU"
    x = torch.randn(16, 1, 8, 8)
    with torch.no_grad():
        ref = model(x)
        for i in range(4):
            torch.testing.assert_close(pm(x[i * 4:(i + 1) * 4]), ref[i * 4:(i + 1) * 4])          # observers do not change values
        feats = model.relu(model.conv(x)).flatten(1)
    assert float(pm.fc.activation_quanter.scales()) == pytest.approx(float(feats.abs().max()), rel=1e-6)
    assert float(pm.fc.weight_quanter.scales()) == pytest.approx(float(model.fc.weight.abs().max()), rel=1e-6)
    cm = PTQ(cfg).convert(pm)
    with torch.no_grad():
        out = cm(x)
    assert float((out - ref).abs().max() / ref.abs().max()) < 0.05 and not torch.equal(out, ref)
    with pytest.raises(AssertionError):
        PTQ(cfg).quantize(model.train())
    # layer > name > type > global
    model = Net()
    model.train()
    g, t, n, l = (FakeQuanterWithAbsMaxObserver(moving_rate=r) for r in (0.1, 0.2, 0.3, 0.4))
    cfg = QuantConfig(activation=g, weight=g)
    cfg.add_type_config(paddle.nn.Linear, activation=t, weight=t)
    cfg.add_name_config("fc", activation=n, weight=n)
    qm = QAT(cfg).quantize(model)
    assert qm.conv.weight_quanter._moving_rate == 0.1 and qm.fc.weight_quanter._moving_rate == 0.3
    cfg2 = QuantConfig(activation=g, weight=g)
    cfg2.add_type_config(paddle.nn.Linear, activation=t, weight=t)
    assert QAT(cfg2).quantize(model).fc.weight_quanter._moving_rate == 0.2
    cfg3 = QuantConfig(activation=None, weight=None)
    cfg3.add_layer_config([model.fc], activation=l, weight=None)
    q3 = QAT(cfg3).quantize(model, inplace=False)
    # deep copies lose object identity, so layer configs are for in-place use
    q3 = QAT(cfg3)
    q3._config.add_layer_config([model.fc], activation=l, weight=None)
    m3 = q3.quantize(model, inplace=True)
    assert m3.fc.activation_quanter._moving_rate == 0.4 and m3.fc.weight_quanter is None and type(m3.conv).__name__ == "Conv2D"
    assert "QuantedLinear" in str(q3) or "fc" in str(q3)
    # group-wise weight observer and a user-defined quanter through the decorator
    obs = GroupWiseWeightObserver(quant_bits=4, group_size=64)._instance(None)
    w = torch.randn(128, 6)
    obs(w)
    torch.testing.assert_close(obs.scales(), w.reshape(2, 64, 6).abs().amax(1))

    @quanter("HalfQuanter")
    class HalfQuanterLayer(BaseQuanter):
        def __init__(self, layer, step=0.5):
            super().__init__()
            self.step = step

        def forward(self, input):
            return torch.round(input / self.step) * self.step

        def scales(self):
            return torch.tensor([self.step * 127])

        def zero_points(self):
            return None

        def quant_axis(self):
            return -1

        def bit_length(self):
            return 8
    factory = globals().get("HalfQuanter") or getattr(__import__(__name__), "HalfQuanter")
    inst = factory(step=0.25)._instance(None)
    assert isinstance(inst, HalfQuanterLayer) and inst(torch.tensor([0.3])).item() == 0.25
    custom = QuantConfig(activation=None, weight=factory(step=0.25))
    mm = Net()
    mm.train()
    cq = QAT(custom).quantize(mm)
    assert isinstance(cq.fc.weight_quanter, HalfQuanterLayer) and cq.fc.activation_quanter is None
