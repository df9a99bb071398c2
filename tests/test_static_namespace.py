"""paddle.static: the names with a meaning in eager mode (reference: python/paddle/static/__init__.py)."""
import os

import numpy as np
import pytest
import torch

import paddle
import paddle.static as st


def test_places_guards_and_autograd():
    assert len(st.cpu_places(3)) == 3 and str(st.cpu_places()[0]).lower().find("cpu") >= 0
    assert [p.get_device_id() if hasattr(p, "get_device_id") else 0 for p in st.cuda_places([0, 1])] in ([0, 1], [0, 0])
    with pytest.raises(NotImplementedError):
        st.xpu_places()
    with st.name_scope("block"):
        with st.device_guard("cpu"):
            t = torch.ones(2)
    assert t.device.type == "cpu"
    with pytest.raises(ValueError):
        with st.device_guard("tpu"):
            pass
    g = st.create_global_var([2, 3], 1.5, "float32", persistable=True, name="gv")
    assert g.shape == (2, 3) and float(g[0, 0]) == 1.5 and g.name == "gv"
    w = st.create_parameter([3, 2], "float32")
    assert w.requires_grad and w.shape == (3, 2)
    x = torch.tensor([1.0, 2.0, 3.0], requires_grad=True)
    y = (x ** 2).sum()
    (gx,) = st.gradients(y, x)
    torch.testing.assert_close(gx, 2 * x.detach())
    (gx2,) = st.gradients([x * 3], [x], [torch.tensor([1.0, 0.0, 2.0])])
    torch.testing.assert_close(gx2, torch.tensor([3.0, 0.0, 6.0]))
    lin = paddle.nn.Linear(3, 1)
    pairs = st.append_backward(lin(x.detach()).sum(), parameter_list=lin.parameters())
    assert len(pairs) == 2 and pairs[0][0] is lin.weight
    torch.testing.assert_close(pairs[0][1], x.detach().reshape(3, 1))
    with pytest.raises(ValueError):
        st.append_backward(y)
    for name in ("Program", "Executor", "data", "program_guard"):
        with pytest.raises(NotImplementedError):
            getattr(st, name)()
    with pytest.raises(NotImplementedError):
        st.IpuStrategy()


def test_ema_print_pyfunc_metrics(capsys):
    p = paddle.create_parameter([2], "float32")
    with torch.no_grad():
        p.copy_(torch.tensor([1.0, 2.0]))
    ema = st.ExponentialMovingAverage(0.5, parameters=[p])
    ema.update()                                           # ema = 0.5 * [1, 2]
    with torch.no_grad():
        p.copy_(torch.tensor([3.0, 4.0]))
    ema.update()                                           # ema = 0.25 * [1, 2] + 0.5 * [3, 4]; correction 1 - 0.25
    with ema.apply():
        torch.testing.assert_close(p.detach(), (0.25 * torch.tensor([1.0, 2.0]) + 0.5 * torch.tensor([3.0, 4.0])) / 0.75)
    torch.testing.assert_close(p.detach(), torch.tensor([3.0, 4.0]))       # restored
    with ema.apply(need_restore=False):
        pass
    assert float(p[0]) != 3.0
    ema.restore()
    torch.testing.assert_close(p.detach(), torch.tensor([3.0, 4.0]))
    warm = st.ExponentialMovingAverage(0.999, thres_steps=lambda: 0, parameters=[p])
    assert warm._current_decay() == pytest.approx(0.1)
    with pytest.raises(ValueError):
        st.ExponentialMovingAverage()
    t = torch.arange(5.0)
    for _ in range(3):
        out = st.Print(t, first_n=2, message="probe", summarize=3)
    assert out is t
    printed = capsys.readouterr().out
    assert printed.count("probe") == 2 and "data: [0.0, 1.0, 2.0]" in printed and "shape: [5]" in printed
    # py_func: forward through numpy-style code, custom backward
    x = torch.tensor([0.5, -1.0], requires_grad=True)
    y = st.py_func(lambda a: torch.tanh(a.detach()), x, backward_func=lambda a, out, dout: dout * (1 - out * out))
    y.sum().backward()
    torch.testing.assert_close(x.grad, 1 - torch.tanh(x.detach()) ** 2)
    holder = torch.zeros(2)
    st.py_func(lambda a: a.detach() * 2, x, out=holder)
    torch.testing.assert_close(holder, x.detach() * 2)
    pred = torch.tensor([[0.1, 0.9], [0.6, 0.4], [0.3, 0.7], [0.8, 0.2]])
    lab = torch.tensor([[1], [0], [1], [0]])
    assert float(st.accuracy(pred, lab)) == 1.0
    a, b, stats = st.auc(pred, lab)
    assert float(a) == pytest.approx(1.0) and float(b) == float(a) and len(stats) == 2
    sq, ab, ps, q, pos, n = st.ctr_metric_bundle(pred[:, 1], lab)
    assert float(n) == 4 and float(pos) == 2 and float(ab) == pytest.approx(0.1 + 0.4 + 0.3 + 0.2)


def test_save_load_inference_and_serialisation(tmp_path):
    net = paddle.nn.Sequential(paddle.nn.Linear(4, 3), paddle.nn.ReLU())
    opt = paddle.optimizer.Adam(parameters=net.parameters())
    prefix = str(tmp_path / "m")
    st.save(net, prefix, optimizer=opt)
    assert os.path.exists(prefix + ".pdparams") and os.path.exists(prefix + ".pdopt")
    other = paddle.nn.Sequential(paddle.nn.Linear(4, 3), paddle.nn.ReLU())
    st.load(other, prefix)
    torch.testing.assert_close(other[0].weight, net[0].weight)
    state = st.load_program_state(prefix)
    assert isinstance(state["0.weight"], np.ndarray) and state["0.weight"].shape == (4, 3)
    state["0.bias"] = np.full(3, 0.5, dtype="float32")
    state["ghost"] = np.zeros(1, dtype="float32")
    with pytest.warns(UserWarning):
        st.set_program_state(other, state)
    assert other[0].bias.tolist() == [0.5, 0.5, 0.5]
    with pytest.raises(TypeError):
        st.save("program", prefix)
    with pytest.raises(ValueError):
        st.load(other, str(tmp_path / "nothing"))
    spec = st.InputSpec([None, 4], "float32", "x")
    st.save_inference_model(str(tmp_path / "infer" / "model"), [spec], None, None, program=net.eval())
    prog, feeds, fetches = st.load_inference_model(str(tmp_path / "infer" / "model"))
    x = torch.randn(5, 4)
    assert feeds == ["x"] and fetches == ["fetch_0"]
    with torch.no_grad():
        torch.testing.assert_close(prog(x), net(x))
    blob, weights = st.serialize_program([spec], None, program=net), st.serialize_persistables([spec], None, program=net)
    assert isinstance(blob, bytes) and isinstance(weights, bytes)
    st.save_to_file(str(tmp_path / "p.bin"), blob)
    restored = st.deserialize_program(st.load_from_file(str(tmp_path / "p.bin")))
    with torch.no_grad():
        net[0].weight.mul_(2.0)                                  # diverge, then bring the serialised weights back
    st.deserialize_persistables(restored, weights)
    fresh = paddle.nn.Sequential(paddle.nn.Linear(4, 3), paddle.nn.ReLU())
    st.deserialize_persistables(fresh, weights)
    with torch.no_grad():
        torch.testing.assert_close(restored(x), fresh(x))
    with pytest.raises(ValueError):
        st.save_to_file(str(tmp_path / "q.bin"), "text")
    assert st.normalize_program(net, [spec], None) is net
    # scopes and strategy holders
    scope = st.Scope()
    with st.scope_guard(scope):
        st.global_scope().var("w").set(np.ones(2, dtype="float32"))
        assert st.global_scope() is scope
    assert scope.find_var("w").get_tensor().tolist() == [1.0, 1.0] and st.global_scope() is not scope
    assert scope.find_var("nothing") is None
    bs = st.BuildStrategy()
    bs.fuse_bn_act_ops = True
    assert st.CompiledProgram(net, bs)._build_strategy.fuse_bn_act_ops
    wn = st.WeightNormParamAttr(dim=0, name="wn", learning_rate=0.5)
    assert wn.dim == 0 and wn.name == "wn" and wn.learning_rate == 0.5
