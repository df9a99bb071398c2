"""Small namespaces: paddle.hub (local hubconf), paddle.sysconfig, paddle.batch, the Hermitian n-d FFTs, hapi callbacks
(ReduceLROnPlateau, VisualDL, EarlyStopping on evaluation results)."""
import json
import os

import numpy as np
import pytest
import scipy.fft as sf
import torch

import paddle


def test_hub_sysconfig_batch(tmp_path):
    (tmp_path / "hubconf.py").write_text('dependencies = ["torch"]\nimport paddle\n\ndef tiny(n=3):\n    """a tiny linear"""\n'
                                         '    return paddle.nn.Linear(n, n)\n\ndef _hidden():\n    pass\n')
    d = str(tmp_path)
    assert paddle.hub.list(d, source="local") == ["tiny"]
    assert paddle.hub.help(d, "tiny", source="local") == "a tiny linear"
    assert paddle.hub.load(d, "tiny", source="local", n=4).weight.shape == (4, 4)
    with pytest.raises(RuntimeError):
        paddle.hub.load(d, "missing", source="local")
    with pytest.raises(ValueError):
        paddle.hub.list(d, source="svn")
    with pytest.raises(RuntimeError, match="no network"):
        paddle.hub.load("owner/repo:dev", "tiny")
    (tmp_path / "needs").mkdir()
    (tmp_path / "needs" / "hubconf.py").write_text('dependencies = ["a_package_that_does_not_exist"]\ndef f():\n    return 1\n')
    with pytest.raises(RuntimeError, match="Missing dependencies"):
        paddle.hub.load(str(tmp_path / "needs"), "f", source="local")
    assert os.path.exists(os.path.join(paddle.sysconfig.get_include(), "pdb200_kernels.h"))
    assert os.path.isdir(paddle.sysconfig.get_lib())
    reader = paddle.batch(lambda: iter(range(5)), 2)
    assert list(reader()) == [[0, 1], [2, 3], [4]] and list(paddle.batch(lambda: iter(range(5)), 2, drop_last=True)()) == [[0, 1], [2, 3]]
    with pytest.raises(ValueError):
        paddle.batch(lambda: iter(()), 0)


def test_hermitian_nd_ffts_match_scipy():
    x = torch.randn(4, 6, 5, dtype=torch.float64)
    for norm in ("backward", "forward", "ortho"):
        spec = sf.ihfft2(x.numpy(), norm=norm)
        np.testing.assert_allclose(paddle.fft.ihfft2(x, norm=norm).numpy(), spec, atol=1e-12)
        np.testing.assert_allclose(paddle.fft.hfft2(torch.from_numpy(spec), s=(6, 5), norm=norm).numpy(), x.numpy(), atol=1e-12)
        specn = sf.ihfftn(x.numpy(), norm=norm)
        np.testing.assert_allclose(paddle.fft.ihfftn(x, norm=norm).numpy(), specn, atol=1e-12)
        np.testing.assert_allclose(paddle.fft.hfftn(torch.from_numpy(specn), s=(4, 6, 5), norm=norm).numpy(), x.numpy(), atol=1e-12)
    np.testing.assert_allclose(paddle.fft.hfftn(torch.from_numpy(sf.ihfftn(x.numpy(), axes=(0, 2))), s=(4, 5), axes=(0, 2)).numpy(), x.numpy(), atol=1e-12)


class _Data(paddle.io.Dataset):
    def __init__(self, n=32):
        g = torch.Generator().manual_seed(0)
        self.x = torch.randn(n, 4, generator=g)
        self.y = (self.x.sum(1, keepdim=True) > 0).long()

    def __len__(self):
        return len(self.x)

    def __getitem__(self, i):
        return self.x[i].numpy(), self.y[i].numpy()


def test_callbacks_follow_evaluation(tmp_path):
    net = paddle.nn.Linear(4, 2)
    model = paddle.Model(net)
    opt = paddle.optimizer.SGD(learning_rate=0.5, parameters=net.parameters())
    model.prepare(opt, paddle.nn.CrossEntropyLoss(), paddle.metric.Accuracy())
    seen = []

    class Spy(paddle.callbacks.Callback):
        def on_train_begin(self, logs=None):
            seen.append(("params", self.params["epochs"], self.params["steps"], self.params["metrics"]))

        def on_eval_end(self, logs=None):
            seen.append(("eval", sorted(logs)))

    opt_lr = []
    # min_delta so large that no evaluation counts as an improvement: every one after the first is a plateau
    reduce = paddle.callbacks.ReduceLROnPlateau(monitor="loss", factor=0.5, patience=1, verbose=0, min_lr=0.1, min_delta=1e9)

    class Track(paddle.callbacks.Callback):
        def on_eval_end(self, logs=None):
            opt_lr.append(opt.get_lr())
    log_dir = str(tmp_path / "vdl")
    model.fit(_Data(), _Data(16), batch_size=8, epochs=5, verbose=0,
              callbacks=[Spy(), reduce, Track(), paddle.callbacks.VisualDL(log_dir)])
    assert seen[0] == ("params", 5, 4, ["loss", "acc"]) and seen[1] == ("eval", ["acc", "loss"])
    # first evaluation sets the best value; every later one is a plateau: 0.5 -> 0.25 -> 0.125 -> 0.1 (min_lr)
    assert opt_lr[0] == 0.5 and opt_lr[-1] == pytest.approx(0.1) and sorted(opt_lr, reverse=True) == opt_lr
    records = [json.loads(l) for l in open(os.path.join(log_dir, "scalars.jsonl"))]
    tags = {r["tag"] for r in records}
    assert {"train/loss", "eval/loss", "eval/acc"} <= tags
    assert sum(r["tag"] == "train/loss" for r in records) == 20 and sum(r["tag"] == "eval/acc" for r in records) == 5
    # EarlyStopping watches the evaluation result
    net2 = paddle.nn.Linear(4, 2)
    m2 = paddle.Model(net2)
    m2.prepare(paddle.optimizer.SGD(learning_rate=0.0, parameters=net2.parameters()), paddle.nn.CrossEntropyLoss(), paddle.metric.Accuracy())
    evals = []

    class Count(paddle.callbacks.Callback):
        def on_eval_end(self, logs=None):
            evals.append(1)
    m2.fit(_Data(), _Data(16), batch_size=8, epochs=10, verbose=0,
           callbacks=[paddle.callbacks.EarlyStopping("acc", mode="max", patience=1, verbose=0), Count()])
    assert len(evals) == 3                      # best at evaluation 1, two evaluations without improvement > patience 1
    with pytest.raises(ValueError):
        paddle.callbacks.ReduceLROnPlateau(factor=1.5)
    import importlib.util
    if importlib.util.find_spec("wandb") is None:
        with pytest.raises(ImportError):
            paddle.callbacks.WandbCallback(project="x")
    else:
        assert paddle.callbacks.WandbCallback(project="x", mode="disabled").wandb_args["project"] == "x"
