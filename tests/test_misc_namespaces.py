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


def test_geometric_reindex_and_sampling():
    import paddle.geometric as G
    x = torch.tensor([0, 1, 2])
    nb, cnt = torch.tensor([8, 9, 0, 4, 7, 6, 7]), torch.tensor([2, 3, 2], dtype=torch.int32)
    src, dst, nodes = G.reindex_graph(x, nb, cnt)
    assert src.tolist() == [3, 4, 0, 5, 6, 7, 6] and dst.tolist() == [0, 0, 1, 1, 1, 2, 2] and nodes.tolist() == [0, 1, 2, 8, 9, 4, 7, 6]
    assert nodes[src].tolist() == nb.tolist()                               # the numbering is invertible
    nb2, cnt2 = torch.tensor([0, 2, 3, 5, 1]), torch.tensor([1, 3, 1], dtype=torch.int32)
    src, dst, nodes = G.reindex_heter_graph(x, [nb, nb2], [cnt, cnt2])
    assert src.tolist() == [3, 4, 0, 5, 6, 7, 6, 0, 2, 8, 9, 1] and dst.tolist() == [0, 0, 1, 1, 1, 2, 2, 0, 1, 1, 1, 2]
    assert nodes.tolist() == [0, 1, 2, 8, 9, 4, 7, 6, 3, 5]
    with pytest.raises(ValueError):
        G.reindex_graph(torch.tensor([0, 0]), nb[:2], torch.tensor([1, 1]))
    with pytest.raises(ValueError):
        G.reindex_graph(x, nb, torch.tensor([2, 3, 3]))
    # CSC graph: neighbours of v are row[colptr[v]:colptr[v + 1]]
    row = torch.tensor([3, 7, 0, 9, 1, 4, 2, 9, 3, 9, 1, 9, 7])
    colptr = torch.tensor([0, 2, 4, 5, 6, 7, 9, 11, 11, 13, 13])
    full, counts = G.sample_neighbors(row, colptr, torch.tensor([0, 8, 1, 2, 7]))
    assert full.tolist() == [3, 7, 9, 7, 0, 9, 1] and counts.tolist() == [2, 2, 2, 1, 0]
    paddle.seed(3)
    n1, c1, e1 = G.sample_neighbors(row, colptr, torch.tensor([0, 8, 1, 2]), sample_size=1, eids=torch.arange(13) + 100, return_eids=True)
    assert c1.tolist() == [1, 1, 1, 1] and (e1 - 100).tolist() == [int(i) for i in (e1 - 100)]
    for node, got, eid in zip([0, 8, 1, 2], n1.tolist(), (e1 - 100).tolist()):
        lo, hi = int(colptr[node]), int(colptr[node + 1])
        assert lo <= eid < hi and int(row[eid]) == got
    paddle.seed(3)
    n2, _, _ = G.sample_neighbors(row, colptr, torch.tensor([0, 8, 1, 2]), sample_size=1, eids=torch.arange(13) + 100, return_eids=True)
    assert n1.tolist() == n2.tolist()                                       # reproducible under paddle.seed
    draws = torch.cat([G.sample_neighbors(row, colptr, torch.tensor([5]), 1)[0] for _ in range(200)])
    assert set(draws.tolist()) == {9, 3} and 60 < int((draws == 9).sum()) < 140
    with pytest.raises(ValueError):
        G.sample_neighbors(row, colptr, torch.tensor([0]), return_eids=True)
    w = torch.ones(13)
    w[0] = 1e-9                                                             # edge 0 (neighbour 3 of node 0) is almost never taken
    picks = [int(G.weighted_sample_neighbors(row, colptr, w, torch.tensor([0]), 1)[0]) for _ in range(50)]
    assert picks == [7] * 50
    allw, cw = G.weighted_sample_neighbors(row, colptr, w, torch.tensor([0, 5]), sample_size=5)
    assert allw.tolist() == [3, 7, 9, 3] and cw.tolist() == [2, 2]


class _Two(paddle.nn.Layer):
    def __init__(self):
        super().__init__()
        self.fc = paddle.nn.Linear(4, 3)

    def forward(self, x):
        h = self.fc(x)
        return h, torch.relu(h).sum(-1)


def test_inference_predictor_onnx_and_cost_model(tmp_path):
    import paddle.inference as inf
    from paddle.static import InputSpec
    net = _Two()
    net.eval()
    prefix = str(tmp_path / "deploy" / "model")
    paddle.jit.save(net, prefix, input_spec=[InputSpec([None, 4], "float32", "feat")])
    config = inf.Config(prefix + ".pdmodel", prefix + ".pdiparams")
    assert config.prog_file() == prefix + ".pdmodel" and config.params_file() == prefix + ".pdiparams" and not config.use_gpu()
    config.disable_gpu()
    config.set_cpu_math_library_num_threads(2)
    config.switch_ir_optim(True)
    config.enable_memory_optim()
    assert "use_gpu" in config.summary() and config.cpu_math_library_num_threads() == 2
    pred = inf.create_predictor(config)
    assert pred.get_input_names() == ["feat"] and len(pred.get_output_names()) == 2
    x = np.random.RandomState(0).randn(5, 4).astype("float32")
    handle = pred.get_input_handle("feat")
    handle.reshape([5, 4])
    handle.copy_from_cpu(x)
    assert pred.run() is True
    outs = [pred.get_output_handle(n).copy_to_cpu() for n in pred.get_output_names()]
    with torch.no_grad():
        ref = net(torch.from_numpy(x))
    np.testing.assert_allclose(outs[0], ref[0].numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(outs[1], ref[1].numpy(), rtol=1e-5, atol=1e-6)
    out_handle = pred.get_output_handle(pred.get_output_names()[0])
    assert out_handle.shape() == [5, 3] and out_handle.type() == inf.DataType.FLOAT32
    listed = pred.run([torch.from_numpy(x[:2])])                            # dynamic batch, list style
    torch.testing.assert_close(listed[0], ref[0][:2], rtol=1e-5, atol=1e-6)
    assert inf.Config(str(tmp_path / "deploy")).prog_file() == prefix + ".pdmodel"        # model dir form
    assert inf.PredictorPool(config, 2).retrieve(1).get_input_names() == ["feat"]
    clone = pred.clone()
    torch.testing.assert_close(clone.run([torch.from_numpy(x)])[1], ref[1], rtol=1e-5, atol=1e-6)
    with pytest.raises(TypeError):
        handle.copy_from_cpu([1, 2, 3])
    with pytest.raises(NotImplementedError):
        config.enable_tensorrt_engine()
    with pytest.raises(RuntimeError):
        gpu = inf.Config(prefix)
        gpu.enable_use_gpu(100, 0)
        if torch.cuda.is_available():
            raise RuntimeError("a GPU is present: nothing to refuse")
        inf.create_predictor(gpu)
    # mixed precision conversion keeps the black-listed parameters in float32
    mixed = str(tmp_path / "mixed" / "model")
    inf.convert_to_mixed_precision(prefix + ".pdmodel", prefix + ".pdiparams", mixed + ".pdmodel", mixed + ".pdiparams",
                                   inf.PrecisionType.Bfloat16, black_list={"bias"})
    state = paddle.load(mixed + ".pdiparams")
    assert state["fc.weight"].dtype == torch.bfloat16 and state["fc.bias"].dtype == torch.float32
    assert os.path.exists(mixed + ".pdmodel") and os.path.exists(mixed + ".pdmodel.json")
    assert inf.get_num_bytes_of_data_type(inf.DataType.FLOAT16) == 2 and "paddle_b200" in inf.get_version()
    assert inf.get_trt_compile_version() == (0, 0, 0)
    # onnx export needs the onnx package
    import importlib.util
    if importlib.util.find_spec("onnx") is None:
        with pytest.raises(ImportError):
            paddle.onnx.export(net, str(tmp_path / "m"), input_spec=[InputSpec([None, 4], "float32", "feat")])
    else:
        assert os.path.exists(paddle.onnx.export(net, str(tmp_path / "m"), input_spec=[InputSpec([None, 4], "float32", "feat")]))
    cm = paddle.cost_model.CostModel()
    peaks = cm.static_cost_data()
    assert peaks["hbm_gbs"] > 1000 and peaks["tflops_bf16"] > 100
    mm = cm.get_static_op_time("matmul", dtype="bfloat16", shape=(4096, 4096, 4096))
    assert mm["op_time"] == pytest.approx(2 * 4096 ** 3 / (peaks["tflops_bf16"] * 1e12) * 1e3, rel=1e-6)
    assert cm.get_static_op_time("matmul", forward=False, dtype="bfloat16", shape=(4096, 4096, 4096))["op_time"] == pytest.approx(2 * mm["op_time"])
    ew = cm.get_static_op_time("relu", dtype="float32", shape=(1 << 24,))
    assert ew["op_time"] == pytest.approx(2 * (1 << 24) * 4 / (peaks["hbm_gbs"] * 1e9) * 1e3, rel=1e-6)
    assert cm.profile_measure(lambda: torch.ones(64).sum(), warmup=1, iters=3, device="cpu") > 0
    with pytest.raises(ValueError):
        cm.get_static_op_time(None)


def test_quasi_newton_minimisers():
    from paddle.incubate.optimizer.functional import minimize_bfgs, minimize_lbfgs

    def rosenbrock(x):
        return (100 * (x[1:] - x[:-1] ** 2) ** 2 + (1 - x[:-1]) ** 2).sum()
    for fn in (minimize_bfgs, minimize_lbfgs):
        out = fn(rosenbrock, torch.tensor([-1.2, 1.0, 0.5]), max_iters=200, dtype="float64")
        converged, calls, pos, val, grad = out[:5]
        assert bool(converged) and int(calls) < 200 and pos.dtype == torch.float64
        torch.testing.assert_close(pos, torch.ones(3, dtype=torch.float64), rtol=1e-4, atol=1e-4)
        assert float(val) < 1e-8 and float(grad.abs().max()) < 1e-3
    # on a quadratic BFGS recovers the inverse Hessian
    A = torch.tensor([[3.0, 1.0], [1.0, 2.0]], dtype=torch.float64)
    b = torch.tensor([1.0, -1.0], dtype=torch.float64)
    out = minimize_bfgs(lambda x: 0.5 * x @ A @ x - b @ x, torch.zeros(2), dtype="float64")
    torch.testing.assert_close(out[2], torch.linalg.solve(A, b), rtol=1e-6, atol=1e-8)
    torch.testing.assert_close(out[5], torch.linalg.inv(A), rtol=1e-5, atol=1e-6)
    short = minimize_lbfgs(rosenbrock, torch.tensor([-1.2, 1.0]), history_size=3, max_iters=2)
    assert not bool(short[0]) and short[2].dtype == torch.float32
    warm = minimize_bfgs(lambda x: (x ** 2).sum(), torch.ones(2), initial_inverse_hessian_estimate=torch.eye(2) * 0.5)
    assert bool(warm[0]) and float(warm[3]) < 1e-10
    with pytest.raises(ValueError):
        minimize_bfgs(rosenbrock, torch.zeros(2), initial_inverse_hessian_estimate=torch.tensor([[1.0, 2.0], [0.0, 1.0]]))
    with pytest.raises(NotImplementedError):
        minimize_lbfgs(rosenbrock, torch.zeros(2), line_search_fn="armijo")
    with pytest.raises(ValueError):
        minimize_bfgs(rosenbrock, torch.zeros(2), dtype="float16")


def test_cpp_extension_and_hdfs_client(tmp_path, monkeypatch):
    from paddle.utils import cpp_extension as C
    monkeypatch.setenv("PADDLE_EXTENSION_DIR", str(tmp_path / "ext"))
    assert C.get_build_directory() == str(tmp_path / "ext") and os.path.isdir(tmp_path / "ext")
    cuda = C.CUDAExtension(["op.cu", "op.cc"], extra_compile_args={"cxx": ["-O2"]})
    assert "arch=compute_100a,code=sm_100a" in cuda.extra_compile_args["nvcc"] and cuda.extra_compile_args["cxx"] == ["-O2"]
    assert paddle.sysconfig.get_include() in cuda.include_dirs
    named = C.CUDAExtension(["op.cu"], extra_compile_args={"nvcc": ["-arch=sm_90"]})
    assert named.extra_compile_args["nvcc"] == ["-arch=sm_90"]                    # a caller-chosen arch is left alone
    cpp = C.CppExtension(["op.cc"], name="mine")
    assert cpp.name == "mine" and cpp.sources == ["op.cc"]
    with pytest.raises(ValueError):
        C.setup(ext_modules=cpp)
    src = tmp_path / "add_one.cc"
    src.write_text('#include <torch/extension.h>\ntorch::Tensor add_one(torch::Tensor x) { return x + 1; }\n'
                   'PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) { m.def("add_one", &add_one); }\n')
    mod = C.load("pd_test_add_one", [str(src)])
    assert mod.add_one(torch.zeros(3)).tolist() == [1.0, 1.0, 1.0]
    assert os.path.isdir(tmp_path / "ext" / "pd_test_add_one")
    # HDFS client: command construction against a stand-in `hadoop` that records its arguments
    from paddle.distributed.fleet.utils import DistributedInfer, HDFSClient
    from paddle.distributed.fleet.utils.fs import ExecuteError, FSFileNotExistsError
    home = tmp_path / "hadoop"
    os.makedirs(home / "bin")
    log = tmp_path / "calls.log"
    (home / "bin" / "hadoop").write_text(f'#!/bin/sh\necho "$@" >> {log}\ncase "$*" in\n  *"-test -e /missing"*) exit 1;;\n'
                                         '  *"-test -d /data/file"*) exit 1;;\n'
                                         '  *"-ls /data"*) echo "Found 2 items"; echo "drwxr-xr-x   - u g          0 2024-01-01 00:00 /data/sub";'
                                         ' echo "-rw-r--r--   3 u g         12 2024-01-01 00:00 /data/file";;\nesac\nexit 0\n')
    os.chmod(home / "bin" / "hadoop", 0o755)
    client = HDFSClient(str(home), {"fs.default.name": "hdfs://nn:9000", "hadoop.job.ugi": "u,p"}, time_out=2000, sleep_inter=100)
    assert client.is_exist("/data") and not client.is_exist("/missing") and client.is_file("/data/file") and client.is_dir("/data")
    assert client.ls_dir("/data") == (["sub"], ["file"]) and client.list_dirs("/data") == ["sub"] and client.ls_dir("/missing") == ([], [])
    client.mkdirs("/missing")
    client.delete("/data")
    local = tmp_path / "local.txt"
    local.write_text("x")
    client.upload(str(local), "/missing")
    with pytest.raises(FSFileNotExistsError):
        client.upload(str(tmp_path / "nope"), "/missing")
    with pytest.raises(FSFileNotExistsError):
        client.download("/missing", str(tmp_path / "dl"))
    calls = log.read_text()
    assert "-D fs.default.name=hdfs://nn:9000 -D hadoop.job.ugi=u,p" in calls
    for piece in ("-mkdir -p /missing", "-rm -r -f /data", f"-put {local} /missing"):
        assert piece in calls, piece
    with pytest.raises(ExecuteError):
        HDFSClient(str(tmp_path / "no_hadoop")).is_exist("/x")
    with pytest.raises(NotImplementedError):
        DistributedInfer()
