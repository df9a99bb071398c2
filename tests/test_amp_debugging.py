"""paddle.amp.debugging: check_numerics statistics, tensor checker modes / step window / log files, operator stats,
accuracy comparison of two logged runs (reference: python/paddle/amp/debugging.py, accuracy_compare.py)."""
import os

import pytest
import torch

import paddle_b200 as paddle
from paddle_b200.amp import debugging as D


def test_check_numerics_counts_and_abort():
    x = torch.tensor([1.0, 0.0, float("nan"), float("inf"), -3.0, 0.0])
    with pytest.raises(RuntimeError, match="NAN or INF"):
        D.check_numerics(x, "relu", "x")
    stats, values = D.check_numerics(x, "relu", "x", D.DebugMode.CHECK_NAN_INF)
    assert stats.tolist() == [1, 1, 2]
    assert values.tolist() == pytest.approx([1.0, -3.0, -0.5])
    stats, _ = D.check_numerics(torch.ones(8), "ones", "y")
    assert stats.tolist() == [0, 0, 0]


def test_tensor_checker_window_logs_and_compare(tmp_path):
    paddle.seed(0)
    net = paddle.nn.Sequential(paddle.nn.Linear(8, 16), paddle.nn.ReLU(), paddle.nn.Linear(16, 4))
    x = torch.randn(5, 8)
    dirs = {}
    for tag, scale in (("fp32", 1.0), ("lowp", 1.0)):
        cfg = D.TensorCheckerConfig(True, D.DebugMode.CHECK_ALL, output_dir=str(tmp_path / tag), debug_step=(1, 3))
        D.TensorCheckerConfig.current_step_id = 0
        for step in range(4):
            D.enable_tensor_checker(cfg)
            y = net(x if tag == "fp32" else x * (1.0 + 0.5 * (step == 1)))
            D.disable_tensor_checker()
        dirs[tag] = str(tmp_path / tag)
        log = open(os.path.join(dirs[tag], "worker_0.log")).read()
        assert "[PRECISION] [INFO]" in log and "op=Linear" in log
        assert log.count("op=Sequential") == 2          # only steps 1 and 2 are inside debug_step=(1, 3)
    rows = D.compare_accuracy(dirs["fp32"], dirs["lowp"], str(tmp_path / "cmp.csv"), dump_all_tensors=True)
    assert os.path.exists(tmp_path / "cmp.csv")
    assert rows and {r["status"] for r in rows} >= {"ok", "diff"}     # the perturbed step shows up as a difference


def test_abort_mode_through_layer_hook_and_decorator():
    class Bad(paddle.nn.Layer):
        @D.check_layer_numerics
        def forward(self, x):
            return x / 0.0
    with pytest.raises(RuntimeError):
        Bad()(torch.ones(3))
    cfg = D.TensorCheckerConfig(True, D.DebugMode.CHECK_NAN_INF_AND_ABORT)
    D.enable_tensor_checker(cfg)
    try:
        with pytest.raises(RuntimeError):
            paddle.nn.Linear(2, 2)(torch.tensor([[float("nan"), 1.0]]))
    finally:
        D.disable_tensor_checker()
    assert not paddle.get_flags(["FLAGS_check_nan_inf"])["FLAGS_check_nan_inf"]


def test_operator_stats_collection(capsys):
    with D.collect_operator_stats():
        a = torch.randn(4, 4)
        (a.bfloat16() @ a.bfloat16()).float().sum()
    out = capsys.readouterr().out
    assert "BF16 Calls" in out and "mm" in out


@pytest.mark.gpu
def test_check_numerics_native_kernel_matches_torch():
    for dtype in (torch.float32, torch.bfloat16, torch.float16):
        x = torch.randn(1 << 16, device="cuda").to(dtype)
        x[5], x[77], x[100:110] = float("nan"), float("-inf"), 0.0
        stats, values = D.check_numerics(x, "op", "t", D.DebugMode.CHECK_NAN_INF)
        f = x.float()
        fin = torch.isfinite(f)
        assert stats.tolist() == [1, 1, int(((f == 0) & fin).sum())]
        torch.testing.assert_close(values.cpu(), torch.stack([f[fin].max(), f[fin].min(), f[fin].mean()]).cpu(), rtol=1e-3, atol=1e-4)
