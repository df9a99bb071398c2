"""paddle.cost_model (reference: python/paddle/cost_model/cost_model.py: per-op GPU timings measured once and shipped
as static_op_benchmark.json, plus a profiler-driven measurement of a static program).

Here ``profile_measure`` times a callable on the current device with CUDA events (wall clock on CPU) and
``get_static_op_time`` answers from the analytic model of the auto-tuner (roofline over the measured B200 peaks in
MEASURED_PEAKS.json) instead of a table measured on other hardware."""
from __future__ import annotations

import time

import torch

__all__ = ["CostModel"]


class CostModel:
    def __init__(self):
        self._static_cost_data = None

    def profile_measure(self, fn, *args, warmup=3, iters=10, device=None, **kwargs):
        """Mean milliseconds per call of ``fn(*args, **kwargs)``."""
        use_cuda = torch.cuda.is_available() and (device is None or str(device).startswith(("gpu", "cuda")))
        for _ in range(warmup):
            fn(*args, **kwargs)
        if use_cuda:
            start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            start.record()
            for _ in range(iters):
                fn(*args, **kwargs)
            end.record()
            torch.cuda.synchronize()
            return start.elapsed_time(end) / iters
        t0 = time.perf_counter()
        for _ in range(iters):
            fn(*args, **kwargs)
        return (time.perf_counter() - t0) * 1e3 / iters

    def static_cost_data(self):
        """The peaks the analytic estimates are made from."""
        from ..distributed.auto_tuner.cost_model import hardware_peaks
        self._static_cost_data = hardware_peaks()
        return self._static_cost_data

    def get_static_op_time(self, op_name, forward=True, dtype="float32", shape=None):
        """Roofline estimate in milliseconds: ``matmul`` with shape (m, n, k) is bounded by tensor-core throughput,
        every other op by memory traffic over ``shape`` elements (read + write; backward = twice the forward)."""
        if op_name is None:
            raise ValueError("op_name should not be empty when you want to get static op time")
        peaks = self._static_cost_data or self.static_cost_data()
        size = {"float32": 4, "float16": 2, "bfloat16": 2, "float64": 8}.get(dtype, 4)
        if op_name in ("matmul", "matmul_v2", "linear", "mul"):
            m, n, k = shape or (4096, 4096, 4096)
            flops = 2.0 * m * n * k * (1 if forward else 2)
            t = max(flops / (peaks["tflops_bf16"] * 1e12 * (1.0 if size <= 2 else 0.125)), (m * k + k * n + m * n) * size / (peaks["hbm_gbs"] * 1e9))
        else:
            numel = 1
            for s in (shape or (1 << 20,)):
                numel *= s
            t = 2.0 * numel * size * (1 if forward else 2) / (peaks["hbm_gbs"] * 1e9)
        return {"op_time": t * 1e3, "config": f"{op_name} {dtype} {tuple(shape) if shape else ''}".strip()}
