"""paddle.device.cuda.graphs (reference: python/paddle/device/cuda/graphs.py:37-110 CUDAGraph, wrap_cuda_graph; native
side paddle/phi/backends/gpu/cuda/cuda_graph.cc with its private memory pool).

CUDA graphs are the replacement for a tracing compiler in this framework: a launch-bound region (decode step, a
micro-batch of small layers) is captured once on a side stream with torch's graph-private allocator pool and replayed
with one launch.  The Paddle API shape is kept: ``g = CUDAGraph(); g.capture_begin(); ...; g.capture_end(); g.replay()``.
Kernels of this library are capture-safe as long as their tensor maps / descriptor tables were built for the same
buffers (pointers are baked into the captured launches), which is what static input / output tensors guarantee."""
from __future__ import annotations

import torch


class CUDAGraph:
    def __init__(self, place=None, mode="thread_local", pool_id=None):
        self._graph = torch.cuda.CUDAGraph()
        self._pool = pool_id
        self._mode = {"thread_local": "thread_local", "global": "global", "relaxed": "relaxed"}.get(mode, "thread_local")
        self._stream = None
        self._ctx = None

    def capture_begin(self):
        self._stream = torch.cuda.Stream()
        self._stream.wait_stream(torch.cuda.current_stream())
        self._ctx = torch.cuda.graph(self._graph, pool=self._pool, stream=self._stream, capture_error_mode=self._mode)
        self._ctx.__enter__()

    def capture_end(self):
        self._ctx.__exit__(None, None, None)
        self._ctx = None
        torch.cuda.current_stream().wait_stream(self._stream)

    def replay(self):
        self._graph.replay()

    def reset(self):
        self._graph.reset()

    def pool(self):
        return self._graph.pool()

    def print_to_dot_files(self, dirname, flags=None):
        self._graph.enable_debug_mode()
        self._graph.debug_dump(dirname)


def is_cuda_graph_supported():
    return torch.cuda.is_available()


def wrap_cuda_graph(function, mode="thread_local", memory_pool="default"):
    """Capture ``function`` on its first call with static copies of the inputs, replay afterwards (inference-style:
    no autograd through the replay)."""
    state = {}

    def run(*args):
        if not torch.cuda.is_available():
            return function(*args)
        if "g" not in state:
            state["inputs"] = [a.clone() if isinstance(a, torch.Tensor) else a for a in args]
            for _ in range(2):                      # warm up outside capture (lazy inits, autotuning, tensor maps)
                function(*state["inputs"])
            torch.cuda.synchronize()
            g = CUDAGraph(mode=mode)
            g.capture_begin()
            state["out"] = function(*state["inputs"])
            g.capture_end()
            state["g"] = g
        for dst, src in zip(state["inputs"], args):
            if isinstance(dst, torch.Tensor):
                dst.copy_(src)
        state["g"].replay()
        return state["out"]
    return run
