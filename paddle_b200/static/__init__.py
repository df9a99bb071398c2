"""paddle.static: only the pieces dygraph code touches.  The static-graph executor, Program/IR and their passes are
out of product scope (the reference's 260 k-line graph stack is replaced by eager execution + CUDA
streams / graphs); ``InputSpec`` is kept because ``paddle.jit.to_static`` / ``paddle.Model`` signatures take it."""
from __future__ import annotations

import torch

from ..framework import dtype as _dt


class InputSpec:
    def __init__(self, shape, dtype="float32", name=None, stop_gradient=False):
        self.shape = tuple(shape)
        self.dtype = _dt.convert_dtype(dtype) if hasattr(_dt, "convert_dtype") else dtype
        self.name = name
        self.stop_gradient = stop_gradient

    @classmethod
    def from_tensor(cls, tensor, name=None):
        return cls(tuple(tensor.shape), tensor.dtype, name or getattr(tensor, "name", None))

    @classmethod
    def from_numpy(cls, ndarray, name=None):
        return cls(tuple(ndarray.shape), torch.from_numpy(ndarray).dtype, name)

    def batch(self, batch_size):
        self.shape = (batch_size,) + self.shape
        return self

    def unbatch(self):
        self.shape = self.shape[1:]
        return self

    def __repr__(self):
        return f"InputSpec(shape={self.shape}, dtype={self.dtype}, name={self.name}, stop_gradient={self.stop_gradient})"


def _no_static(*a, **k):
    raise NotImplementedError("the static-graph executor is out of scope of this framework (dygraph only); "
                              "see paddle_b200/static/__init__.py")


Program = Executor = program_guard = default_main_program = default_startup_program = data = _no_static
