"""PyLayer on torch.autograd.Function (reference: python/paddle/autograd/py_layer.py:29-420,
paddle/fluid/eager/pylayer/).  Paddle spells ``ctx.saved_tensor()`` (method) where torch has the
``saved_tensors`` property; everything else has the same shape, so the context is patched once."""
from __future__ import annotations

import torch
from torch.autograd.function import FunctionCtx

saved_tensors_hooks = torch.autograd.graph.saved_tensors_hooks


def _saved_tensor(self):
    return self.saved_tensors


# torch creates a fresh backward-ctx class per Function; all derive from FunctionCtx.
if not hasattr(FunctionCtx, "saved_tensor"):
    FunctionCtx.saved_tensor = _saved_tensor
if not hasattr(FunctionCtx, "mark_not_inplace"):
    FunctionCtx.mark_not_inplace = lambda self, *a: None
if not hasattr(FunctionCtx, "set_materialize_grads_paddle"):
    FunctionCtx.set_materialize_grads_paddle = FunctionCtx.set_materialize_grads

PyLayerContext = FunctionCtx


class PyLayer(torch.autograd.Function):
    """Subclass with ``@staticmethod forward(ctx, ...)`` / ``backward(ctx, *grads)`` and call ``.apply``."""
