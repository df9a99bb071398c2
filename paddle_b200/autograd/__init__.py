"""paddle.autograd (reference: python/paddle/autograd/{py_layer,backward_mode,autograd}.py; C++ engine
paddle/fluid/eager/backward.cc).  torch.autograd is the engine; PyLayer maps onto torch.autograd.Function."""
from __future__ import annotations

import torch

from .py_layer import PyLayer, PyLayerContext, saved_tensors_hooks  # noqa: F401

no_grad = torch.no_grad
enable_grad = torch.enable_grad
set_grad_enabled = torch.set_grad_enabled
is_grad_enabled = torch.is_grad_enabled


def backward(tensors, grad_tensors=None, retain_graph=False):
    if isinstance(tensors, torch.Tensor):
        tensors = [tensors]
    if grad_tensors is not None and isinstance(grad_tensors, torch.Tensor):
        grad_tensors = [grad_tensors]
    torch.autograd.backward(tensors, grad_tensors, retain_graph=retain_graph)


def grad(outputs, inputs, grad_outputs=None, retain_graph=None, create_graph=False, only_inputs=True,
         allow_unused=False, no_grad_vars=None):
    single = isinstance(inputs, torch.Tensor)
    res = torch.autograd.grad(outputs, [inputs] if single else list(inputs), grad_outputs,
                              retain_graph=retain_graph, create_graph=create_graph, allow_unused=allow_unused)
    return list(res)


class Jacobian:
    """Jacobian of already-computed ``ys`` with respect to ``xs`` (reference: python/paddle/autograd/autograd.py
    Jacobian :34-160): shape [ny, nx], or [B, ny, nx] with ``batch_axis=0`` (rows of a batch are independent).  Rows are
    evaluated through the recorded graph when first indexed; ``J[:]`` gives the whole matrix."""

    def __init__(self, ys, xs, batch_axis=None):
        assert batch_axis in (None, 0), "batch_axis is None or 0"
        self._ys, self._xs, self._batch = ys, xs, batch_axis
        self._full = None
        if batch_axis is None:
            self.shape = (ys.numel(), xs.numel())
        else:
            self.shape = (ys.shape[0], ys[0].numel(), xs[0].numel())

    def _evaluate(self):
        if self._full is not None:
            return self._full
        ys, xs = self._ys, self._xs
        rows = []
        if self._batch is None:
            flat = ys.reshape(-1)
            for i in range(flat.numel()):
                g, = torch.autograd.grad(flat[i], xs, retain_graph=True, create_graph=torch.is_grad_enabled(), allow_unused=True)
                rows.append((torch.zeros_like(xs) if g is None else g).reshape(-1))
            self._full = torch.stack(rows) if rows else xs.new_zeros(self.shape)
        else:
            B = ys.shape[0]
            flat = ys.reshape(B, -1)
            for j in range(flat.shape[1]):
                g, = torch.autograd.grad(flat[:, j].sum(), xs, retain_graph=True, create_graph=torch.is_grad_enabled(), allow_unused=True)
                rows.append((torch.zeros_like(xs) if g is None else g).reshape(B, -1))
            self._full = torch.stack(rows, 1) if rows else xs.new_zeros(self.shape)
        return self._full

    def __getitem__(self, idx):
        return self._evaluate()[idx]

    def numpy(self):
        return self._evaluate().detach().cpu().numpy()


class Hessian(Jacobian):
    """Hessian of a scalar (per batch row, with ``batch_axis=0``) ``ys``: the Jacobian of its gradient."""

    def __init__(self, ys, xs, batch_axis=None):
        if batch_axis is None:
            assert ys.numel() == 1, "hessian needs a scalar output"
            g, = torch.autograd.grad(ys.reshape(()), xs, create_graph=True)
        else:
            g, = torch.autograd.grad(ys.sum(), xs, create_graph=True)
        super().__init__(g, xs, batch_axis)


def jacobian(ys, xs, batch_axis=None):
    """``paddle.autograd.jacobian``: a tuple of xs (or ys) gives a tuple (of tuples) of Jacobians."""
    if isinstance(ys, (list, tuple)):
        return tuple(jacobian(y, xs, batch_axis) for y in ys)
    if isinstance(xs, (list, tuple)):
        return tuple(Jacobian(ys, x, batch_axis) for x in xs)
    return Jacobian(ys, xs, batch_axis)


def hessian(ys, xs, batch_axis=None):
    if isinstance(xs, (list, tuple)):
        grads = torch.autograd.grad(ys.sum(), list(xs), create_graph=True)
        return tuple(tuple(Jacobian(g, x, batch_axis) for x in xs) for g in grads)
    return Hessian(ys, xs, batch_axis)
