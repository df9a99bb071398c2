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


def jacobian(ys, xs, batch_axis=None):
    return torch.autograd.functional.jacobian(lambda *a: ys, xs)


def hessian(ys, xs, batch_axis=None):
    return torch.autograd.functional.hessian(lambda *a: ys, xs)
