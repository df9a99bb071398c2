"""paddle.incubate.autograd functional API (reference: python/paddle/incubate/autograd/functional.py vjp :26, jvp :88,
Jacobian :170, Hessian :270; primapi.py forward_grad / grad).  torch.func is the engine."""
from __future__ import annotations

import torch

__all__ = ["vjp", "jvp", "Jacobian", "Hessian", "enable_prim", "disable_prim", "forward_grad", "grad"]
_prim = [False]


def _tup(x):
    return tuple(x) if isinstance(x, (list, tuple)) else (x,)


def vjp(func, xs, v=None):
    """(func(xs), v^T J): ``v`` defaults to ones like the output."""
    xs_t = _tup(xs)
    ys, pull = torch.func.vjp(func, *xs_t)
    v = v if v is not None else (tuple(torch.ones_like(y) for y in ys) if isinstance(ys, (tuple, list)) else torch.ones_like(ys))
    out = pull(tuple(v) if isinstance(v, list) else v)
    return ys, (out if isinstance(xs, (list, tuple)) else out[0])


def jvp(func, xs, v=None):
    """(func(xs), J v): ``v`` defaults to ones like the inputs."""
    xs_t = _tup(xs)
    v_t = _tup(v) if v is not None else tuple(torch.ones_like(x) for x in xs_t)
    ys, out = torch.func.jvp(func, xs_t, v_t)
    return ys, out


class Jacobian:
    """Jacobian of ``func`` at ``xs`` (inputs and outputs flattened and concatenated): [ny, nx] or, ``is_batched``,
    [B, ny, nx].  Evaluated on first indexing."""

    def __init__(self, func, xs, is_batched=False):
        self._func, self._xs, self._batched = func, _tup(xs), is_batched
        self._full = None

    def _flat_out(self, *xs):
        ys = _tup(self._func(*xs))
        if self._batched:
            return torch.cat([y.reshape(y.shape[0], -1) for y in ys], 1)
        return torch.cat([y.reshape(-1) for y in ys])

    def _evaluate(self):
        if self._full is None:
            jac = torch.func.jacrev(self._flat_out, argnums=tuple(range(len(self._xs))))(*self._xs)
            if self._batched:
                B = self._xs[0].shape[0]
                # jacrev gives [B, ny, B, ...]: batch rows are independent, take the diagonal over the two batch dims
                parts = [j.reshape(B, -1, B, x[0].numel())[torch.arange(B), :, torch.arange(B)] for j, x in zip(jac, self._xs)]
                self._full = torch.cat(parts, 2)
            else:
                ny = jac[0].shape[0] if jac[0].dim() else 1
                self._full = torch.cat([j.reshape(ny, -1) for j in jac], 1)
        return self._full

    @property
    def shape(self):
        return tuple(self._evaluate().shape)

    def __getitem__(self, idx):
        return self._evaluate()[idx]


class Hessian(Jacobian):
    def __init__(self, func, xs, is_batched=False):
        xs_t = _tup(xs)

        def grad_fn(*a):
            if is_batched:
                g = torch.func.grad(lambda *b: func(*b).sum(), argnums=tuple(range(len(a))))(*a)
            else:
                g = torch.func.grad(lambda *b: func(*b).reshape(()), argnums=tuple(range(len(a))))(*a)
            return torch.cat([t.reshape(t.shape[0], -1) for t in g], 1) if is_batched else torch.cat([t.reshape(-1) for t in g])
        super().__init__(grad_fn, xs_t, is_batched)


def enable_prim():
    _prim[0] = True


def disable_prim():
    _prim[0] = False


def prim_enabled():
    return _prim[0]


def forward_grad(outputs, inputs, grad_inputs=None):
    raise RuntimeError("forward_grad works on static-graph programs with primitive operators, which this framework does "
                       "not have: use paddle.incubate.autograd.jvp(func, xs, v)")


def grad(outputs, inputs, grad_outputs=None):
    single = isinstance(inputs, torch.Tensor)
    res = torch.autograd.grad(outputs, [inputs] if single else list(inputs), grad_outputs, allow_unused=True)
    return res[0] if single else list(res)
