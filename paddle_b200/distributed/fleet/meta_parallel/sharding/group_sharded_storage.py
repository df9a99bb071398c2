"""Flat storages of GroupSharded stage 2 (reference: .../sharding/group_sharded_storage.py:35-363 InternalStorage /
ParamStorage / GradStorage).

* ``ParamStorage``: the parameters one rank owns (per dtype) live in one flat buffer, so the post-step parameter
  exchange is ONE broadcast per (owner, dtype) instead of one per parameter.
* ``GradStorage``: the gradients headed for one owner (per dtype, up to ``buffer_max_size`` elements) are views of one
  flat bucket; autograd accumulates straight into the views, and when the last gradient of the bucket has been
  produced the bucket leaves in ONE ``reduce``.  Non-owners only keep the bucket as a send buffer.
Differences from the reference by design: buckets are zeroed with one fill per step (not per parameter), and views are
re-attached by the wrapper when the user drops gradients with ``clear_grad(set_to_zero=False)``.
"""
from __future__ import annotations

import torch

ALIGN = 256   # bytes; every parameter starts on a 256-byte boundary inside a flat buffer


def _aligned_numel(p):
    a = ALIGN // p.element_size()
    return (p.numel() + a - 1) // a * a


class InternalStorage:
    def __init__(self, size, dtype, device):
        self._params = []
        self._fill = 0
        self._dtype = dtype
        self._device = device
        self.buffer = torch.zeros(int(size), dtype=dtype, device=device)

    def to(self, device, dtype=None):
        self.buffer = self.buffer.to(device=device, dtype=dtype or self._dtype)
        self._device = self.buffer.device
        if dtype is not None:
            self._dtype = dtype
        self._rebind()
        return self

    def _rebind(self):
        pass

    def can_add(self, p):
        return self._fill + _aligned_numel(p) <= self.buffer.numel() and p.dtype == self._dtype


class ParamStorage(InternalStorage):
    """Parameters of one owner packed into a flat buffer; ``param.data`` becomes a view of it."""

    def __init__(self, size, dtype, device):
        super().__init__(size, dtype, device)
        self._offsets = []

    @torch.no_grad()
    def add_rank_params(self, params):
        for p in params:
            assert self.can_add(p), "ParamStorage overflow / dtype mismatch"
            off = self._fill
            view = self.buffer[off:off + p.numel()].view(p.shape)
            view.copy_(p.data)
            p.data = view
            self._params.append(p)
            self._offsets.append(off)
            self._fill += _aligned_numel(p)

    @torch.no_grad()
    def _rebind(self):
        for p, off in zip(self._params, self._offsets):
            p.data = self.buffer[off:off + p.numel()].view(p.shape)


class GradStorage(InternalStorage):
    """Gradient bucket for one destination rank.  ``params_checked_in`` counts the gradients produced in the current
    backward; ``all_checked_in`` says the bucket can be reduced."""

    def __init__(self, size, dtype, device, destination, parm2align=None):
        super().__init__(size, dtype, device)
        self.destination = destination
        self._offsets = []
        self.params_checked_in = 0
        self.sent = False
        self.callback = None
        self._max_size = int(size)

    def reset_checked_in(self):
        self.params_checked_in = 0
        self.sent = False

    @property
    def all_checked_in(self):
        return len(self._params) == self.params_checked_in

    def can_add_grad_view(self, p, align=None):
        return self.can_add(p) and not any(q is p for q in self._params)

    @torch.no_grad()
    def add_grad(self, p, align=None):
        assert self.can_add_grad_view(p)
        off = self._fill
        self._params.append(p)
        self._offsets.append(off)
        self._fill += _aligned_numel(p)
        self._attach(p, off)

    def _attach(self, p, off):
        view = self.buffer[off:off + p.numel()].view(p.shape)
        if p.grad is not None and p.grad.data_ptr() != view.data_ptr():
            view.copy_(p.grad)
        p.grad = view
        p._gs_grad_bucket = self

    @torch.no_grad()
    def rebuild(self):
        """Re-attach the views (after ``clear_grad(set_to_zero=False)`` dropped them, or after ``to``)."""
        for p, off in zip(self._params, self._offsets):
            view = self.buffer[off:off + p.numel()].view(p.shape)
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                p.grad = view

    _rebind = rebuild

    @torch.no_grad()
    def zero_(self):
        self.buffer.zero_()

    def manumal_relase(self):     # spelling follows the reference API (group_sharded_storage.py:332)
        for p in self._params:
            p.grad = None
        self.buffer = torch.empty(0, dtype=self._dtype, device=self._device)

    manual_release = manumal_relase
