"""Activation recompute (reference: python/paddle/distributed/fleet/recompute/recompute.py:109-620).
A reentrant autograd Function that drops the block's activations, replays forward in backward with the saved RNG
state (global stream AND the model-parallel tracker's named states) and the saved autocast state; the
non-reentrant variant uses saved-tensor hooks (torch.utils.checkpoint)."""
from __future__ import annotations

import contextlib

import torch
import torch.utils.checkpoint as tcp

from paddle_b200.framework import random as _rnd


def _tracker():
    from paddle_b200.distributed.fleet.layers.mpu.random import get_rng_state_tracker
    return get_rng_state_tracker()


@contextlib.contextmanager
def switch_rng_state_tracker(rng_state, tracker_states):
    tr = _tracker()
    orig_rng = _rnd.current_rng_state()
    orig_tracker = tr.get_states_tracker()
    _rnd.set_current_rng_state(rng_state)
    tr.set_states_tracker(tracker_states)
    try:
        yield
    finally:
        _rnd.set_current_rng_state(orig_rng)
        tr.set_states_tracker(orig_tracker)


class RecomputeFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, run_function, preserve_rng_state, n_args, *args):
        ctx.run_function = run_function
        ctx.preserve_rng_state = preserve_rng_state
        tensors, others, idx = [], [], []
        for i, a in enumerate(args):
            if isinstance(a, torch.Tensor):
                tensors.append(a)
                idx.append(i)
                others.append(None)
            else:
                others.append(a)
        ctx.tensor_idx, ctx.others = idx, others
        ctx.save_for_backward(*tensors)
        if preserve_rng_state:
            ctx.fw_rng_state = _rnd.current_rng_state()
            ctx.fw_tracker = _tracker().get_states_tracker()
        ctx.autocast = (torch.is_autocast_enabled("cuda"), torch.get_autocast_dtype("cuda")) \
            if torch.cuda.is_available() else (False, torch.bfloat16)
        with torch.no_grad():
            out = run_function(*args)
        return out

    @staticmethod
    def backward(ctx, *grads):
        tensors = ctx.saved_tensors
        args = list(ctx.others)
        detached = []
        for i, t in zip(ctx.tensor_idx, tensors):
            d = t.detach()
            d.requires_grad_(t.requires_grad)
            args[i] = d
            detached.append(d)
        rng_ctx = switch_rng_state_tracker(ctx.fw_rng_state, ctx.fw_tracker) if ctx.preserve_rng_state \
            else contextlib.nullcontext()
        ac_ctx = torch.autocast("cuda", dtype=ctx.autocast[1], enabled=ctx.autocast[0]) if torch.cuda.is_available() \
            else contextlib.nullcontext()
        with rng_ctx, torch.enable_grad(), ac_ctx:
            outputs = ctx.run_function(*args)
        if isinstance(outputs, torch.Tensor):
            outputs = (outputs,)
        outs, gs = [], []
        for o, g in zip(outputs, grads):
            if isinstance(o, torch.Tensor) and o.requires_grad and g is not None:
                outs.append(o)
                gs.append(g)
        if not outs:
            raise RuntimeError("none of output has requires_grad=True, this recompute() is not necessary")
        torch.autograd.backward(outs, gs)
        in_grads = [d.grad if d.requires_grad else None for d in detached]
        full = [None] * len(ctx.others)
        for i, g in zip(ctx.tensor_idx, in_grads):
            full[i] = g
        return (None, None, None, *full)


def recompute(function, *args, **kwargs):
    preserve = kwargs.pop("preserve_rng_state", True)
    use_reentrant = kwargs.pop("use_reentrant", True)
    if not torch.is_grad_enabled():
        return function(*args, **kwargs)
    if kwargs and use_reentrant:
        fn = lambda *a: function(*a, **kwargs)  # noqa: E731
    else:
        fn = function
    if use_reentrant:
        return RecomputeFunction.apply(fn, preserve, len(args), *args)
    return tcp.checkpoint(function, *args, use_reentrant=False, preserve_rng_state=preserve, **kwargs)


def recompute_sequential(ctx, functions, *args, **kwargs):
    segments = ctx.get("segments", 1)
    preserve = ctx.get("preserve_rng_state", True)
    if isinstance(functions, torch.nn.Module) and not isinstance(functions, (list, tuple)):
        functions = list(functions.children())

    def run(begin, end, funcs):
        def do_run(inp):
            for i in range(begin, end + 1):
                inp = funcs[i](inp)
            return inp
        return do_run

    seg_size = len(functions) // segments
    end = -1
    x = args[0] if len(args) == 1 else args
    for begin in range(0, seg_size * (segments - 1), seg_size):
        end = begin + seg_size - 1
        x = recompute(run(begin, end, functions), x, preserve_rng_state=preserve, **kwargs)
    return run(end + 1, len(functions) - 1, functions)(x)
