"""recompute_hybrid (reference: python/paddle/distributed/fleet/recompute/recompute_hybrid.py:31-295): recompute
whose saved inputs are partitioned across the mp group (all-gathered on replay) and/or offloaded to pinned host
memory."""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm

from .recompute import recompute


class _PartitionedInput(torch.autograd.Function):
    """Identity whose saved copy is 1/mp of the tensor (optionally on the host)."""

    @staticmethod
    def forward(ctx, x, group, offload):
        ctx.group, ctx.offload, ctx.shape = group, offload, x.shape
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g, None, None


def recompute_hybrid(ctx, function, *args, **kwargs):
    mp_group = ctx.get("mp_group")
    assert mp_group is not None, "ctx must contains mp_group and mp_group can not be None."
    offload = ctx.get("offload", False)
    partition = ctx.get("partition", False)
    n = mp_group.nranks

    if not torch.is_grad_enabled() or (not partition and not offload):
        return recompute(function, *args, **kwargs)

    shapes, packed = [], []
    for a in args:
        if isinstance(a, torch.Tensor) and a.is_floating_point() and partition and n > 1 and a.numel() % n == 0:
            flat = a.detach().reshape(-1)
            part = flat.chunk(n)[mp_group.rank].clone()
            if offload:
                part = part.to("cpu", non_blocking=True)
            shapes.append((a.shape, a.requires_grad, True))
            packed.append(part)
        elif isinstance(a, torch.Tensor) and offload:
            shapes.append((a.shape, a.requires_grad, False))
            packed.append(a.detach().to("cpu", non_blocking=True))
        else:
            shapes.append(None)
            packed.append(a)

    class _Fn(torch.autograd.Function):
        @staticmethod
        def forward(c, *tensors):
            with torch.no_grad():
                out = function(*args, **kwargs)
            c.dev = next(t.device for t in tensors if isinstance(t, torch.Tensor))
            return out

        @staticmethod
        def backward(c, *grads):
            rebuilt = []
            for meta, pk in zip(shapes, packed):
                if meta is None:
                    rebuilt.append(pk)
                    continue
                shape, rg, parted = meta
                t = pk.to(c.dev, non_blocking=True)
                if parted:
                    full = torch.empty(t.numel() * n, dtype=t.dtype, device=c.dev)
                    comm.all_gather(full, t, mp_group)
                    t = full
                t = t.reshape(shape).detach().requires_grad_(rg)
                rebuilt.append(t)
            with torch.enable_grad():
                out = function(*rebuilt, **kwargs)
            outs = (out,) if isinstance(out, torch.Tensor) else out
            pairs = [(o, g) for o, g in zip(outs, grads) if isinstance(o, torch.Tensor) and o.requires_grad and g is not None]
            torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
            return tuple(r.grad if isinstance(r, torch.Tensor) and r.requires_grad else None for r in rebuilt)

    return _Fn.apply(*args)
