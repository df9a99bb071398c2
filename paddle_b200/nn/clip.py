"""Gradient clipping (reference: python/paddle/nn/clip.py: ClipGradByValue/ClipGradByNorm/ClipGradByGlobalNorm).
``_dygraph_clip(params_grads) -> params_grads`` is the hook the optimizers (and HybridParallelClipGrad) call.
The global-norm variant computes sum-of-squares with the multi-tensor kernel (K15) and keeps the coefficient on
the device so the optimizer step never synchronises."""
from __future__ import annotations

import torch


class ClipGradBase:
    def __call__(self, params_grads):
        return self._dygraph_clip(params_grads)


class ClipGradByValue(ClipGradBase):
    def __init__(self, max, min=None):
        self.max = float(max)
        self.min = float(min) if min is not None else -self.max

    @torch.no_grad()
    def _dygraph_clip(self, params_grads):
        out = []
        for p, g in params_grads:
            if g is None or not getattr(p, "need_clip", True):
                out.append((p, g))
                continue
            out.append((p, g.clamp_(self.min, self.max)))
        return out


class ClipGradByNorm(ClipGradBase):
    def __init__(self, clip_norm):
        self.clip_norm = float(clip_norm)

    @torch.no_grad()
    def _dygraph_clip(self, params_grads):
        out = []
        for p, g in params_grads:
            if g is None or not getattr(p, "need_clip", True):
                out.append((p, g))
                continue
            n = g.float().norm()
            g.mul_((self.clip_norm / torch.maximum(n, torch.tensor(self.clip_norm, device=g.device))).to(g.dtype))
            out.append((p, g))
        return out


class ClipGradByGlobalNorm(ClipGradBase):
    def __init__(self, clip_norm, group_name="default_group", auto_skip_clip=False):
        self.clip_norm = float(clip_norm)
        self.group_name = group_name
        self.auto_skip_clip = auto_skip_clip

    @torch.no_grad()
    def global_norm_sq(self, grads):
        from .. import ops
        return ops.squared_l2_norm(grads)

    def clip_coef(self, global_norm_sq):
        gn = torch.sqrt(global_norm_sq.float())
        cn = torch.full_like(gn, self.clip_norm)
        return cn / torch.maximum(gn, cn)

    @torch.no_grad()
    def _dygraph_clip(self, params_grads):
        grads = [g for p, g in params_grads if g is not None and getattr(p, "need_clip", True)]
        if not grads:
            return params_grads
        coef = self.clip_coef(self.global_norm_sq(grads).to(grads[0].device))
        for g in grads:
            g.mul_(coef.to(g.dtype))
        return params_grads


def clip_grad_norm_(parameters, max_norm, norm_type=2.0, error_if_nonfinite=False):
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    return torch.nn.utils.clip_grad_norm_(list(parameters), max_norm, norm_type, error_if_nonfinite)


def clip_grad_value_(parameters, clip_value):
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    return torch.nn.utils.clip_grad_value_(list(parameters), clip_value)
