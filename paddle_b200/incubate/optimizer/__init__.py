"""paddle.incubate.optimizer (reference: python/paddle/incubate/optimizer/{lookahead,modelaverage}.py)."""
from __future__ import annotations

import torch


class LookAhead:
    """k fast steps of the inner optimizer, then slow <- slow + alpha (fast - slow), fast <- slow."""

    def __init__(self, inner_optimizer, alpha=0.5, k=5, name=None):
        assert 0.0 <= alpha <= 1.0 and k >= 1
        self.inner_optimizer, self.alpha, self.k = inner_optimizer, alpha, k
        self._step = 0
        self._slow = {}

    def _params(self):
        return [p for p, _ in self.inner_optimizer._all_params()]

    @torch.no_grad()
    def step(self):
        if not self._slow:
            self._slow = {id(p): p.detach().clone() for p in self._params()}
        self.inner_optimizer.step()
        self._step += 1
        if self._step % self.k == 0:
            for p in self._params():
                s = self._slow[id(p)]
                s.add_(self.alpha * (p.detach() - s))
                p.copy_(s)

    def minimize(self, loss, startup_program=None, parameters=None, no_grad_set=None):
        loss.backward()
        self.step()

    def clear_grad(self, set_to_zero=True):
        self.inner_optimizer.clear_grad(set_to_zero)

    def get_lr(self):
        return self.inner_optimizer.get_lr()

    def state_dict(self):
        sd = self.inner_optimizer.state_dict()
        sd["@lookahead_step"] = self._step
        return sd

    def set_state_dict(self, sd):
        self._step = int(sd.pop("@lookahead_step", 0)) if isinstance(sd, dict) else 0
        self.inner_optimizer.set_state_dict(sd)


class ModelAverage:
    """Running average of the parameters over a sliding window; ``apply()`` swaps the averages in for evaluation and
    ``restore()`` puts the trained values back."""

    def __init__(self, average_window_rate, parameters=None, min_average_window=10000, max_average_window=10000, name=None):
        self.rate, self.min_w, self.max_w = average_window_rate, min_average_window, max_average_window
        self._params = list(parameters or [])
        self._sum = {id(p): torch.zeros_like(p, dtype=torch.float32) for p in self._params}
        self._n = 0
        self._updates = 0
        self._backup = None

    @torch.no_grad()
    def step(self):
        self._updates += 1
        window = min(self.max_w, max(self.min_w, int(self._updates * self.rate)))
        if self._n >= window:                     # restart the window, as the reference does when it is exceeded
            for s in self._sum.values():
                s.zero_()
            self._n = 0
        for p in self._params:
            self._sum[id(p)].add_(p.detach().float())
        self._n += 1

    def minimize(self, loss, startup_program=None, parameters=None, no_grad_set=None):
        self.step()

    def apply(self, executor=None, need_restore=True):
        outer = self

        class _Ctx:
            def __enter__(self_inner):
                outer._swap_in()
                return outer

            def __exit__(self_inner, *exc):
                if need_restore:
                    outer.restore()
                return False
        return _Ctx()

    @torch.no_grad()
    def _swap_in(self):
        self._backup = {id(p): p.detach().clone() for p in self._params}
        if self._n:
            for p in self._params:
                p.copy_((self._sum[id(p)] / self._n).to(p.dtype))

    @torch.no_grad()
    def restore(self, executor=None):
        if self._backup is not None:
            for p in self._params:
                p.copy_(self._backup[id(p)])
            self._backup = None


from ...optimizer import LBFGS  # noqa: E402,F401  (the reference keeps the old import path alive)
from . import functional  # noqa: E402,F401
