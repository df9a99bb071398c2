"""LR schedulers (reference: python/paddle/optimizer/lr.py).  ``scheduler()`` returns the current lr,
``step()`` advances the epoch counter; state_dict carries ``last_epoch``/``last_lr`` like the reference."""
from __future__ import annotations

import math


class LRScheduler:
    def __init__(self, learning_rate=0.1, last_epoch=-1, verbose=False):
        self.base_lr = float(learning_rate)
        self.last_lr = float(learning_rate)
        self.last_epoch = last_epoch
        self.verbose = verbose
        self.step()

    def __call__(self):
        return self.last_lr

    def step(self, epoch=None):
        if epoch is None:
            self.last_epoch += 1
        else:
            self.last_epoch = epoch
        self.last_lr = self.get_lr()

    def get_lr(self):
        return self.base_lr

    def state_keys(self):
        """Names of the attributes ``state_dict`` saves; subclasses with more state override this."""
        self.keys = ["last_epoch", "last_lr"]

    def state_dict(self):
        """Reference: python/paddle/optimizer/lr.py:178-199 -- the attributes ``state_keys`` names, tensors as floats."""
        self.state_keys()
        out = {}
        for key in self.keys:
            if key not in self.__dict__:
                continue
            value = self.__dict__[key]
            if hasattr(value, "numel") and hasattr(value, "item"):
                assert value.numel() == 1, "numel of Tensor in state_dict must be 1"
                value = float(value.item())
            out[key] = value
        return out

    def set_state_dict(self, sd):
        self.state_keys()
        for key in self.keys:
            if key in sd:
                self.__dict__[key] = sd[key]
            else:
                raise RuntimeError(f"Please check whether state_dict is correct for optimizer. Can't find [ {key} ] in state_dict")

    set_dict = set_state_dict


class NoamDecay(LRScheduler):
    def __init__(self, d_model, warmup_steps, learning_rate=1.0, last_epoch=-1, verbose=False):
        self.d_model, self.warmup_steps = d_model, warmup_steps
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        a = 1 if self.last_epoch == 0 else self.last_epoch ** -0.5
        b = self.warmup_steps ** -1.5 * self.last_epoch
        return self.base_lr * (self.d_model ** -0.5) * min(a, b)


class PiecewiseDecay(LRScheduler):
    def __init__(self, boundaries, values, last_epoch=-1, verbose=False):
        self.boundaries, self.values = boundaries, values
        super().__init__(values[0], last_epoch, verbose)

    def get_lr(self):
        for i, b in enumerate(self.boundaries):
            if self.last_epoch < b:
                return self.values[i]
        return self.values[len(self.values) - 1]


class NaturalExpDecay(LRScheduler):
    def __init__(self, learning_rate, gamma, last_epoch=-1, verbose=False):
        self.gamma = gamma
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        return self.base_lr * math.exp(-1 * self.gamma * self.last_epoch)


class ExponentialDecay(LRScheduler):
    def __init__(self, learning_rate, gamma, last_epoch=-1, verbose=False):
        self.gamma = gamma
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        return self.base_lr * (self.gamma ** self.last_epoch)


class InverseTimeDecay(LRScheduler):
    def __init__(self, learning_rate, gamma, last_epoch=-1, verbose=False):
        self.gamma = gamma
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        return self.base_lr / (1 + self.gamma * self.last_epoch)


class PolynomialDecay(LRScheduler):
    def __init__(self, learning_rate, decay_steps, end_lr=0.0001, power=1.0, cycle=False, last_epoch=-1, verbose=False):
        self.decay_steps, self.end_lr, self.power, self.cycle = decay_steps, end_lr, power, cycle
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        e, ds = self.last_epoch, self.decay_steps
        if self.cycle:
            div = math.ceil(e / float(ds)) or 1
            ds = ds * div
        else:
            e = min(e, ds)
        return (self.base_lr - self.end_lr) * ((1 - float(e) / float(ds)) ** self.power) + self.end_lr


class LinearWarmup(LRScheduler):
    def __init__(self, learning_rate, warmup_steps, start_lr, end_lr, last_epoch=-1, verbose=False):
        self.learning_rate_obj = learning_rate
        self.warmup_steps, self.start_lr, self.end_lr = warmup_steps, start_lr, end_lr
        super().__init__(start_lr, last_epoch, verbose)

    def get_lr(self):
        if self.last_epoch < self.warmup_steps:
            return (self.end_lr - self.start_lr) * float(self.last_epoch) / float(self.warmup_steps) + self.start_lr
        if isinstance(self.learning_rate_obj, LRScheduler):
            self.learning_rate_obj.step(self.last_epoch - self.warmup_steps)
            return self.learning_rate_obj()
        return float(self.learning_rate_obj)

    def state_dict(self):
        sd = super().state_dict()
        if isinstance(self.learning_rate_obj, LRScheduler):
            sd["LinearWarmup_LR"] = self.learning_rate_obj.state_dict()
        return sd

    def set_state_dict(self, sd):
        super().set_state_dict(sd)
        if isinstance(self.learning_rate_obj, LRScheduler) and "LinearWarmup_LR" in sd:
            self.learning_rate_obj.set_state_dict(sd["LinearWarmup_LR"])


class StepDecay(LRScheduler):
    def __init__(self, learning_rate, step_size, gamma=0.1, last_epoch=-1, verbose=False):
        self.step_size, self.gamma = step_size, gamma
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        return self.base_lr * (self.gamma ** (self.last_epoch // self.step_size))


class MultiStepDecay(LRScheduler):
    def __init__(self, learning_rate, milestones, gamma=0.1, last_epoch=-1, verbose=False):
        self.milestones, self.gamma = milestones, gamma
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        n = sum(1 for m in self.milestones if self.last_epoch >= m)
        return self.base_lr * (self.gamma ** n)


class LambdaDecay(LRScheduler):
    def __init__(self, learning_rate, lr_lambda, last_epoch=-1, verbose=False):
        self.lr_lambda = lr_lambda
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        return self.base_lr * self.lr_lambda(self.last_epoch)


class CosineAnnealingDecay(LRScheduler):
    def __init__(self, learning_rate, T_max, eta_min=0, last_epoch=-1, verbose=False):
        self.T_max, self.eta_min = T_max, float(eta_min)
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        return self.eta_min + (self.base_lr - self.eta_min) * (1 + math.cos(math.pi * self.last_epoch / self.T_max)) / 2


class OneCycleLR(LRScheduler):
    """One-cycle policy.  Phases are (first step, last step, start lr, end lr): two phases - up to the peak at step
    ``phase_pct * total - 1``, then down to ``end_learning_rate`` at the last step - or, ``three_phase``, up, back down to the
    initial rate in as many steps, then down to the end rate (boundaries as in the reference, lr.py OneCycleLR)."""

    def __init__(self, max_learning_rate, total_steps, divide_factor=25.0, end_learning_rate=0.0001, phase_pct=0.3,
                 anneal_strategy="cos", three_phase=False, last_epoch=-1, verbose=False):
        if total_steps <= 0 or not 0 <= phase_pct <= 1:
            raise ValueError("OneCycleLR: total_steps must be positive and phase_pct in [0, 1]")
        if anneal_strategy not in ("cos", "linear"):
            raise ValueError(f"'anneal_strategy' must by one of 'cos' or 'linear', but received {anneal_strategy}")
        init, peak, low, last = max_learning_rate / float(divide_factor), float(max_learning_rate), float(end_learning_rate), total_steps - 1
        top = phase_pct * total_steps - 1
        if three_phase:
            if phase_pct >= 0.5:
                raise ValueError("When three_phase is True, 'phase_pct' must be less than 0.5")
            back = 2 * phase_pct * total_steps - 2
            self._phases = [(0, top, init, peak), (top, back, peak, init), (back, last, init, low)]
        else:
            self._phases = [(0, top, init, peak), (top, last, peak, low)]
        self.total_steps, self._cos = total_steps, anneal_strategy == "cos"
        super().__init__(init, last_epoch, verbose)

    def get_lr(self):
        e = self.last_epoch
        if e > self.total_steps:
            raise ValueError(f"Tried to step {e} times. However the number of total steps is {self.total_steps}")
        for k, (a, b, lr_a, lr_b) in enumerate(self._phases):
            if e <= b or k == len(self._phases) - 1:
                pct = (e - a) / (b - a) if b != a else 1.0
                if self._cos:
                    return lr_b + (lr_a - lr_b) / 2.0 * (math.cos(math.pi * pct) + 1)
                return (lr_b - lr_a) * pct + lr_a


class ReduceOnPlateau(LRScheduler):
    def __init__(self, learning_rate, mode="min", factor=0.1, patience=10, threshold=1e-4, threshold_mode="rel",
                 cooldown=0, min_lr=0, epsilon=1e-8, verbose=False):
        self.mode, self.factor, self.patience, self.threshold = mode, factor, patience, threshold
        self.threshold_mode, self.cooldown, self.min_lr, self.epsilon = threshold_mode, cooldown, min_lr, epsilon
        self.best, self.num_bad_epochs, self.cooldown_counter = None, 0, 0
        self.base_lr = self.last_lr = float(learning_rate)
        self.last_epoch = 0
        self.verbose = verbose

    def state_keys(self):
        self.keys = ["cooldown_counter", "best", "num_bad_epochs", "last_epoch", "last_lr"]

    def _is_better(self, current, best):
        # reference: python/paddle/optimizer/lr.py:1556-1567 -- the relative margin is best * threshold, sign and all
        margin = best * self.threshold if self.threshold_mode == "rel" else self.threshold
        return current < best - margin if self.mode == "min" else current > best + margin

    def step(self, metrics=None, epoch=None):
        """Reference: python/paddle/optimizer/lr.py:1501-1554.  While cooling down the metric is not looked at at all
        (neither ``best`` nor the bad-epoch count moves); otherwise a non-improving epoch counts as bad and the rate
        drops once more than ``patience`` of them have piled up."""
        self.last_epoch = self.last_epoch + 1 if epoch is None else epoch
        if metrics is None:
            return
        if hasattr(metrics, "size") and not isinstance(metrics, (int, float)):
            size = metrics.numel() if hasattr(metrics, "numel") else metrics.size
            assert size == 1, f"the size of metrics must be 1, but the current metrics.size is {size}"
        elif not isinstance(metrics, (int, float)):
            raise TypeError(f"metrics must be a number, a numpy.ndarray or a Tensor, but received {type(metrics)}")
        m = float(metrics.item() if hasattr(metrics, "item") else metrics)
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            return
        if self.best is None or self._is_better(m, self.best):
            self.best, self.num_bad_epochs = m, 0
        else:
            self.num_bad_epochs += 1
        if self.num_bad_epochs > self.patience:
            self.cooldown_counter, self.num_bad_epochs = self.cooldown, 0
            new_lr = max(self.last_lr * self.factor, self.min_lr)
            if self.last_lr - new_lr > self.epsilon:
                self.last_lr = new_lr
                if self.verbose:
                    print(f"Epoch {self.last_epoch}: {type(self).__name__} set learning rate to {self.last_lr}.")


class MultiplicativeDecay(LRScheduler):
    def __init__(self, learning_rate, lr_lambda, last_epoch=-1, verbose=False):
        self.lr_lambda = lr_lambda
        self._cur = float(learning_rate)
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        cur = self.base_lr
        for e in range(1, self.last_epoch + 1):
            cur = cur * self.lr_lambda(e)
        return cur


class LinearLR(LRScheduler):
    def __init__(self, learning_rate, total_steps, start_factor=1.0 / 3, end_factor=1.0, last_epoch=-1, verbose=False):
        self.total_steps, self.start_factor, self.end_factor = total_steps, start_factor, end_factor
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        t = min(max(self.last_epoch, 0), self.total_steps)
        return self.base_lr * (self.start_factor + (self.end_factor - self.start_factor) * t / self.total_steps)


class CosineAnnealingWarmRestarts(LRScheduler):
    def __init__(self, learning_rate, T_0, T_mult=1, eta_min=0, last_epoch=-1, verbose=False):
        self.T_0, self.T_mult, self.eta_min = T_0, T_mult, eta_min
        super().__init__(learning_rate, last_epoch, verbose)

    def get_lr(self):
        import math
        e, T = max(self.last_epoch, 0), self.T_0
        if self.T_mult == 1:
            cur = e % T
        else:
            n = int(math.log(e / self.T_0 * (self.T_mult - 1) + 1, self.T_mult)) if e >= self.T_0 else 0
            cur = e - self.T_0 * (self.T_mult ** n - 1) // (self.T_mult - 1)
            T = self.T_0 * self.T_mult ** n
        return self.eta_min + (self.base_lr - self.eta_min) * (1 + math.cos(math.pi * cur / T)) / 2


class CyclicLR(LRScheduler):
    def __init__(self, base_learning_rate, max_learning_rate, step_size_up, step_size_down=None, mode="triangular",
                 exp_gamma=1.0, scale_fn=None, scale_mode="cycle", last_epoch=-1, verbose=False):
        self.max_lr = max_learning_rate
        self.up = float(step_size_up)
        self.down = float(step_size_down) if step_size_down is not None else self.up
        self.mode, self.gamma, self.scale_fn, self.scale_mode = mode, exp_gamma, scale_fn, scale_mode
        super().__init__(base_learning_rate, last_epoch, verbose)

    def get_lr(self):
        import math
        total = self.up + self.down
        e = max(self.last_epoch, 0)
        cycle = math.floor(1 + e / total)
        x = 1.0 + e / total - cycle
        pct = x / (self.up / total) if x <= self.up / total else (x - 1) / (self.up / total - 1)
        amp = (self.max_lr - self.base_lr) * pct
        if self.scale_fn is not None:
            return self.base_lr + amp * self.scale_fn(cycle if self.scale_mode == "cycle" else e)
        if self.mode == "triangular2":
            return self.base_lr + amp / (2.0 ** (cycle - 1))
        if self.mode == "exp_range":
            return self.base_lr + amp * self.gamma ** e
        return self.base_lr + amp
