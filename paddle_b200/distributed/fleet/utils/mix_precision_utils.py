"""main_grad mixed precision (reference: python/paddle/distributed/fleet/utils/mix_precision_utils.py:35-250).
``MixPrecisionLayer`` gives every low-precision parameter an fp32 ``main_grad`` that accumulates the bf16 grads;
Linear layers of this framework accumulate straight into it in the wgrad GEMM epilogue (K13) when
``param._fuse_main_grad`` is set, so the bf16 dW tensor is never materialised."""
from __future__ import annotations

import torch

from paddle_b200.nn.layer.layers import Layer


class MixPrecisionLayer(Layer):
    def __init__(self, layers, dtype="float16"):
        super().__init__()
        self._layers = layers
        self._dtype_str = dtype
        for p in self._layers.parameters():
            if not p.requires_grad:
                continue
            p.main_grad = None
            if p.dim() == 2 and p.dtype in (torch.bfloat16, torch.float16):
                p._fuse_main_grad = True
            p.register_post_accumulate_grad_hook(_main_grad_hook)

    def forward(self, *inputs, **kwargs):
        return self._layers(*inputs, **kwargs)

    def state_dict(self, *a, **k):
        return self._layers.state_dict(*a, **k)

    def set_state_dict(self, sd, use_structured_name=True):
        return self._layers.set_state_dict(sd, use_structured_name)

    def parameters(self, *a, **k):
        return self._layers.parameters(*a, **k)


@torch.no_grad()
def _main_grad_hook(p):
    if p.grad is None:
        return
    if getattr(p, "main_grad", None) is None:
        p.main_grad = p.grad.float()
    else:
        p.main_grad.add_(p.grad)
    p.grad = None


def ensure_main_grad(p):
    if getattr(p, "main_grad", None) is None:
        p.main_grad = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
    return p.main_grad


class MixPrecisionOptimizer:
    def __init__(self, optimizer):
        self._inner_opt = optimizer
        optimizer._multi_precision = True

    def step(self):
        return self._inner_opt.step()

    def clear_grad(self, set_to_zero=True):
        for p in self._inner_opt._parameter_list if not isinstance(self._inner_opt._parameter_list[0], dict) else \
                [q for g in self._inner_opt._parameter_list for q in g["params"]]:
            mg = getattr(p, "main_grad", None)
            if mg is not None:
                if set_to_zero:
                    mg.zero_()
                else:
                    p.main_grad = None
            p.grad = None

    clear_gradients = clear_grad

    def __getattr__(self, item):
        return getattr(self.__dict__["_inner_opt"], item)


class MixPrecisionScaler:
    def __init__(self, scaler):
        self._inner = scaler

    def __getattr__(self, item):
        return getattr(self.__dict__["_inner"], item)


def unscale_method(self, optimizer):
    return self.unscale_(optimizer)
