"""paddle.amp (reference: python/paddle/amp/{auto_cast,grad_scaler}.py; kernels paddle/phi/kernels/gpu/amp_kernel.cu).

O1 = torch.autocast with the white/black lists; O2 = parameters cast by ``decorate`` + fp32 master weights inside
the optimizer (``multi_precision``).  GradScaler keeps scale / found_inf on the device: unscale+check is one
multi-tensor kernel, the optimizer kernel reads found_inf to skip, and the scale update is a few device ops -
there is no ``.item()`` in the step."""
from __future__ import annotations

import contextlib

import torch

from .. import ops
from ..framework import dtype as _dt
from ..framework.place import current_place
from . import debugging  # noqa: F401

_state = {"enable": False, "level": "O0", "dtype": "float32", "white": set(), "black": set()}


def is_float16_supported(device=None):
    return torch.cuda.is_available()


def is_bfloat16_supported(device=None):
    return True


# Paddle op names (what the white / black lists hold) -> names of the torch-level functions that implement them here
_OP_ALIASES = {
    "matmul": {"matmul", "mm", "bmm", "linear", "addmm", "baddbmm", "einsum"}, "matmul_v2": {"matmul", "mm", "bmm", "linear"},
    "mul": {"matmul", "mm"}, "conv2d": {"conv2d"}, "conv3d": {"conv3d"}, "conv1d": {"conv1d"},
    "conv2d_transpose": {"conv_transpose2d"}, "depthwise_conv2d": {"conv2d"},
    "elementwise_add": {"add", "__add__", "__radd__"}, "elementwise_sub": {"sub", "__sub__", "__rsub__"},
    "elementwise_mul": {"mul", "multiply", "__mul__", "__rmul__"}, "elementwise_div": {"div", "true_divide", "__truediv__"},
    "softmax": {"softmax"}, "log_softmax": {"log_softmax"}, "softmax_with_cross_entropy": {"cross_entropy"},
    "cross_entropy": {"cross_entropy", "nll_loss"}, "layer_norm": {"layer_norm"}, "batch_norm": {"batch_norm"},
    "group_norm": {"group_norm"}, "reduce_sum": {"sum"}, "reduce_mean": {"mean"}, "mean": {"mean"}, "sum": {"sum"},
    "exp": {"exp"}, "log": {"log"}, "pow": {"pow", "__pow__"}, "square": {"square"}, "sqrt": {"sqrt"}, "rsqrt": {"rsqrt"},
    "tanh": {"tanh"}, "sigmoid": {"sigmoid"}, "relu": {"relu"}, "gelu": {"gelu"}, "silu": {"silu"}, "dropout": {"dropout"},
    "cumsum": {"cumsum"}, "scale": {"mul"}, "lookup_table_v2": {"embedding"}, "embedding": {"embedding"},
    "flash_attn": {"scaled_dot_product_attention"}, "scaled_dot_product_attention": {"scaled_dot_product_attention"},
}


def _torch_names(op_names):
    out = set()
    for n in op_names or ():
        out |= _OP_ALIASES.get(n, {n})
    return out


class _AmpListMode(torch.overrides.TorchFunctionMode):
    """User lists on top of torch.autocast: a black-listed op runs in float32 with autocast off, a white-listed op has
    its floating-point tensor arguments cast to the AMP dtype."""

    def __init__(self, white, black, dtype, device_type):
        super().__init__()
        self.white, self.black, self.dtype, self.device_type = white, black, dtype, device_type

    @staticmethod
    def _cast(obj, dtype):
        if isinstance(obj, torch.Tensor):
            return obj.to(dtype) if obj.is_floating_point() and obj.dtype != dtype else obj
        if isinstance(obj, (list, tuple)):
            return type(obj)(_AmpListMode._cast(o, dtype) for o in obj)
        if isinstance(obj, dict):
            return {k: _AmpListMode._cast(v, dtype) for k, v in obj.items()}
        return obj

    def __torch_function__(self, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        name = getattr(func, "__name__", "")
        if name in self.black:
            with torch.autocast(device_type=self.device_type, enabled=False):
                return func(*self._cast(args, torch.float32), **self._cast(kwargs, torch.float32))
        if name in self.white:
            return func(*self._cast(args, self.dtype), **self._cast(kwargs, self.dtype))
        return func(*args, **kwargs)


@contextlib.contextmanager
def auto_cast(enable=True, custom_white_list=None, custom_black_list=None, level="O1", dtype="float16", use_promote=True):
    """reference: python/paddle/amp/auto_cast.py amp_guard.  ``custom_white_list`` / ``custom_black_list`` hold Paddle op
    names (``matmul``, ``softmax``, ``elementwise_add`` ...); an op may not be in both."""
    if level not in ("O0", "OD", "O1", "O2"):
        raise ValueError("level should be O0, OD, O1 or O2.")
    if dtype not in ("float16", "bfloat16") and _dt.to_torch_dtype(dtype) not in (torch.float16, torch.bfloat16):
        raise ValueError("dtype should be 'float16' or 'bfloat16'.")
    white_ops, black_ops = set(custom_white_list or ()), set(custom_black_list or ())
    both = white_ops & black_ops
    if both:
        raise ValueError(f"Custom white list overlap custom black list: {sorted(both)}")
    td = _dt.to_torch_dtype(dtype)
    active = bool(enable) and level != "O0"
    dev = "cuda" if current_place().is_gpu_place() else "cpu"
    if dev == "cpu" and td == torch.float16:
        td = torch.bfloat16
    prev = dict(_state)
    _state.update(enable=active, level=level if active else "O0", dtype=_dt.convert_dtype(td),
                  white=white_ops if active else set(), black=black_ops if active else set())
    try:
        with torch.autocast(device_type=dev, dtype=td, enabled=active):
            if active and (white_ops or black_ops):
                with _AmpListMode(_torch_names(white_ops), _torch_names(black_ops), td, dev):
                    yield
            else:
                yield
    finally:
        _state.clear()
        _state.update(prev)


amp_guard = auto_cast


def amp_state():
    return dict(_state)


def decorate(models, optimizers=None, level="O1", dtype="float16", master_weight=None, save_dtype=None,
             master_grad=False, excluded_layers=None):
    """O2: cast model parameters to ``dtype`` (norm layers stay fp32 unless bf16, like the reference keeps
    BatchNorm/LayerNorm in fp32 for fp16) and switch the optimizers to multi_precision (fp32 master weights)."""
    single_model = not isinstance(models, (list, tuple))
    model_list = [models] if single_model else list(models)
    if level == "O2":
        from ..nn.layer.common import _BatchNormBase, LayerNorm
        td = _dt.to_torch_dtype(dtype)
        excluded = tuple(excluded_layers) if excluded_layers else ()
        for m in model_list:
            keep = set()
            for sub in m.sublayers(include_self=True):
                is_excl = any((isinstance(e, type) and isinstance(sub, e)) or sub is e for e in excluded)
                if isinstance(sub, _BatchNormBase) or (td == torch.float16 and isinstance(sub, LayerNorm)) or is_excl:
                    for p in sub.parameters(include_sublayers=False):
                        keep.add(id(p))
            for p in m.parameters():
                if id(p) not in keep and p.is_floating_point():
                    p.data = p.data.to(td)
            if master_grad:
                from ..distributed.fleet.utils.mix_precision_utils import MixPrecisionLayer  # noqa: F401
                for p in m.parameters():
                    p._fuse_main_grad = False
    if optimizers is None:
        return models
    single_opt = not isinstance(optimizers, (list, tuple))
    opt_list = [optimizers] if single_opt else list(optimizers)
    if level == "O2" and master_weight is not False:
        for o in opt_list:
            o._multi_precision = True
    return (models, optimizers)


def _optimizer_chain(optimizer):
    """The optimizer and every optimizer it wraps (HybridParallelOptimizer -> sharding -> base ...)."""
    out, o, seen = [], optimizer, set()
    while o is not None and id(o) not in seen:
        seen.add(id(o))
        out.append(o)
        d = getattr(o, "__dict__", {})
        o = d.get("_inner_opt") or d.get("_optim") or d.get("_inner_optimizer")
    return out


class GradScaler:
    def __init__(self, enable=True, init_loss_scaling=2.0 ** 16, incr_ratio=2.0, decr_ratio=0.5,
                 incr_every_n_steps=2000, decr_every_n_nan_or_inf=1, use_dynamic_loss_scaling=True):
        self._enable = bool(enable)
        self._init_loss_scaling = float(init_loss_scaling)
        self._incr_ratio, self._decr_ratio = float(incr_ratio), float(decr_ratio)
        self._incr_every_n_steps, self._decr_every_n_nan_or_inf = int(incr_every_n_steps), int(decr_every_n_nan_or_inf)
        self._use_dynamic_loss_scaling = bool(use_dynamic_loss_scaling)
        self._scale = None
        self._found_inf = None
        self._good = None
        self._bad = None
        self._unscaled = set()
        self._tables = {}

    # -- device state is created lazily on the device of the first loss
    # --- the reference's accessors for the dynamic-scaling knobs
    def get_incr_ratio(self):
        return self._incr_ratio

    def set_incr_ratio(self, new_incr_ratio):
        assert new_incr_ratio > 1.0, "The new_incr_ratio must be > 1.0."
        self._incr_ratio = float(new_incr_ratio)

    def get_decr_ratio(self):
        return self._decr_ratio

    def set_decr_ratio(self, new_decr_ratio):
        assert new_decr_ratio < 1.0, "The new_decr_ratio must be < 1.0."
        self._decr_ratio = float(new_decr_ratio)

    def get_incr_every_n_steps(self):
        return self._incr_every_n_steps

    def set_incr_every_n_steps(self, new_incr_every_n_steps):
        self._incr_every_n_steps = int(new_incr_every_n_steps)

    def get_decr_every_n_nan_or_inf(self):
        return self._decr_every_n_nan_or_inf

    def set_decr_every_n_nan_or_inf(self, new_decr_every_n_nan_or_inf):
        self._decr_every_n_nan_or_inf = int(new_decr_every_n_nan_or_inf)

    def _init_state(self, device):
        if self._scale is None:
            self._scale = torch.full((1,), self._init_loss_scaling, dtype=torch.float32, device=device)
            self._found_inf = torch.zeros(1, dtype=torch.float32, device=device)
            self._good = torch.zeros(1, dtype=torch.int32, device=device)
            self._bad = torch.zeros(1, dtype=torch.int32, device=device)

    def scale(self, var):
        if not self._enable:
            return var
        self._init_state(var.device)
        return var * self._scale.to(var.dtype)

    @torch.no_grad()
    def unscale_(self, optimizer):
        if not self._enable or id(optimizer) in self._unscaled:
            return
        # A sharding optimizer whose gradients are not reduced yet (ZeRO flat buffers) takes over: unscaling a local
        # shard of an unreduced buffer would corrupt the reduction, so it unscales + checks the REDUCED shard in step().
        defer = getattr(optimizer, "_amp_defer_unscale", None)
        if defer is not None and defer(self):
            self._unscaled.add(id(optimizer))
            return
        grads = [g for _, g in optimizer._collect_params_grads()]
        if not grads:
            return
        self._unscale_grads(grads)
        self._sync_found_inf(optimizer)
        self._unscaled.add(id(optimizer))

    @torch.no_grad()
    def _unscale_grads(self, grads):
        """grads *= 1/scale in place; found_inf |= any non-finite value (device flag, no host sync)."""
        if not grads:
            return
        self._init_state(grads[0].device)
        inv = (1.0 / self._scale).float()
        from ..ops._util import use_native
        if use_native(grads[0]):
            tab = ops.grads_table(grads)
            ops.unscale_and_check(tab, inv, self._found_inf)
        else:
            for g in grads:
                bad = ~torch.isfinite(g.float()).all()
                self._found_inf.copy_(torch.maximum(self._found_inf, bad.float().reshape(1).to(self._found_inf.device)))
                g.mul_(inv.to(g.dtype).to(g.device))

    def _sync_found_inf(self, optimizer):
        """Hook for HybridParallelGradScaler: all-reduce MAX of found_inf across the check group."""

    def step(self, optimizer):
        if not self._enable:
            return optimizer.step()
        self.unscale_(optimizer)
        chain = _optimizer_chain(optimizer)      # wrappers forward reads, not writes: set the flag on every level
        for o in chain:
            o._found_inf = self._found_inf
        try:
            optimizer.step()
        finally:
            for o in chain:
                o._found_inf = None
        self._unscaled.discard(id(optimizer))

    def minimize(self, optimizer, *args, **kwargs):
        self.step(optimizer)
        self.update()

    @torch.no_grad()
    def update(self):
        if not self._enable or self._scale is None:
            return
        if self._use_dynamic_loss_scaling:
            inf = self._found_inf > 0
            self._bad.copy_(torch.where(inf, self._bad + 1, torch.zeros_like(self._bad)))
            self._good.copy_(torch.where(inf, torch.zeros_like(self._good), self._good + 1))
            dec = self._bad >= self._decr_every_n_nan_or_inf
            inc = self._good >= self._incr_every_n_steps
            new_scale = torch.where(dec, torch.clamp(self._scale * self._decr_ratio, min=1.0),
                                    torch.where(inc, self._scale * self._incr_ratio, self._scale))
            self._scale.copy_(new_scale)
            self._bad.copy_(torch.where(dec, torch.zeros_like(self._bad), self._bad))
            self._good.copy_(torch.where(inc, torch.zeros_like(self._good), self._good))
        self._found_inf.zero_()

    def is_enable(self):
        return self._enable

    def is_use_dynamic_loss_scaling(self):
        return self._use_dynamic_loss_scaling

    def get_init_loss_scaling(self):
        return self._init_loss_scaling

    def set_init_loss_scaling(self, v):
        self._init_loss_scaling = float(v)
        if self._scale is not None:
            self._scale.fill_(float(v))

    def get_loss_scaling(self):
        return self._scale if self._scale is not None else torch.tensor([self._init_loss_scaling])

    def state_dict(self):
        if not self._enable:
            return {}
        return {"scale": self.get_loss_scaling().detach().cpu().numpy(), "incr_ratio": self._incr_ratio,
                "decr_ratio": self._decr_ratio, "incr_every_n_steps": self._incr_every_n_steps,
                "decr_every_n_nan_or_inf": self._decr_every_n_nan_or_inf,
                "incr_count": int(self._good.item()) if self._good is not None else 0,
                "decr_count": int(self._bad.item()) if self._bad is not None else 0,
                "use_dynamic_loss_scaling": self._use_dynamic_loss_scaling}

    def load_state_dict(self, sd):
        if not sd:
            return
        self._init_loss_scaling = float(torch.as_tensor(sd["scale"]).reshape(-1)[0])
        if self._scale is not None:
            self._scale.fill_(self._init_loss_scaling)
        self._incr_ratio, self._decr_ratio = sd["incr_ratio"], sd["decr_ratio"]
        self._incr_every_n_steps, self._decr_every_n_nan_or_inf = sd["incr_every_n_steps"], sd["decr_every_n_nan_or_inf"]
        self._use_dynamic_loss_scaling = sd["use_dynamic_loss_scaling"]


AmpScaler = GradScaler
