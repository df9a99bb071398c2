"""paddle.static (reference: python/paddle/static/__init__.py).

The static-graph executor, Program / IR and their passes are out of product scope (the reference's
260 k-line graph stack is replaced by eager execution + CUDA streams / graphs), so ``Program`` / ``Executor`` /
``program_guard`` / ``data`` raise with that explanation.  Everything in the namespace that has a meaning without a
graph is implemented eagerly, with a ``Layer`` standing where the reference takes a ``Program``: places, ``name_scope``
/ ``device_guard``, ``gradients`` / ``append_backward`` (autograd), ``ExponentialMovingAverage``, ``Print`` /
``py_func``, ``accuracy`` / ``auc``, ``save`` / ``load`` / ``*_program_state``, ``save_inference_model`` /
``load_inference_model`` (over ``paddle.jit``), the (de)serialisation helpers and the strategy / scope holders."""
from __future__ import annotations

import torch

from ..framework import dtype as _dt


class InputSpec:
    def __init__(self, shape, dtype="float32", name=None, stop_gradient=False):
        self.shape = tuple(shape)
        self.dtype = _dt.convert_dtype(dtype) if hasattr(_dt, "convert_dtype") else dtype
        self.name = name
        self.stop_gradient = stop_gradient

    @classmethod
    def from_tensor(cls, tensor, name=None):
        return cls(tuple(tensor.shape), tensor.dtype, name or getattr(tensor, "name", None))

    @classmethod
    def from_numpy(cls, ndarray, name=None):
        return cls(tuple(ndarray.shape), torch.from_numpy(ndarray).dtype, name)

    def batch(self, batch_size):
        self.shape = (batch_size,) + self.shape
        return self

    def unbatch(self):
        self.shape = self.shape[1:]
        return self

    def __repr__(self):
        return f"InputSpec(shape={self.shape}, dtype={self.dtype}, name={self.name}, stop_gradient={self.stop_gradient})"


def _no_static(*a, **k):
    raise NotImplementedError("the static-graph executor is out of scope of this framework (dygraph only); "
                              "see paddle_b200/static/__init__.py")


Program = Executor = program_guard = default_main_program = default_startup_program = data = _no_static


# ------------------------------------------------------------------------------------------------------------- places
def cpu_places(device_count=None):
    import os
    from ..framework.place import CPUPlace
    n = int(device_count) if device_count is not None else int(os.environ.get("CPU_NUM", 1))
    return [CPUPlace() for _ in range(n)]


def cuda_places(device_ids=None):
    from ..framework.place import CUDAPlace
    if device_ids is None:
        import os
        sel = os.environ.get("FLAGS_selected_gpus")
        device_ids = [int(i) for i in sel.split(",")] if sel else list(range(torch.cuda.device_count()))
    elif isinstance(device_ids, int):
        device_ids = [device_ids]
    return [CUDAPlace(int(i)) for i in device_ids]


def xpu_places(device_ids=None):
    raise NotImplementedError("there is no XPU back end in this build")


import contextlib as _contextlib  # noqa: E402


@_contextlib.contextmanager
def name_scope(prefix=None):
    """Names are not used for anything here (there is no graph to print); kept so that model code runs unchanged."""
    yield


@_contextlib.contextmanager
def device_guard(device=None):
    """Run the block's tensor creations on ``device`` ("cpu", "gpu", "gpu:1")."""
    if device is None:
        yield
        return
    name = str(device).replace("gpu", "cuda")
    if name.split(":")[0] not in ("cpu", "cuda"):
        raise ValueError(f"The Attr(device) should be 'cpu', 'gpu' or 'gpu:x', but received {device}")
    with torch.device(name):
        yield


Variable = torch.Tensor


def create_parameter(shape, dtype, name=None, attr=None, is_bias=False, default_initializer=None):
    from .. import create_parameter as _cp
    return _cp(shape, dtype, name=name, attr=attr, is_bias=is_bias, default_initializer=default_initializer)


def create_global_var(shape, value, dtype, persistable=False, force_cpu=False, name=None):
    t = torch.full([int(s) for s in shape], value, dtype=_dt.to_torch_dtype(dtype), device="cpu" if force_cpu else None)
    if name:
        t.name = name
    t.persistable = persistable
    return t


# ----------------------------------------------------------------------------------------------------------- autograd
def gradients(targets, inputs, target_gradients=None, no_grad_set=None):
    """d(sum targets) / d(inputs), evaluated now (reference: base/backward.py gradients builds the backward graph)."""
    targets = list(targets) if isinstance(targets, (list, tuple)) else [targets]
    inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
    if target_gradients is not None:
        target_gradients = list(target_gradients) if isinstance(target_gradients, (list, tuple)) else [target_gradients]
    else:
        target_gradients = [torch.ones_like(t) for t in targets]
    return list(torch.autograd.grad(targets, inputs, target_gradients, allow_unused=True, retain_graph=True))


def append_backward(loss, parameter_list=None, no_grad_set=None, callbacks=None, checkpoints=None, distop_context=None):
    """Back-propagate ``loss`` now and return [(parameter, gradient)].  Without a graph to collect parameters from,
    ``parameter_list`` is required."""
    if parameter_list is None:
        raise ValueError("append_backward needs parameter_list here: there is no program to collect the parameters from")
    params = [p for p in parameter_list if isinstance(p, torch.Tensor) and p.requires_grad]
    grads = torch.autograd.grad(loss, params, allow_unused=True, retain_graph=True)
    return [(p, g) for p, g in zip(params, grads) if g is not None]


class ExponentialMovingAverage:
    """Shadow copies ``ema = decay * ema + (1 - decay) * param`` with the reference's bias correction
    (``ema / (1 - decay^t)``) applied when the averages are swapped in.  ``thres_steps``: a callable / number giving the
    global step, making the decay ``min(decay, (1 + t) / (10 + t))``.  ``parameters`` takes the place of the program the
    reference collects trainable parameters from."""

    def __init__(self, decay=0.999, thres_steps=None, name=None, parameters=None):
        if parameters is None:
            raise ValueError("ExponentialMovingAverage needs parameters=<iterable of Parameters> (no program to collect them from)")
        self._decay, self._thres_steps, self._name = decay, thres_steps, name or ""
        self._params = [p for p in parameters if p.requires_grad]
        self._ema = [torch.zeros_like(p, dtype=torch.float32) for p in self._params]
        self._decay_pow = 1.0
        self._backup = None

    def _current_decay(self):
        if self._thres_steps is None:
            return self._decay
        t = float(self._thres_steps() if callable(self._thres_steps) else self._thres_steps)
        return min(self._decay, (1.0 + t) / (10.0 + t))

    @torch.no_grad()
    def update(self):
        d = self._current_decay()
        self._decay_pow *= d
        for e, p in zip(self._ema, self._params):
            e.mul_(d).add_(p.detach().float(), alpha=1.0 - d)

    @_contextlib.contextmanager
    def apply(self, executor=None, need_restore=True):
        with torch.no_grad():
            self._backup = [p.detach().clone() for p in self._params]
            corr = 1.0 - self._decay_pow
            for e, p in zip(self._ema, self._params):
                p.copy_((e / corr if corr > 0 else e).to(p.dtype))
        try:
            yield
        finally:
            if need_restore:
                self.restore(executor)

    @torch.no_grad()
    def restore(self, executor=None):
        if self._backup is not None:
            for b, p in zip(self._backup, self._params):
                p.copy_(b)
            self._backup = None


# ------------------------------------------------------------------------------------------------------ debug helpers
def Print(input, first_n=-1, message=None, summarize=20, print_tensor_name=True, print_tensor_type=True,  # noqa: N802
          print_tensor_shape=True, print_tensor_layout=True, print_tensor_lod=True, print_phase="both"):
    """Print a tensor (at most ``first_n`` times per call site when positive) and pass it through."""
    counts = Print.__dict__.setdefault("_counts", {})
    key = (message, tuple(input.shape))
    counts[key] = counts.get(key, 0) + 1
    if first_n < 0 or counts[key] <= first_n:
        parts = [message] if message else []
        if print_tensor_name and getattr(input, "name", None):
            parts.append(f"name: {input.name}")
        if print_tensor_type:
            parts.append(f"dtype: {input.dtype}")
        if print_tensor_shape:
            parts.append(f"shape: {list(input.shape)}")
        flat = input.detach().reshape(-1)
        parts.append(f"data: {flat[:summarize].tolist() if summarize > 0 else flat.tolist()}")
        print("  ".join(str(p) for p in parts), flush=True)
    return input


def py_func(func, x, out=None, backward_func=None, skip_vars_in_backward_input=None):
    """Call a Python function on tensors.  With ``backward_func`` the call is an autograd node whose backward is
    ``backward_func(*inputs, *outputs, *output_grads)`` (minus ``skip_vars_in_backward_input``)."""
    xs = list(x) if isinstance(x, (list, tuple)) else [x]
    if backward_func is None:
        res = func(*xs)
    else:
        skip = {id(v) for v in (skip_vars_in_backward_input or [])} if not isinstance(skip_vars_in_backward_input, torch.Tensor) \
            else {id(skip_vars_in_backward_input)}

        class _Fn(torch.autograd.Function):
            @staticmethod
            def forward(ctx, *args):
                outs = func(*args)
                outs_l = list(outs) if isinstance(outs, (list, tuple)) else [outs]
                ctx.save_for_backward(*args, *outs_l)
                ctx.n_in = len(args)
                return outs if not isinstance(outs, list) else tuple(outs)

            @staticmethod
            def backward(ctx, *grads):
                saved = ctx.saved_tensors
                ins, outs = saved[:ctx.n_in], saved[ctx.n_in:]
                args = [t for t, orig in zip(ins, xs) if id(orig) not in skip] + list(outs) + list(grads)
                g = backward_func(*args)
                return tuple(g) if isinstance(g, (list, tuple)) else (g,)
        res = _Fn.apply(*xs)
    if out is not None:
        outs = list(out) if isinstance(out, (list, tuple)) else [out]
        ress = list(res) if isinstance(res, (list, tuple)) else [res]
        for o, r in zip(outs, ress):
            if isinstance(o, torch.Tensor) and o.shape == r.shape and not r.requires_grad:
                with torch.no_grad():
                    o.copy_(r)
    return res


def accuracy(input, label, k=1, correct=None, total=None):
    from ..metric import accuracy as _acc
    return _acc(input, label, k=k, correct=correct, total=total)


def auc(input, label, curve="ROC", num_thresholds=4095, topk=1, slide_steps=1, ins_tag_weight=None):
    """(auc, batch auc, [statistics]) of positive-class probabilities ``input[:, 1]`` - one batch, so both values agree."""
    from ..metric import Auc
    m = Auc(curve=curve, num_thresholds=num_thresholds)
    m.update(input.detach().cpu().numpy(), label.detach().cpu().numpy())
    value = torch.tensor(m.accumulate(), dtype=torch.float32)
    return value, value.clone(), [torch.as_tensor(m._pos), torch.as_tensor(m._neg)]


def ctr_metric_bundle(input, label, ins_tag_weight=None):
    """(squared error sum, absolute error sum, probability sum, q sum, positive count, instance count) of CTR predictions."""
    p, y = input.reshape(-1).float(), label.reshape(-1).float()
    return ((p - y).square().sum(), (p - y).abs().sum(), p.sum(), p.sum(), y.sum(), torch.tensor(float(p.numel())))


# --------------------------------------------------------------------------------------------------------- save / load
def _as_layer(program, what):
    if isinstance(program, torch.nn.Module):
        return program
    raise TypeError(f"{what}: there are no static Programs in this framework - pass the Layer whose state should be used")


def save(program, model_path, protocol=4, **configs):
    """``<model_path>.pdparams`` (parameters and buffers) and ``<model_path>.pdopt`` when ``optimizer=`` is given."""
    from ..framework.io import save as _save
    layer = _as_layer(program, "paddle.static.save")
    _save(layer.state_dict(), model_path + ".pdparams", protocol=protocol)
    opt = configs.get("optimizer")
    if opt is not None:
        _save(opt.state_dict(), model_path + ".pdopt", protocol=protocol)


def load(program, model_path, executor=None, var_list=None):
    import os
    from ..framework.io import load as _load
    layer = _as_layer(program, "paddle.static.load")
    path = model_path if model_path.endswith(".pdparams") else model_path + ".pdparams"
    if not os.path.exists(path):
        raise ValueError(f"{path} does not exist")
    layer.set_state_dict(_load(path))


def load_program_state(model_path, var_list=None):
    """{name: ndarray} of a saved model."""
    import os
    import numpy as np
    from ..framework.io import load as _load
    path = model_path if os.path.isfile(model_path) else model_path + ".pdparams"
    state = _load(path)
    out = {}
    for k, v in state.items():
        if isinstance(v, torch.Tensor):
            out[k] = (v.float() if v.dtype == torch.bfloat16 else v).cpu().numpy()
        elif isinstance(v, np.ndarray):
            out[k] = v
    return out


def set_program_state(program, state_dict):
    layer = _as_layer(program, "paddle.static.set_program_state")
    own = layer.state_dict()
    unused = [k for k in state_dict if k not in own]
    with torch.no_grad():
        for k, v in state_dict.items():
            if k in own:
                own[k].copy_(torch.as_tensor(v).to(own[k].dtype))
    if unused:
        import warnings
        warnings.warn(f"This list is not set, Because of Paramerter not found in program. There are: {' '.join(unused)}")


def save_inference_model(path_prefix, feed_vars, fetch_vars, executor=None, **kwargs):
    """``paddle.jit.save(program, path_prefix, input_spec=feed_vars)``: ``program=`` is the Layer to deploy, ``feed_vars``
    its InputSpecs / example tensors.  ``fetch_vars`` is not needed (the outputs are what ``forward`` returns)."""
    from .. import jit
    layer = _as_layer(kwargs.get("program"), "paddle.static.save_inference_model")
    feeds = list(feed_vars) if isinstance(feed_vars, (list, tuple)) else [feed_vars]
    jit.save(layer, path_prefix, input_spec=feeds)


def load_inference_model(path_prefix, executor=None, **kwargs):
    """[callable program, feed names, fetch names]."""
    from .. import jit
    layer = jit.load(path_prefix)
    feeds = [s.get("name") or f"x{i}" for i, s in enumerate(layer.input_spec or [])]
    fetches = [f"fetch_{i}" for i in range(layer.num_outputs or 1)]
    return [layer, feeds, fetches]


def serialize_program(feed_vars, fetch_vars, **kwargs):
    """Bytes of the traced program of ``program=<Layer>`` for the given feeds."""
    import os
    import tempfile
    layer = _as_layer(kwargs.get("program"), "paddle.static.serialize_program")
    with tempfile.TemporaryDirectory() as d:
        save_inference_model(os.path.join(d, "m"), feed_vars, fetch_vars, program=layer)
        with open(os.path.join(d, "m.pdmodel"), "rb") as f:
            return f.read()


def serialize_persistables(feed_vars, fetch_vars, executor=None, **kwargs):
    import io
    layer = _as_layer(kwargs.get("program"), "paddle.static.serialize_persistables")
    buf = io.BytesIO()
    torch.save({k: v.detach().cpu() for k, v in layer.state_dict().items()}, buf)
    return buf.getvalue()


def deserialize_program(data):
    import io
    return torch.jit.load(io.BytesIO(data))


def deserialize_persistables(program, data, executor=None):
    import io
    state = torch.load(io.BytesIO(data), weights_only=True)
    if isinstance(program, torch.nn.Module) and not isinstance(program, torch.jit.ScriptModule):
        program.set_state_dict(state) if hasattr(program, "set_state_dict") else program.load_state_dict(state)
    else:
        own = dict(program.named_parameters())
        own.update(dict(program.named_buffers()))
        with torch.no_grad():
            for k, v in state.items():
                if k in own:
                    own[k].copy_(v)
    return program


def save_to_file(path, content):
    if not isinstance(content, bytes):
        raise ValueError("'content' type should be bytes.")
    with open(path, "wb") as f:
        f.write(content)


def load_from_file(path):
    with open(path, "rb") as f:
        return f.read()


def normalize_program(program, feed_vars, fetch_vars, **kwargs):
    """Pruning a graph to the feed -> fetch slice has no counterpart: a Layer already is that slice."""
    return _as_layer(program, "paddle.static.normalize_program")


# ---------------------------------------------------------------------------------------------- scopes and strategies
class _ScopeVar:
    def __init__(self, name):
        self.name, self._value = name, None

    def get_tensor(self):
        return self._value

    def set(self, value, place=None):
        self._value = value if isinstance(value, torch.Tensor) else torch.as_tensor(value)


class Scope:
    """Name -> variable table (the reference's C++ Scope, without the hierarchy)."""

    def __init__(self):
        self._vars = {}

    def var(self, name):
        return self._vars.setdefault(name, _ScopeVar(name))

    def find_var(self, name):
        return self._vars.get(name)

    def local_var_names(self):
        return list(self._vars)

    def drop_kids(self):
        pass


_scope_stack = [Scope()]


def global_scope():
    return _scope_stack[-1]


@_contextlib.contextmanager
def scope_guard(scope):
    _scope_stack.append(scope)
    try:
        yield
    finally:
        _scope_stack.pop()


class BuildStrategy:
    """Graph-build options of the reference's parallel executor; an attribute holder here (nothing reads them)."""

    def __init__(self):
        self.enable_inplace = True
        self.fuse_all_reduce_ops = True
        self.fuse_elewise_add_act_ops = False
        self.fuse_bn_act_ops = False
        self.fuse_relu_depthwise_conv = False
        self.fuse_broadcast_ops = False
        self.fuse_all_optimizer_ops = False
        self.sync_batch_norm = False
        self.memory_optimize = None
        self.debug_graphviz_path = ""
        self.build_cinn_pass = False


class CompiledProgram:
    """Holds (program, build strategy); 'compilation' is the identity."""

    def __init__(self, program_or_graph, build_strategy=None):
        self._program, self._build_strategy = program_or_graph, build_strategy or BuildStrategy()


class WeightNormParamAttr:
    """ParamAttr asking for weight normalisation along ``dim`` (w = g * v / ||v||).  Apply it to a built layer with
    ``paddle.nn.utils.weight_norm(layer, dim=attr.dim)``; layers do not reparameterise themselves from the attr."""

    params_with_weight_norm = []

    def __init__(self, dim=None, name=None, initializer=None, learning_rate=1.0, regularizer=None, trainable=True,
                 do_model_average=False, need_clip=True):
        from ..framework.param_attr import ParamAttr
        self._attr = ParamAttr(name=name, initializer=initializer, learning_rate=learning_rate, regularizer=regularizer,
                               trainable=trainable, do_model_average=do_model_average, need_clip=need_clip)
        self.dim = dim

    def __getattr__(self, item):
        return getattr(self.__dict__["_attr"], item)


def _no_ipu(*a, **k):
    raise NotImplementedError("there is no IPU back end in this build")


ipu_shard_guard = IpuCompiledProgram = IpuStrategy = set_ipu_shard = _no_ipu
