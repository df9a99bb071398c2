"""paddle.jit for dygraph code (reference: python/paddle/jit/api.py ``to_static`` :171-330, ``save`` :742-1200,
``load`` :1300-1500, translated_layer.py TranslatedLayer).

There is no tracing compiler on the training path by design: hot loops are hand-written kernels.  What a "static" function
buys on this hardware is the removal of launch overhead, and the tool for that is a CUDA graph, so

* ``to_static(fn_or_layer)`` returns a ``StaticFunction``.  Under autograd (training) and on CPU it simply runs the
  function.  Under ``no_grad`` with CUDA inputs it captures one CUDA graph per input signature (shapes, dtypes) with
  static input buffers and replays it afterwards - one launch per call.
* ``save(layer, path, input_spec)`` writes the parameters (``<path>.pdiparams``), a meta file (``<path>.pdmodel.json``)
  and - when ``input_spec`` (or a previous call) tells the input shapes - the program itself as a TorchScript trace
  (``<path>.pdmodel``), traced through the library ops (this framework's custom kernels are switched off while tracing, a
  trace cannot see through a ctypes launch), so the file loads and runs on any device.
* ``load(path)`` returns a ``TranslatedLayer``: callable when a program was saved, always a holder of the parameters.
"""
from __future__ import annotations

import inspect
import json
import os

import torch

from ..framework.io import load as _load
from ..framework.io import save as _save

_enabled = [True]


def _signature(args):
    sig = []
    for a in args:
        if isinstance(a, torch.Tensor):
            sig.append((tuple(a.shape), a.dtype, a.device.type))
        else:
            sig.append(("py", repr(a)))
    return tuple(sig)


class StaticFunction:
    """Callable returned by ``to_static`` (reference: dy2static/program_translator.py StaticFunction)."""

    def __init__(self, function, input_spec=None, layer=None, full_graph=True, backend=None):
        self._fn = function
        self._layer = layer
        self._input_spec = list(input_spec) if input_spec is not None else None
        self._graphs = {}
        self._last_args = None
        self.__name__ = getattr(function, "__name__", type(function).__name__)
        self.__doc__ = getattr(function, "__doc__", None)

    # the reference exposes these on the wrapper
    @property
    def dygraph_function(self):
        return self._fn

    @property
    def code(self):
        try:
            return inspect.getsource(self._fn)
        except (OSError, TypeError):
            return ""

    @property
    def input_spec(self):
        return self._input_spec

    def rollback(self):
        """Give the eager function back (and, for a layer's forward, put it back on the layer)."""
        if self._layer is not None:
            self._layer.forward = self._fn
        return self._fn

    def concrete_program_specify_input_spec(self, input_spec=None, **kw):
        self._input_spec = list(input_spec) if input_spec is not None else self._input_spec
        return self

    def _graph_ok(self, args, kwargs):
        if not _enabled[0] or kwargs or not torch.cuda.is_available() or torch.is_grad_enabled():
            return False
        tensors = [a for a in args if isinstance(a, torch.Tensor)]
        return bool(tensors) and all(t.is_cuda for t in tensors) and all(
            isinstance(a, (torch.Tensor, int, float, bool, str, type(None))) for a in args)

    def __call__(self, *args, **kwargs):
        self._last_args = args
        if not self._graph_ok(args, kwargs):
            return self._fn(*args, **kwargs)
        key = _signature(args)
        st = self._graphs.get(key)
        if st is None:
            from ..device.cuda.graphs import CUDAGraph
            static_in = [a.clone() if isinstance(a, torch.Tensor) else a for a in args]
            for _ in range(2):                  # lazy initialisation (tensor maps, descriptor tables) outside capture
                self._fn(*static_in)
            torch.cuda.synchronize()
            g = CUDAGraph()
            g.capture_begin()
            out = self._fn(*static_in)
            g.capture_end()
            st = self._graphs[key] = (g, static_in, out)
        g, static_in, out = st
        for dst, src in zip(static_in, args):
            if isinstance(dst, torch.Tensor):
                dst.copy_(src)
        g.replay()
        return out

    def __get__(self, instance, owner):          # decorated methods
        if instance is None:
            return self
        bound = StaticFunction(self._fn.__get__(instance, owner), self._input_spec, None)
        return bound


def to_static(function=None, input_spec=None, build_strategy=None, backend=None, full_graph=True, **kwargs):
    """Decorator / converter (reference: api.py:171).  A ``Layer`` argument keeps being that layer, with its ``forward``
    replaced by a ``StaticFunction``."""
    def convert(obj):
        if isinstance(obj, torch.nn.Module):
            if not isinstance(obj.forward, StaticFunction):
                obj.forward = StaticFunction(obj.forward, input_spec, layer=obj, full_graph=full_graph, backend=backend)
            elif input_spec is not None:
                obj.forward._input_spec = list(input_spec)
            return obj
        if isinstance(obj, StaticFunction):
            return obj
        return StaticFunction(obj, input_spec, full_graph=full_graph, backend=backend)
    return convert(function) if function is not None else convert


def not_to_static(func=None):
    return func if func is not None else (lambda f: f)


def ignore_module(modules):
    return None


_verbosity = {"level": 0, "code_level": 0, "also_to_stdout": False}


def set_verbosity(level=0, also_to_stdout=False):
    """Logging level of the translator.  Nothing is transformed here (a StaticFunction replays a CUDA graph), so the
    level only gates the messages ``StaticFunction`` prints about capture / fallback."""
    _verbosity["level"], _verbosity["also_to_stdout"] = int(level), bool(also_to_stdout)


def set_code_level(level=100, also_to_stdout=False):
    """The reference prints the transformed code of AST pass ``level``; the source is untransformed here, so any level
    > 0 makes ``StaticFunction.code`` the thing to look at.  Kept for scripts that call it."""
    _verbosity["code_level"], _verbosity["also_to_stdout"] = int(level), bool(also_to_stdout)


def enable_to_static(enable_to_static_bool):
    """Global switch: off -> StaticFunctions run eagerly (no graph capture)."""
    _enabled[0] = bool(enable_to_static_bool)


# --------------------------------------------------------------------------------------------------------- save / load
def _example_inputs(specs, device):
    """Concrete tensors for tracing: unknown dims (None / -1) become 2, so that the trace does not specialise on 1."""
    out = []
    for s in specs:
        if isinstance(s, torch.Tensor):
            out.append(s.detach().to(device))
            continue
        shape = [2 if (d is None or (isinstance(d, int) and d < 0)) else int(d) for d in s.shape]
        dt = s.dtype if isinstance(s.dtype, torch.dtype) else getattr(torch, str(s.dtype).replace("paddle.", "").replace("torch.", ""), torch.float32)
        if dt.is_floating_point:
            out.append(torch.randn(shape, device=device).to(dt))
        elif dt == torch.bool:
            out.append(torch.zeros(shape, dtype=dt, device=device))
        else:
            out.append(torch.zeros(shape, dtype=dt, device=device))
    return out


def save(layer, path, input_spec=None, **configs):
    """``paddle.jit.save`` (reference: api.py:742).  Always writes the parameters and the meta file; writes the program
    when the input shapes are known (``input_spec``, the spec given to ``to_static``, or the arguments of the last call
    through a ``StaticFunction``)."""
    from ..framework.flags import get_flags, set_flags
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    _save(layer.state_dict(), path + ".pdiparams")
    specs = input_spec
    fwd = getattr(layer, "forward", None)
    if specs is None and isinstance(fwd, StaticFunction):
        specs = fwd.input_spec or ([a for a in fwd._last_args if isinstance(a, torch.Tensor)] if fwd._last_args else None)
    meta_specs = [dict(shape=[None if d is None else int(d) for d in getattr(s, "shape", ())], dtype=str(getattr(s, "dtype", "")),
                       name=getattr(s, "name", None)) for s in (specs or [])]
    has_program = False
    num_outputs = None
    if specs:
        params = [p for p in layer.parameters()]
        device = params[0].device if params else torch.device("cpu")
        examples = _example_inputs(specs, device)
        was_training = layer.training
        old = get_flags(["FLAGS_use_native_kernels", "FLAGS_use_tcgen05_gemm", "FLAGS_use_fused_comm"])
        set_flags({"FLAGS_use_native_kernels": False, "FLAGS_use_tcgen05_gemm": False, "FLAGS_use_fused_comm": False})
        saved_forward = layer.__dict__.get("forward")
        try:
            layer.eval()
            if isinstance(saved_forward, StaticFunction):
                layer.forward = saved_forward.dygraph_function
            with torch.no_grad():
                traced = torch.jit.trace(layer, tuple(examples), check_trace=False, strict=False)
            torch.jit.save(traced, path + ".pdmodel")
            has_program = True
            with torch.no_grad():
                sample_out = traced(*examples)
            num_outputs = len(sample_out) if isinstance(sample_out, (list, tuple)) else 1
        finally:
            if saved_forward is not None:
                layer.forward = saved_forward
            set_flags(old)
            layer.train(was_training)
    with open(path + ".pdmodel.json", "w") as f:
        json.dump({"class": type(layer).__name__, "input_spec": meta_specs, "program": has_program, "num_outputs": num_outputs}, f)


class TranslatedLayer(torch.nn.Module):
    """What ``paddle.jit.load`` returns (reference: translated_layer.py:1280): runs the saved program, exposes its
    parameters; ``program()`` gives the underlying TorchScript module."""

    def __init__(self, state, specs, module=None):
        super().__init__()
        self._state, self.input_spec = state, specs
        self._module = module

    def state_dict(self, *a, **k):
        return self._state

    def set_state_dict(self, sd, *a, **k):
        self._state = dict(sd)
        if self._module is not None:
            own = dict(self._module.named_parameters())
            own.update(dict(self._module.named_buffers()))
            with torch.no_grad():
                for k_, v in sd.items():
                    if k_ in own:
                        own[k_].copy_(torch.as_tensor(v))

    def parameters(self, *a, **k):
        return list(self._module.parameters()) if self._module is not None else []

    def program(self, method_name="forward"):
        return self._module

    def forward(self, *args, **kwargs):
        if self._module is None:
            raise RuntimeError("this model was saved without a program (no input_spec was known at paddle.jit.save): "
                               "build the Layer and call set_state_dict(loaded.state_dict())")
        return self._module(*args, **kwargs)


def load(path, **configs):
    with open(path + ".pdmodel.json") as f:
        meta = json.load(f)
    module = None
    if meta.get("program") and os.path.exists(path + ".pdmodel"):
        module = torch.jit.load(path + ".pdmodel", map_location=configs.get("map_location"))
    layer = TranslatedLayer(_load(path + ".pdiparams"), meta.get("input_spec"), module)
    layer.num_outputs = meta.get("num_outputs")
    layer.eval()
    return layer
