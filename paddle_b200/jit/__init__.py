"""paddle.jit surface for dygraph code (reference: python/paddle/jit/api.py).

There is no tracing compiler here by design (hot loops are hand-written kernels, launch-bound loops are captured in
CUDA graphs), so ``to_static`` is the identity: decorated layers / functions keep running eagerly.  ``save`` / ``load``
persist the parameters (``<path>.pdiparams``) and the input specs (``<path>.pdmodel.json``) so that training scripts
which end with ``paddle.jit.save`` keep working; the loaded object is a state dict holder, not an inference program."""
from __future__ import annotations

import json

from ..framework.io import load as _load
from ..framework.io import save as _save


def to_static(function=None, input_spec=None, build_strategy=None, backend=None, **kwargs):
    def deco(fn):
        try:
            fn._input_spec = input_spec
        except Exception:  # noqa: BLE001
            pass
        return fn
    return deco(function) if function is not None else deco


def not_to_static(func=None):
    return func if func is not None else (lambda f: f)


def ignore_module(modules):
    return None


def enable_to_static(enable_to_static_bool):
    return None


class TranslatedLayer:
    def __init__(self, state, specs):
        self._state, self.input_spec = state, specs

    def state_dict(self):
        return self._state

    def __call__(self, *a, **k):
        raise NotImplementedError("paddle.jit.load returns the saved parameters only: build the Layer and call "
                                  "set_state_dict(loaded.state_dict())")


def save(layer, path, input_spec=None, **configs):
    _save(layer.state_dict(), path + ".pdiparams")
    specs = [dict(shape=list(getattr(s, "shape", ())), dtype=str(getattr(s, "dtype", "")), name=getattr(s, "name", None))
             for s in (input_spec or [])]
    with open(path + ".pdmodel.json", "w") as f:
        json.dump({"class": type(layer).__name__, "input_spec": specs}, f)


def load(path, **configs):
    with open(path + ".pdmodel.json") as f:
        meta = json.load(f)
    return TranslatedLayer(_load(path + ".pdiparams"), meta.get("input_spec"))
