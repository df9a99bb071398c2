"""``paddle.nn.Layer`` on ``torch.nn.Module`` (reference: python/paddle/nn/layer/layers.py:332-2400).

Kept from the reference: structured state-dict keys ``prefix + sublayer + "." + param`` (layers.py:1889-1917),
``create_parameter`` with ParamAttr/initializer and unique parameter names ``<layer full name>.w_0 / .b_0``,
forward pre/post hooks with removable handles, ``parameters()`` returning a list, ``set_state_dict`` matching by
structured name, ``to``/``astype``, ``clear_gradients``.
"""
from __future__ import annotations

import re
from collections import OrderedDict

import numpy as np
import torch

from ...framework import dtype as _dt
from ...framework.param_attr import ParamAttr
from ...framework.place import current_torch_device, to_torch_device
from ...framework.tensor_patch import Parameter
from ...utils import unique_name
from .. import initializer as I


def _camel_to_snake(name):
    s1 = re.sub("(.)([A-Z][a-z]+)", r"\1_\2", name)
    return re.sub("([a-z])([A-Z])", r"\1_\2", s1).lower()


class HookRemoveHelper:
    def __init__(self, handle):
        self._h = handle

    def remove(self):
        self._h.remove()


class Layer(torch.nn.Module):
    def __init__(self, name_scope=None, dtype="float32"):
        super().__init__()
        object.__setattr__(self, "_full_name", unique_name.generate(name_scope or _camel_to_snake(type(self).__name__)))
        # layers follow paddle.set_default_dtype unless a dtype is given explicitly (LayerHelper.get_default_dtype)
        object.__setattr__(self, "_dtype", dtype if dtype != "float32" else _dt.get_default_dtype())
        object.__setattr__(self, "_param_counter", {"w": 0, "b": 0})
        object.__setattr__(self, "_non_persistable", set())

    # torch.nn.Module subclasses that forget super().__init__ would break; make attribute access forgiving
    def __getattr__(self, name):
        if name in ("_full_name", "_dtype", "_param_counter", "_non_persistable"):
            defaults = {"_full_name": _camel_to_snake(type(self).__name__), "_dtype": "float32",
                        "_param_counter": {"w": 0, "b": 0}, "_non_persistable": set()}
            object.__setattr__(self, name, defaults[name])
            return defaults[name]
        return super().__getattr__(name)

    # ------------------------------------------------------------- naming
    def full_name(self):
        return self._full_name

    # ------------------------------------------------------------- parameters
    def create_parameter(self, shape, attr=None, dtype=None, is_bias=False, default_initializer=None):
        attr = ParamAttr._to_attr(attr)
        if attr is False:
            return None
        td = _dt.to_torch_dtype(dtype or self._dtype or "float32")
        kind = "b" if is_bias else "w"
        n = self._param_counter[kind]
        self._param_counter[kind] = n + 1
        name = attr.name or f"{self._full_name}.{kind}_{n}"
        data = torch.empty([int(s) for s in shape], dtype=td, device=current_torch_device())
        p = Parameter(data, requires_grad=attr.trainable, name=name, trainable=attr.trainable,
                      optimize_attr={"learning_rate": attr.learning_rate}, regularizer=attr.regularizer,
                      need_clip=attr.need_clip, do_model_average=attr.do_model_average)
        init = attr.initializer or default_initializer
        if init is None:
            if is_bias:
                init = I._global_bias_init or I.Constant(0.0)
            else:
                init = I._global_weight_init or I.XavierUniform()
        if td in (torch.bfloat16, torch.float16) and not isinstance(init, (I.Constant, I.Assign)):
            tmp = torch.empty_like(p, dtype=torch.float32)
            init(tmp)
            with torch.no_grad():
                p.copy_(tmp)
        else:
            init(p)
        return p

    def create_tensor(self, name=None, persistable=None, dtype=None):
        t = torch.empty(0, dtype=_dt.to_torch_dtype(dtype or self._dtype), device=current_torch_device())
        if name:
            t.name = name
        return t

    create_variable = create_tensor

    def add_parameter(self, name, parameter):
        self.register_parameter(name, parameter)
        return parameter

    def add_sublayer(self, name, sublayer):
        self.add_module(str(name), sublayer)
        return sublayer

    def parameters(self, include_sublayers=True):
        return list(torch.nn.Module.parameters(self, recurse=include_sublayers))

    def named_parameters(self, prefix="", include_sublayers=True, remove_duplicate=True, recurse=None):
        if recurse is not None:   # torch-internal spelling
            include_sublayers = recurse
        return torch.nn.Module.named_parameters(self, prefix=prefix, recurse=include_sublayers,
                                                remove_duplicate=remove_duplicate)

    def sublayers(self, include_self=False):
        mods = list(torch.nn.Module.modules(self))
        return mods if include_self else mods[1:]

    def named_sublayers(self, prefix="", include_self=False, layers_set=None):
        for n, m in torch.nn.Module.named_modules(self, prefix=prefix):
            if m is self and not include_self:
                continue
            yield n, m

    def children(self):
        return torch.nn.Module.children(self)

    def named_children(self):
        return torch.nn.Module.named_children(self)

    def buffers(self, include_sublayers=True):
        return list(torch.nn.Module.buffers(self, recurse=include_sublayers))

    def named_buffers(self, prefix="", include_sublayers=True, recurse=None, remove_duplicate=True):
        if recurse is not None:
            include_sublayers = recurse
        return torch.nn.Module.named_buffers(self, prefix=prefix, recurse=include_sublayers,
                                             remove_duplicate=remove_duplicate)

    def register_buffer(self, name, tensor, persistable=True):
        torch.nn.Module.register_buffer(self, name, tensor, persistent=bool(persistable))

    # ------------------------------------------------------------- hooks
    def register_forward_pre_hook(self, hook, **kw):
        return HookRemoveHelper(torch.nn.Module.register_forward_pre_hook(self, hook, **kw))

    def register_forward_post_hook(self, hook, **kw):
        return HookRemoveHelper(torch.nn.Module.register_forward_hook(self, hook, **kw))

    # ------------------------------------------------------------- state dict
    def state_dict(self, destination=None, include_sublayers=True, structured_name_prefix="", use_hook=True,
                   keep_vars=True, prefix=None):
        if prefix is not None:  # torch-style call from torch internals
            structured_name_prefix = prefix
        dest = destination if destination is not None else OrderedDict()
        for n, p in self.named_parameters(prefix="", include_sublayers=include_sublayers):
            dest[structured_name_prefix + n] = p
        for n, b in self._named_persistent_buffers(include_sublayers):
            dest[structured_name_prefix + n] = b
        if use_hook:
            for hook in self.__dict__.get("_pd_state_dict_hooks", {}).values():
                hook(dest)
        return dest

    def _named_persistent_buffers(self, include_sublayers=True):
        mods = torch.nn.Module.named_modules(self) if include_sublayers else [("", self)]
        seen = set()
        for mp, m in mods:
            for bn, b in m._buffers.items():
                if b is None or bn in m._non_persistent_buffers_set or id(b) in seen:
                    continue
                seen.add(id(b))
                yield (mp + "." if mp else "") + bn, b

    def to_static_state_dict(self, *a, **k):
        return self.state_dict(*a, **k)

    def register_state_dict_hook(self, hook):
        """``hook(state_dict)`` runs on (and may edit) every dict ``state_dict()`` returns; the returned helper's
        ``remove()`` unregisters it."""
        hooks = self.__dict__.setdefault("_pd_state_dict_hooks", OrderedDict())
        key = len(hooks)
        while key in hooks:
            key += 1
        hooks[key] = hook

        class _Remove:
            def remove(self_inner):
                hooks.pop(key, None)
        return _Remove()

    def backward(self, *inputs):
        raise ValueError("Layer shouldn't implement backward")

    def set_state_dict(self, state_dict, use_structured_name=True):
        own = self.state_dict()
        if not use_structured_name:
            by_name = {v.name: v for v in own.values()}
        missing, unexpected = [], []
        matched = set()
        for k, v in state_dict.items():
            if k in ("StructuredToParameterName@@", "UnpackBigParamInfor@@"):
                continue
            tgt = own.get(k) if use_structured_name else by_name.get(k)
            if tgt is None:
                unexpected.append(k)
                continue
            matched.add(k)
            if isinstance(v, np.ndarray) or not isinstance(v, torch.Tensor):
                tgt.set_value(np.asarray(v))
            else:
                if tuple(v.shape) != tuple(tgt.shape):
                    raise ValueError(f"{k} receives a shape {tuple(v.shape)}, but the expected shape is {tuple(tgt.shape)}.")
                with torch.no_grad():
                    tgt.copy_(v.to(device=tgt.device, dtype=tgt.dtype))
        for k in own:
            if k not in matched:
                missing.append(k)
        return missing, unexpected

    set_dict = set_state_dict
    load_dict = set_state_dict

    def load_state_dict(self, state_dict, strict=True, **kw):  # torch spelling
        return self.set_state_dict(state_dict)

    # ------------------------------------------------------------- misc
    def clear_gradients(self, set_to_zero=False):
        for p in self.parameters():
            p.clear_grad(set_to_zero)

    def train(self, mode=True):
        torch.nn.Module.train(self, mode)
        return self

    def eval(self):
        torch.nn.Module.eval(self)
        return self

    def to(self, device=None, dtype=None, blocking=None):
        kw = {}
        if device is not None:
            kw["device"] = to_torch_device(device)
        if dtype is not None:
            kw["dtype"] = _dt.to_torch_dtype(dtype)
            object.__setattr__(self, "_dtype", _dt.convert_dtype(dtype))
        if kw.get("dtype") is not None:
            # cast floating parameters only (torch semantics already)
            pass
        self._apply_keep_param_attrs(lambda t: t.to(**{k: v for k, v in kw.items()
                                                          if not (k == "dtype" and not t.is_floating_point())}))
        return self

    def _apply_keep_param_attrs(self, fn):
        """Move/cast in place so Parameter objects (and the attributes hung on them) survive."""
        for m in torch.nn.Module.modules(self):
            for k, p in m._parameters.items():
                if p is None:
                    continue
                with torch.no_grad():
                    new = fn(p.data)
                p.data = new
                if p.grad is not None:
                    p.grad.data = fn(p.grad.data)
            for k, b in m._buffers.items():
                if b is not None:
                    m._buffers[k] = fn(b)

    def astype(self, dtype=None):
        return self.to(dtype=dtype)

    def float(self, excluded_layers=None):
        return self.to(dtype="float32")

    def float16(self, excluded_layers=None):
        return self.to(dtype="float16")

    def bfloat16(self, excluded_layers=None):
        return self.to(dtype="bfloat16")

    def cuda(self, device=None):
        return self.to(device=torch.device("cuda", device if device is not None else torch.cuda.current_device()))

    def cpu(self):
        return self.to(device="cpu")


# ----------------------------------------------------------------- containers
class LayerList(Layer):
    def __init__(self, sublayers=None):
        super().__init__()
        if sublayers is not None:
            for i, l in enumerate(sublayers):
                self.add_sublayer(str(i), l)

    def _idx(self, i):
        n = len(self)
        if i < 0:
            i += n
        return str(i)

    def __getitem__(self, idx):
        if isinstance(idx, slice):
            return LayerList(list(self._modules.values())[idx])
        return self._modules[self._idx(idx)]

    def __setitem__(self, idx, layer):
        self._modules[self._idx(idx)] = layer

    def __delitem__(self, idx):
        del self._modules[self._idx(idx)]
        vals = list(self._modules.values())
        self._modules.clear()
        for i, l in enumerate(vals):
            self._modules[str(i)] = l

    def __len__(self):
        return len(self._modules)

    def __iter__(self):
        return iter(self._modules.values())

    def append(self, sublayer):
        self.add_sublayer(str(len(self)), sublayer)
        return self

    def insert(self, index, sublayer):
        vals = list(self._modules.values())
        vals.insert(index, sublayer)
        self._modules.clear()
        for i, l in enumerate(vals):
            self._modules[str(i)] = l

    def extend(self, sublayers):
        for l in sublayers:
            self.append(l)
        return self


class Sequential(Layer):
    def __init__(self, *layers):
        super().__init__()
        if len(layers) > 0 and isinstance(layers[0], (list, tuple)) and len(layers[0]) == 2 and isinstance(layers[0][0], str):
            for name, l in layers:
                self.add_sublayer(name, l)
        else:
            for i, l in enumerate(layers):
                self.add_sublayer(str(i), l)

    def __getitem__(self, name):
        if isinstance(name, slice):
            return Sequential(*list(self._modules.values())[name])
        if isinstance(name, str):
            return self._modules[name]
        vals = list(self._modules.values())
        return vals[name]

    def __len__(self):
        return len(self._modules)

    def __iter__(self):
        return iter(self._modules.values())

    def forward(self, input):
        for l in self._modules.values():
            input = l(input)
        return input


class ParameterList(Layer):
    def __init__(self, parameters=None):
        super().__init__()
        if parameters is not None:
            for i, p in enumerate(parameters):
                self.add_parameter(str(i), p)

    def __getitem__(self, idx):
        return self._parameters[str(idx if idx >= 0 else len(self) + idx)]

    def __setitem__(self, idx, p):
        self._parameters[str(idx)] = p

    def __len__(self):
        return len(self._parameters)

    def __iter__(self):
        return iter(self._parameters.values())

    def append(self, parameter):
        self.add_parameter(str(len(self)), parameter)
        return self


class LayerDict(Layer):
    def __init__(self, sublayers=None):
        super().__init__()
        if sublayers:
            self.update(sublayers)

    def __getitem__(self, key):
        return self._modules[key]

    def __setitem__(self, key, layer):
        self.add_sublayer(key, layer)

    def __delitem__(self, key):
        del self._modules[key]

    def __len__(self):
        return len(self._modules)

    def __iter__(self):
        return iter(self._modules)

    def __contains__(self, key):
        return key in self._modules

    def keys(self):
        return self._modules.keys()

    def values(self):
        return self._modules.values()

    def items(self):
        return self._modules.items()

    def pop(self, key):
        v = self._modules[key]
        del self._modules[key]
        return v

    def clear(self):
        self._modules.clear()

    def update(self, sublayers):
        items = sublayers.items() if hasattr(sublayers, "items") else sublayers
        for k, v in items:
            self.add_sublayer(k, v)
