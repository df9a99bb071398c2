"""paddle.quantization (reference: python/paddle/quantization/{config,factory,base_quanter,base_observer,quantize,qat,
ptq,wrapper}.py, quanters/abs_max.py, observers/{abs_max,groupwise}.py, nn/quant/qat/*, nn/quant/format.py).

The moving parts:
  * a *quanter* is a Layer that simulates quantisation in the forward pass (fake quant, straight-through gradient) and
    reports ``scales() / zero_points() / quant_axis() / bit_length()``; an *observer* only watches values (PTQ) and
    produces scales from what it saw;
  * ``QuanterFactory`` (made by the ``@quanter`` decorator) stores constructor arguments so that one configuration
    object can stamp out a fresh quanter per layer;
  * ``QuantConfig`` resolves, per layer, which factories apply (by layer object > by full name > by type > global) and
    which quantised layer class replaces which float layer;
  * ``QAT.quantize`` / ``PTQ.quantize`` swap ``Linear`` / ``Conv2D`` for ``QuantedLinear`` / ``QuantedConv2D`` and wrap
    other configured leaves in ``ObserveWrapper``; ``convert`` freezes everything into explicit quantise -> dequantise
    pairs (``LinearQuanterDequanter``) around weights stored on the integer grid.
"""
from __future__ import annotations

import abc
import copy

import torch

from ..nn.layer.layers import Layer
from ..nn.quant import QuanterStub, Stub

__all__ = ["QuantConfig", "BaseQuanter", "BaseObserver", "quanter", "QAT", "PTQ"]


# ------------------------------------------------------------------------------------------------------------ factory
class ClassWithArguments(metaclass=abc.ABCMeta):
    def __init__(self, *args, **kwargs):
        self._args, self._kwargs = args, kwargs

    @property
    def args(self):
        return self._args

    @property
    def kwargs(self):
        return self._kwargs

    @abc.abstractmethod
    def _get_class(self):
        pass

    def __str__(self):
        args = ",".join([str(a) for a in self.args] + [f"{k}={v}" for k, v in self.kwargs.items()])
        return f"{self.__class__.__name__}({args})"

    __repr__ = __str__


class QuanterFactory(ClassWithArguments):
    """Holds a quanter class and its arguments; ``_instance(layer)`` builds the quanter for one layer."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.partial_class = None

    def _instance(self, layer):
        return self._get_class()(layer, *self.args, **self.kwargs)


ObserverFactory = QuanterFactory


def quanter(class_name):
    """Class decorator: ``@quanter("MyQuanter") class MyQuanterLayer(BaseQuanter)`` declares, in the layer's module, a
    factory class ``MyQuanter(*args, **kwargs)`` whose instances create ``MyQuanterLayer(layer, *args, **kwargs)``."""
    import sys

    def wrapper(target_class):
        def _get_class(self):
            return target_class
        factory = type(class_name, (QuanterFactory,), {"_get_class": _get_class, "__doc__": target_class.__doc__})
        module = sys.modules[target_class.__module__]
        setattr(module, class_name, factory)
        if hasattr(module, "__all__") and class_name not in module.__all__:
            module.__all__.append(class_name)
        return target_class
    return wrapper


# ------------------------------------------------------------------------------------------------------- base classes
class BaseQuanter(Layer, metaclass=abc.ABCMeta):
    def __init__(self):
        super().__init__()

    @abc.abstractmethod
    def forward(self, input):
        pass

    @abc.abstractmethod
    def scales(self):
        pass

    @abc.abstractmethod
    def zero_points(self):
        pass

    @abc.abstractmethod
    def quant_axis(self):
        pass

    @abc.abstractmethod
    def bit_length(self):
        pass


class BaseObserver(BaseQuanter, metaclass=abc.ABCMeta):
    def __init__(self):
        super().__init__()

    @abc.abstractmethod
    def cal_thresholds(self):
        pass


class _FakeQuantSTE(torch.autograd.Function):
    """round(clip(x, -s, s) / s * range) * s / range, gradient passed straight through."""

    @staticmethod
    def forward(ctx, x, scale, qrange):
        s = scale.clamp_min(1e-30)
        return torch.round(torch.maximum(torch.minimum(x, s), -s) / s * qrange) * s / qrange

    @staticmethod
    def backward(ctx, g):
        return g, None, None


def fake_quant_dequant(x, scale, bit_length=8):
    return _FakeQuantSTE.apply(x, scale.to(x.dtype), float(2 ** (bit_length - 1) - 1))


# ----------------------------------------------------------------------------------------------------------- quanters
class FakeQuanterWithAbsMaxObserverLayer(BaseQuanter):
    """QAT quanter with a moving-average abs-max scale: while training
    ``state = rate * state + 1; accum = rate * accum + max|x|; scale = accum / state``; the output is the fake-quantised
    input with that scale.  In eval mode the stored scale is used unchanged."""

    def __init__(self, layer, name=None, moving_rate=0.9, bit_length=8, dtype="float32"):
        super().__init__()
        self._moving_rate, self._bit_length = moving_rate, bit_length
        from ..framework import dtype as _dt
        td = _dt.to_torch_dtype(dtype)
        self.register_buffer("_scale", torch.full([1], 0.001, dtype=td))
        self.register_buffer("_state", torch.ones([1], dtype=td))
        self.register_buffer("_accum", torch.ones([1], dtype=td))

    def forward(self, input):
        if self.training:
            with torch.no_grad():
                self._state.mul_(self._moving_rate).add_(1.0)
                self._accum.mul_(self._moving_rate).add_(input.detach().abs().max().to(self._accum.dtype))
                self._scale.copy_(self._accum / self._state)
        return fake_quant_dequant(input, self._scale, self._bit_length)

    def bit_length(self):
        return self._bit_length

    def quant_axis(self):
        return -1

    def scales(self):
        return self._scale

    def zero_points(self):
        return None


quanter("FakeQuanterWithAbsMaxObserver")(FakeQuanterWithAbsMaxObserverLayer)


# ---------------------------------------------------------------------------------------------------------- observers
class AbsmaxObserverLayer(BaseObserver):
    """PTQ observer: remembers the largest |x| seen; the values pass through untouched."""

    def __init__(self, layer, quant_bits=8):
        super().__init__()
        self._quant_bits = quant_bits
        self._max = 1e-7
        self._scale = None

    def forward(self, inputs):
        self._max = max(self._max, float(inputs.detach().abs().max()))
        return inputs

    def cal_thresholds(self):
        self._scale = torch.tensor([self._max], dtype=torch.float32)

    def bit_length(self):
        return self._quant_bits

    def quant_axis(self):
        return -1

    def scales(self):
        if self._scale is None:
            self.cal_thresholds()
        return self._scale

    def zero_points(self):
        return None


quanter("AbsmaxObserver")(AbsmaxObserverLayer)


class GroupWiseWeightObserverLayer(BaseObserver):
    """Weight observer with one abs-max per (group of ``group_size`` input rows, output column) of a [k, n] weight."""

    def __init__(self, layer, quant_bits=8, group_size=128):
        super().__init__()
        self._quant_bits, self.group_size = quant_bits, group_size
        self._scale = None

    def forward(self, inputs):
        k, n = inputs.shape
        if k % self.group_size:
            raise ValueError(f"group_size {self.group_size} must be divisible by input_channel {k}")
        self._scale = inputs.detach().abs().reshape(k // self.group_size, self.group_size, n).amax(1)
        return inputs

    def cal_thresholds(self):
        pass

    def bit_length(self):
        return self._quant_bits

    def quant_axis(self):
        return -1

    def scales(self):
        return self._scale

    def zero_points(self):
        return None


quanter("GroupWiseWeightObserver")(GroupWiseWeightObserverLayer)


# ------------------------------------------------------------------------------------------------------ configuration
class SingleLayerConfig:
    def __init__(self, activation, weight):
        self._activation, self._weight = activation, weight

    @property
    def activation(self):
        return self._activation

    @property
    def weight(self):
        return self._weight

    def __str__(self):
        return f"activation: {self._activation}\nweight: {self._weight}"


class ObserveWrapper(Layer):
    """Observe the INPUT (default) or the output of a leaf layer that has no quantised counterpart."""

    def __init__(self, observer, observed, observe_input=True):
        super().__init__()
        self._observer, self._observed, self._observe_input = observer, observed, observe_input

    def forward(self, *inputs, **kwargs):
        if self._observe_input:
            return self._observed(self._observer(*inputs, **kwargs))
        return self._observer(self._observed(*inputs, **kwargs))


class ConvertibleQuantedLayer(Layer, metaclass=abc.ABCMeta):
    """A quantised layer that ``convert`` can freeze: ``weights_to_quanters()`` lists (weight attribute, quanter
    attribute) pairs, ``activation_quanters()`` the activation quanter attributes."""

    def __init__(self):
        super().__init__()
        self.converted = False

    @abc.abstractmethod
    def weights_to_quanters(self):
        pass

    @abc.abstractmethod
    def activation_quanters(self):
        pass

    def _convert(self, remain_weight=False):
        assert not self.converted, "The model should be converted only once."
        for weight_name, quanter_name in self.weights_to_quanters():
            q = getattr(self, quanter_name)
            if q is None:
                continue
            qdq = LinearQuanterDequanter.from_quanter(q)
            weight = getattr(self, weight_name)
            if not remain_weight:
                # store the weight on the integer grid (as floats): dequantising it reproduces the fake-quant value
                with torch.no_grad():
                    weight.copy_(qdq._quanter(weight))
                qdq = qdq._dequanter
            setattr(self, quanter_name, qdq)
        for quanter_name in self.activation_quanters():
            q = getattr(self, quanter_name)
            if q is not None:
                setattr(self, quanter_name, LinearQuanterDequanter.from_quanter(q))
        self.converted = True


class QuantedLinear(ConvertibleQuantedLayer):
    def __init__(self, layer, q_config):
        super().__init__()
        self.weight, self.bias, self.name = layer.weight, layer.bias, getattr(layer, "name", None)
        self.weight_quanter = q_config.weight._instance(layer) if q_config.weight is not None else None
        self.activation_quanter = q_config.activation._instance(layer) if q_config.activation is not None else None

    def forward(self, input):
        from ..nn import functional as F
        x = self.activation_quanter(input) if self.activation_quanter is not None else input
        w = self.weight_quanter(self.weight) if self.weight_quanter is not None else self.weight
        return F.linear(x, w, self.bias)

    def weights_to_quanters(self):
        return [("weight", "weight_quanter")]

    def activation_quanters(self):
        return ["activation_quanter"]


class QuantedConv2D(ConvertibleQuantedLayer):
    def __init__(self, layer, q_config):
        super().__init__()
        self._inner = [layer]                                   # configuration source, not a sub-layer
        self.weight, self.bias = layer.weight, layer.bias
        self.weight_quanter = q_config.weight._instance(layer) if q_config.weight is not None else None
        self.activation_quanter = q_config.activation._instance(layer) if q_config.activation is not None else None

    def forward(self, input):
        from ..nn import functional as F
        src = self._inner[0]
        x = self.activation_quanter(input) if self.activation_quanter is not None else input
        w = self.weight_quanter(self.weight) if self.weight_quanter is not None else self.weight
        padding = src._padding
        if src._padding_mode != "zeros":
            x = F.pad(x, src._explicit_pad, mode=src._padding_mode, data_format=src._data_format)
            padding = 0
        return F.conv2d(x, w, self.bias, stride=src._stride, padding=padding, dilation=src._dilation, groups=src._groups,
                        data_format=src._data_format)

    def weights_to_quanters(self):
        return [("weight", "weight_quanter")]

    def activation_quanters(self):
        return ["activation_quanter"]


class LinearQuanter(Layer):
    """x -> clip(round(x / scale * bound) + zero point, -bound, bound): the symmetric grid the fake quanters train on, so a
    converted model computes exactly what the quantised model computed in eval mode."""

    def __init__(self, scales, zero_point=None, quant_axis=None, bit_length=8):
        super().__init__()
        self._scales = torch.as_tensor(scales, dtype=torch.float32)
        self._zero_point = torch.zeros(1) if zero_point is None else torch.as_tensor(zero_point, dtype=torch.float32)
        self._quant_axis = -1 if quant_axis is None else quant_axis
        self._bit_length = bit_length

    def _shaped(self, t, x):
        if self._quant_axis in (-1, None) or t.numel() == 1:
            return t.to(x.device, x.dtype)
        shape = [1] * x.dim()
        shape[self._quant_axis] = -1
        return t.reshape(shape).to(x.device, x.dtype)

    def forward(self, input):
        bound = 2 ** (self._bit_length - 1) - 1
        s = self._shaped(self._scales, input).clamp_min(1e-30)
        return torch.clamp(torch.round(input / s * bound) + self._shaped(self._zero_point, input), -bound, bound)

    @staticmethod
    def from_quanter(q):
        return LinearQuanter(q.scales(), q.zero_points(), q.quant_axis(), q.bit_length())


class LinearDequanter(LinearQuanter):
    """q -> (q - zero point) * scale / bound."""

    def forward(self, input):
        bound = 2 ** (self._bit_length - 1) - 1
        return (input - self._shaped(self._zero_point, input)) * self._shaped(self._scales, input) / bound

    @staticmethod
    def from_quanter(q):
        return LinearDequanter(q.scales(), q.zero_points(), q.quant_axis(), q.bit_length())


class LinearQuanterDequanter(Layer):
    def __init__(self, quanter, dequanter):
        super().__init__()
        self._quanter, self._dequanter = quanter, dequanter

    def forward(self, input):
        return self._dequanter(self._quanter(input))

    @staticmethod
    def from_quanter(q):
        assert q is not None
        return LinearQuanterDequanter(LinearQuanter.from_quanter(q), LinearDequanter.from_quanter(q))


def _default_qat_mapping():
    from .. import nn
    return {Stub: QuanterStub, nn.Linear: QuantedLinear, nn.Conv2D: QuantedConv2D}


class QuantConfig:
    """Which quanters go where.  Priority: ``add_layer_config`` > ``add_name_config`` > ``add_type_config`` > the global
    (``activation``, ``weight``) pair given to the constructor."""

    def __init__(self, activation, weight):
        self._global_config = None if activation is None and weight is None else SingleLayerConfig(activation, weight)
        self._layer2config, self._prefix2config, self._type2config = {}, {}, {}
        self._model = None
        self._qat_layer_mapping = _default_qat_mapping()
        self._customized_qat_layer_mapping = {}
        self._customized_leaves = []

    def add_layer_config(self, layer, activation=None, weight=None):
        for one in (layer if isinstance(layer, list) else [layer]):
            self._layer2config[id(one)] = SingleLayerConfig(activation, weight)

    def add_name_config(self, layer_name, activation=None, weight=None):
        for one in (layer_name if isinstance(layer_name, list) else [layer_name]):
            self._prefix2config[one] = SingleLayerConfig(activation, weight)

    def add_type_config(self, layer_type, activation=None, weight=None):
        for one in (layer_type if isinstance(layer_type, list) else [layer_type]):
            self._type2config[one] = SingleLayerConfig(activation, weight)

    def add_qat_layer_mapping(self, source, target):
        assert isinstance(source, type) and issubclass(source, torch.nn.Module), "The source layer to be placed should be a subclass of paddle.nn.Layer"
        assert isinstance(target, type) and issubclass(target, torch.nn.Module), "The target layer should be a subclass of paddle.nn.qat.Layer"
        self._qat_layer_mapping[source] = target
        self._customized_qat_layer_mapping[source] = target

    def add_customized_leaf(self, layer_type):
        self._customized_leaves.append(layer_type)

    @property
    def customized_leaves(self):
        return self._customized_leaves

    @property
    def qat_layer_mappings(self):
        return self._qat_layer_mapping

    @property
    def default_qat_layer_mapping(self):
        return _default_qat_mapping()

    @property
    def global_config(self):
        return self._global_config

    # --- resolution
    def _get_config_by_layer(self, layer):
        return self._layer2config.get(id(layer))

    def _is_quantifiable(self, layer):
        return id(layer) in self._layer2config

    def _is_leaf(self, layer):
        return not list(layer.children()) or type(layer) in self._customized_leaves

    def _need_observe(self, layer):
        return self._is_leaf(layer) and self._has_observer_config(layer)

    def _has_observer_config(self, layer):
        cfg = self._get_config_by_layer(layer)
        return cfg is not None and cfg.activation is not None

    def _get_qat_layer(self, layer):
        return self._qat_layer_mapping[type(layer)](layer, self._get_config_by_layer(layer))

    def _get_observe_wrapper(self, layer):
        return ObserveWrapper(self._get_config_by_layer(layer).activation._instance(layer), layer)

    def _specify(self, model):
        """Resolve the configuration of every sub-layer of ``model`` (children inherit their parent's)."""
        self._model = model
        self._walk(model, "", self._global_config)

    def _walk(self, layer, prefix, inherited):
        for name, child in layer.named_children():
            full = f"{prefix}.{name}" if prefix else name
            cfg = inherited
            if type(child) in self._type2config:
                cfg = self._type2config[type(child)]
            if full in self._prefix2config:
                cfg = self._prefix2config[full]
            if id(child) in self._layer2config:
                cfg = self._layer2config[id(child)]
            if cfg is not None:
                self._layer2config[id(child)] = cfg
            self._walk(child, full, cfg)

    def details(self):
        lines = []

        def walk(layer, indent):
            for name, child in layer.named_children():
                cfg = self._layer2config.get(id(child))
                lines.append("  " * indent + f"({name}): {type(child).__name__}" + (f" [{cfg.activation}, {cfg.weight}]" if cfg else ""))
                walk(child, indent + 1)
        if self._model is not None:
            lines.append(type(self._model).__name__)
            walk(self._model, 1)
        return "\n".join(lines)

    def __str__(self):
        out = f"Global config:\n{self._global_config}\n" if self._global_config is not None else ""
        if self._type2config:
            out += f"Layer type config:\n{ {k.__name__: str(v) for k, v in self._type2config.items()} }\n"
        if self._prefix2config:
            out += f"Layer prefix config: \n{ {k: str(v) for k, v in self._prefix2config.items()} }\n"
        return out


# ------------------------------------------------------------------------------------------------------------ drivers
class Quantization(metaclass=abc.ABCMeta):
    def __init__(self, config):
        self._config = copy.deepcopy(config)

    @abc.abstractmethod
    def quantize(self, model, inplace=False):
        pass

    def convert(self, model, inplace=False, remain_weight=False):
        """Freeze a quantised model: quanters / observers become explicit quantise-dequantise pairs with the scales
        they ended up with, weights are stored on the integer grid unless ``remain_weight``."""
        _model = model if inplace else copy.deepcopy(model)
        replaced = {}
        for name, child in _model.named_children():
            if isinstance(child, ConvertibleQuantedLayer):
                if child.converted:
                    continue
                wq = getattr(child, "weight_quanter", None)
                if hasattr(child, "weight_quanter") and (wq is None or wq.scales() is None):
                    continue
                child._convert(remain_weight=remain_weight)
            elif isinstance(child, BaseQuanter):
                replaced[name] = LinearQuanterDequanter.from_quanter(child)
            else:
                self.convert(child, inplace=True, remain_weight=remain_weight)
        for key, value in replaced.items():
            setattr(_model, key, value)
        return _model

    def _convert_to_quant_layers(self, model, config):
        replaced = {}
        for name, child in model.named_children():
            if config._is_quantifiable(child) and type(child) in config.qat_layer_mappings:
                replaced[name] = config._get_qat_layer(child)
            else:
                self._convert_to_quant_layers(child, config)
        for key, value in replaced.items():
            setattr(model, key, value)

    def _insert_activation_observers(self, model, config):
        replaced = {}
        targets = set(config._qat_layer_mapping.values()) | set(config._customized_qat_layer_mapping.values())
        for name, child in model.named_children():
            if type(child) in targets:
                continue
            if config._need_observe(child):
                replaced[name] = config._get_observe_wrapper(child)
            else:
                self._insert_activation_observers(child, config)
        for key, value in replaced.items():
            setattr(model, key, value)

    def __str__(self):
        return self._config.details()

    __repr__ = __str__


class QAT(Quantization):
    """Quantisation-aware training: ``quantize`` puts fake quanters into a model that must be in training mode."""

    def quantize(self, model, inplace=False):
        assert model.training, "Quantization-Aware Training should work on training models. Please set training mode by model.train()."
        _model = model if inplace else copy.deepcopy(model)
        self._config._specify(_model)
        self._convert_to_quant_layers(_model, self._config)
        self._insert_activation_observers(_model, self._config)
        return _model


class PTQ(Quantization):
    """Post-training quantisation: ``quantize`` inserts observers into a model in eval mode; run calibration data through
    it, then ``convert``."""

    def quantize(self, model, inplace=False):
        assert not model.training, "Post-Training Quantization should not work on training models. Please set evaluation mode by model.eval()."
        _model = model if inplace else copy.deepcopy(model)
        self._config._specify(_model)
        self._convert_to_quant_layers(_model, self._config)
        self._insert_activation_observers(_model, self._config)
        return _model


from . import observers, quanters  # noqa: E402,F401
