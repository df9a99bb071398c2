"""paddle.inference (reference: python/paddle/inference/{__init__,wrapper}.py; C++ AnalysisPredictor behind it).

The serving path of this framework: a model saved with ``paddle.jit.save`` (parameters + traced program) is loaded into
a ``Predictor`` with the reference's handle-based API (``get_input_handle(name).copy_from_cpu(...)``, ``run()``,
``get_output_handle(name).copy_to_cpu()``) or the newer ``predictor.run([tensors])``.  On GPU the program is replayed
as one CUDA graph per input signature - the role the reference gives to its IR passes / TensorRT sub-graphs is launch
overhead removal here; precision is chosen by converting the saved weights (``convert_to_mixed_precision``).
"""
from __future__ import annotations

import enum
import os
import shutil

import numpy as np
import torch

__all__ = ["Config", "DataType", "PlaceType", "PrecisionType", "Tensor", "Predictor", "create_predictor", "get_version",
           "_get_phi_kernel_name", "get_trt_compile_version", "convert_to_mixed_precision", "get_trt_runtime_version",
           "get_num_bytes_of_data_type", "PredictorPool", "XpuConfig"]


class DataType(enum.IntEnum):
    FLOAT64 = 0
    FLOAT32 = 1
    INT64 = 2
    INT32 = 3
    UINT8 = 4
    INT8 = 5
    FLOAT16 = 6
    BOOL = 7
    BFLOAT16 = 8


_TORCH_OF = {DataType.FLOAT64: torch.float64, DataType.FLOAT32: torch.float32, DataType.INT64: torch.int64, DataType.INT32: torch.int32,
             DataType.UINT8: torch.uint8, DataType.INT8: torch.int8, DataType.FLOAT16: torch.float16, DataType.BOOL: torch.bool,
             DataType.BFLOAT16: torch.bfloat16}
_DATATYPE_OF = {v: k for k, v in _TORCH_OF.items()}


class PlaceType(enum.IntEnum):
    UNK = -1
    CPU = 0
    GPU = 1
    XPU = 2
    CUSTOM = 4


class PrecisionType(enum.IntEnum):
    Float32 = 0
    Int8 = 1
    Half = 2
    Bfloat16 = 3


def get_num_bytes_of_data_type(dtype):
    return torch.empty(0, dtype=_TORCH_OF[DataType(dtype)]).element_size()


def get_version():
    from .. import __version__ as v
    return f"version: {v}\nbackend: paddle_b200 (CUDA graphs over hand-written sm_100a kernels)"


def get_trt_compile_version():
    return (0, 0, 0)                       # TensorRT is not part of this build


def get_trt_runtime_version():
    return (0, 0, 0)


def _get_phi_kernel_name(op_name):
    return op_name


class XpuConfig:
    """Accepted for source compatibility; there is no XPU back end."""

    def __init__(self):
        self.device_id = 0


class Config:
    """``Config(path_prefix)``, ``Config(model_dir)`` or ``Config(prog_file, params_file)`` - the files written by
    ``paddle.jit.save(layer, path_prefix, input_spec)``: ``.pdmodel`` (program), ``.pdiparams``, ``.pdmodel.json``."""

    def __init__(self, *args):
        self._prefix = None
        self._use_gpu, self._device_id, self._memory_pool_mb = False, 0, 0
        self._precision = PrecisionType.Float32
        self._ir_optim, self._memory_optim, self._threads, self._glog = True, False, 1, True
        self._cuda_graph = True
        if len(args) == 1:
            self.set_model(args[0])
        elif len(args) == 2:
            self.set_model(args[0], args[1])
        elif args:
            raise TypeError("Config() takes a path prefix / model dir, or (prog_file, params_file)")

    # --- model location
    def set_model(self, first, params_file=None):
        if params_file is not None:
            self._prefix = first[:-len(".pdmodel")] if first.endswith(".pdmodel") else os.path.splitext(first)[0]
            return
        if os.path.isdir(first):
            found = sorted(f for f in os.listdir(first) if f.endswith(".pdmodel"))
            if not found:
                raise ValueError(f"no .pdmodel file under {first}")
            self._prefix = os.path.join(first, found[0][:-len(".pdmodel")])
        else:
            self._prefix = first[:-len(".pdmodel")] if first.endswith(".pdmodel") else first

    def set_prog_file(self, path):
        self.set_model(path, self.params_file())

    def set_params_file(self, path):
        pass                                # the parameters travel with the prefix

    def model_dir(self):
        return os.path.dirname(self._prefix) if self._prefix else ""

    def prog_file(self):
        return self._prefix + ".pdmodel" if self._prefix else ""

    def params_file(self):
        return self._prefix + ".pdiparams" if self._prefix else ""

    # --- device / precision
    def enable_use_gpu(self, memory_pool_init_size_mb=100, device_id=0, precision_mode=PrecisionType.Float32):
        self._use_gpu, self._memory_pool_mb, self._device_id = True, memory_pool_init_size_mb, device_id
        self._precision = PrecisionType(precision_mode)

    def disable_gpu(self):
        self._use_gpu = False

    def use_gpu(self):
        return self._use_gpu

    def gpu_device_id(self):
        return self._device_id

    def memory_pool_init_size_mb(self):
        return self._memory_pool_mb

    def enable_tensorrt_engine(self, *args, **kwargs):
        raise NotImplementedError("TensorRT is not part of this build: on B200 the predictor replays CUDA graphs over the "
                                  "framework's own kernels (Config.enable_cuda_graph, on by default)")

    def tensorrt_engine_enabled(self):
        return False

    def enable_cuda_graph(self, flag=True):
        self._cuda_graph = bool(flag)

    def cuda_graph_enabled(self):
        return self._cuda_graph

    # --- knobs kept for source compatibility
    def switch_ir_optim(self, flag=True):
        self._ir_optim = bool(flag)

    def ir_optim(self):
        return self._ir_optim

    def enable_memory_optim(self, flag=True):
        self._memory_optim = bool(flag)

    def set_cpu_math_library_num_threads(self, n):
        self._threads = int(n)

    def cpu_math_library_num_threads(self):
        return self._threads

    def disable_glog_info(self):
        self._glog = False

    def enable_mkldnn(self):
        pass

    def switch_use_feed_fetch_ops(self, flag=False):
        pass

    def switch_specify_input_names(self, flag=True):
        pass

    def summary(self):
        rows = [("model file", self.prog_file()), ("params file", self.params_file()), ("use_gpu", self._use_gpu),
                ("gpu_device_id", self._device_id), ("precision", self._precision.name), ("cuda_graph", self._cuda_graph),
                ("cpu_math_thread", self._threads)]
        width = max(len(k) for k, _ in rows)
        return "\n".join(f"{k.ljust(width)}  {v}" for k, v in rows)


class Tensor:
    """An input / output handle of a Predictor."""

    def __init__(self, name, device, is_input):
        self._name, self._device, self._is_input = name, device, is_input
        self._value = None
        self._shape = None

    def name(self):
        return self._name

    def reshape(self, shape):
        self._shape = [int(s) for s in shape]

    def copy_from_cpu(self, data):
        if not isinstance(data, np.ndarray):
            raise TypeError("In copy_from_cpu, we only support numpy ndarray data type.")
        if not self._is_input:
            raise RuntimeError("copy_from_cpu is only legal on input handles")
        t = torch.from_numpy(np.ascontiguousarray(data))
        if self._shape is not None and list(t.shape) != self._shape:
            t = t.reshape(self._shape)
        self._value = t.to(self._device, non_blocking=True)

    def share_external_data(self, data):
        if not isinstance(data, torch.Tensor):
            raise TypeError("In share_external_data, we only support Tensor data type.")
        self._value = data if data.device == self._device else data.to(self._device)

    def copy_to_cpu(self):
        if self._value is None:
            raise RuntimeError(f"handle {self._name} holds no data yet (run the predictor first)")
        v = self._value.detach()
        if v.dtype == torch.bfloat16:
            v = v.float()
        return v.cpu().numpy()

    def shape(self):
        return list(self._value.shape) if self._value is not None else (self._shape or [])

    def type(self):
        return _DATATYPE_OF[self._value.dtype] if self._value is not None else DataType.FLOAT32

    def lod(self):
        return []

    def set_lod(self, lod):
        pass


class Predictor:
    def __init__(self, config):
        from .. import jit
        if not isinstance(config, Config) or not config._prefix:
            raise ValueError("create_predictor needs a Config with a model set")
        self._config = config
        layer = jit.load(config._prefix)
        if layer.program() is None:
            raise RuntimeError(f"{config._prefix} was saved without a program (no input_spec at paddle.jit.save): nothing to run")
        self._device = torch.device("cuda", config._device_id) if config._use_gpu and torch.cuda.is_available() else torch.device("cpu")
        if config._use_gpu and not torch.cuda.is_available():
            raise RuntimeError("Config.enable_use_gpu was called but no CUDA device is visible")
        self._module = layer.program().to(self._device)
        dtype = {PrecisionType.Half: torch.float16, PrecisionType.Bfloat16: torch.bfloat16}.get(config._precision)
        self._cast = dtype
        if dtype is not None:
            self._module = self._module.to(dtype)
        specs = layer.input_spec or []
        self._input_names = [s.get("name") or f"x{i}" for i, s in enumerate(specs)]
        n_out = layer.num_outputs or 1
        self._output_names = [f"save_infer_model/scale_{i}.tmp_0" for i in range(n_out)]
        self._inputs = {n: Tensor(n, self._device, True) for n in self._input_names}
        self._outputs = {n: Tensor(n, self._device, False) for n in self._output_names}
        self._runner = self._module
        if self._device.type == "cuda" and config._cuda_graph:
            self._runner = jit.to_static(lambda *a: self._module(*a))       # CUDA graph per input signature under no_grad
        torch.set_num_threads(max(config._threads, 1)) if self._device.type == "cpu" and config._threads > 1 else None

    def get_input_names(self):
        return list(self._input_names)

    def get_output_names(self):
        return list(self._output_names)

    def get_input_handle(self, name):
        return self._inputs[name]

    def get_output_handle(self, name):
        return self._outputs[name]

    def run(self, inputs=None):
        """Handle style: fill the input handles, ``run()``, read the output handles.  List style:
        ``outs = run([tensor, ...])`` returns the output tensors."""
        if inputs is not None:
            if len(inputs) != len(self._input_names):
                raise ValueError(f"the model takes {len(self._input_names)} inputs, got {len(inputs)}")
            for n, t in zip(self._input_names, inputs):
                self._inputs[n].share_external_data(t if isinstance(t, torch.Tensor) else torch.as_tensor(t))
        args = []
        for n in self._input_names:
            v = self._inputs[n]._value
            if v is None:
                raise RuntimeError(f"input {n} was not set")
            if self._cast is not None and v.dtype.is_floating_point:
                v = v.to(self._cast)
            args.append(v)
        with torch.no_grad():
            out = self._runner(*args)
        outs = list(out) if isinstance(out, (list, tuple)) else [out]
        for n, o in zip(self._output_names, outs):
            self._outputs[n]._value = o
        return outs if inputs is not None else True

    def clone(self):
        return Predictor(self._config)

    def clear_intermediate_tensor(self):
        pass

    def try_shrink_memory(self):
        if self._device.type == "cuda":
            torch.cuda.empty_cache()
        return 0


def create_predictor(config):
    return Predictor(config)


class PredictorPool:
    """``size`` predictors over one configuration (one per serving thread)."""

    def __init__(self, config, size=1):
        self._predictors = [Predictor(config) for _ in range(max(int(size), 1))]

    def retrieve(self, idx):
        return self._predictors[idx]

    retrive = retrieve                      # the reference's (misspelled) name


def convert_to_mixed_precision(model_file, params_file, mixed_model_file, mixed_params_file, mixed_precision, backend=PlaceType.GPU,
                               keep_io_types=True, black_list=None, **kwargs):
    """Rewrite a saved model with its floating-point parameters in half / bfloat16.  ``black_list``: parameter-name
    substrings to leave in float32.  The program file is copied; the Predictor casts activations at its boundary, so
    inputs and outputs keep their types (``keep_io_types``)."""
    from ..framework.io import load, save
    dtype = {PrecisionType.Half: torch.float16, PrecisionType.Bfloat16: torch.bfloat16}.get(PrecisionType(mixed_precision))
    if dtype is None:
        raise ValueError("mixed_precision should be PrecisionType.Half or PrecisionType.Bfloat16")
    black = set(black_list or ())
    state = load(params_file)
    out = {}
    for k, v in state.items():
        t = v if isinstance(v, torch.Tensor) else torch.as_tensor(v)
        keep = any(b in k for b in black) or not t.dtype.is_floating_point
        out[k] = t if keep else t.to(dtype)
    for path in (mixed_model_file, mixed_params_file):
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
    save(out, mixed_params_file)
    if os.path.abspath(model_file) != os.path.abspath(mixed_model_file):
        shutil.copyfile(model_file, mixed_model_file)
        meta = model_file + ".json"
        if os.path.exists(meta):
            shutil.copyfile(meta, mixed_model_file + ".json")
