"""Paddle dtype names on torch dtypes (reference: python/paddle/framework/dtype.py, base/data_feeder.py:convert_dtype)."""
import numpy as np
import torch

bool = torch.bool
uint8 = torch.uint8
int8 = torch.int8
int16 = torch.int16
int32 = torch.int32
int64 = torch.int64
float16 = torch.float16
bfloat16 = torch.bfloat16
float32 = torch.float32
float64 = torch.float64
complex64 = torch.complex64
complex128 = torch.complex128
float8_e4m3fn = torch.float8_e4m3fn
float8_e5m2 = torch.float8_e5m2
dtype = torch.dtype

_STR2DTYPE = {
    "bool": torch.bool, "uint8": torch.uint8, "int8": torch.int8, "int16": torch.int16,
    "int32": torch.int32, "int64": torch.int64, "float16": torch.float16, "half": torch.float16,
    "bfloat16": torch.bfloat16, "uint16": torch.bfloat16,  # paddle stores bf16 as uint16 ndarrays
    "float32": torch.float32, "float": torch.float32, "float64": torch.float64, "double": torch.float64,
    "complex64": torch.complex64, "complex128": torch.complex128,
    "float8_e4m3fn": torch.float8_e4m3fn, "float8_e5m2": torch.float8_e5m2,
}
_DTYPE2STR = {
    torch.bool: "bool", torch.uint8: "uint8", torch.int8: "int8", torch.int16: "int16",
    torch.int32: "int32", torch.int64: "int64", torch.float16: "float16", torch.bfloat16: "bfloat16",
    torch.float32: "float32", torch.float64: "float64", torch.complex64: "complex64",
    torch.complex128: "complex128", torch.float8_e4m3fn: "float8_e4m3fn", torch.float8_e5m2: "float8_e5m2",
}
_NP2TORCH = {
    np.dtype("bool"): torch.bool, np.dtype("uint8"): torch.uint8, np.dtype("int8"): torch.int8,
    np.dtype("int16"): torch.int16, np.dtype("int32"): torch.int32, np.dtype("int64"): torch.int64,
    np.dtype("float16"): torch.float16, np.dtype("float32"): torch.float32,
    np.dtype("float64"): torch.float64, np.dtype("complex64"): torch.complex64,
    np.dtype("complex128"): torch.complex128, np.dtype("uint16"): torch.bfloat16,
}

_default_dtype = torch.float32


def to_torch_dtype(d):
    """Accepts a torch dtype, a paddle-style string, a numpy dtype/type or None."""
    if d is None:
        return None
    if isinstance(d, torch.dtype):
        return d
    if isinstance(d, str):
        key = d.replace("paddle.", "")
        if key in _STR2DTYPE:
            return _STR2DTYPE[key]
        raise TypeError(f"unsupported dtype string {d!r}")
    try:
        return _NP2TORCH[np.dtype(d)]
    except Exception as e:  # noqa: BLE001
        raise TypeError(f"unsupported dtype {d!r}") from e


def convert_dtype(d):
    """dtype -> paddle-style string name."""
    return _DTYPE2STR[to_torch_dtype(d)]


def set_default_dtype(d):
    global _default_dtype
    d = to_torch_dtype(d)
    if d not in (torch.float16, torch.bfloat16, torch.float32, torch.float64):
        raise TypeError("set_default_dtype only supports float16/bfloat16/float32/float64")
    _default_dtype = d
    torch.set_default_dtype(d)


def get_default_dtype():
    return convert_dtype(_default_dtype)


def default_torch_dtype():
    return _default_dtype


def is_floating_point_dtype(d):
    return to_torch_dtype(d).is_floating_point


def finfo(d):
    return torch.finfo(to_torch_dtype(d))


def iinfo(d):
    return torch.iinfo(to_torch_dtype(d))
