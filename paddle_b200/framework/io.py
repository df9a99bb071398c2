"""paddle.save / paddle.load (reference: python/paddle/framework/io.py:130-150, 395-462, 743-982, 1150-1241;
io_utils.py:218-275).

On-disk format is the reference's: a pickle (protocol 2..4) of ``{structured_key: numpy.ndarray}`` plus
``"StructuredToParameterName@@": {structured_key: tensor.name}`` for state dicts; bf16 tensors travel as uint16
ndarrays; arrays above 2**30 bytes are split under protocol 2/3 with an ``"UnpackBigParamInfor@@"`` table;
non-state-dict objects have every Tensor reduced to the tuple ``(name, ndarray)``.
"""
from __future__ import annotations

import copyreg
import io as _io
import os
import pickle
import threading

import numpy as np
import torch

_MAX_BYTES = 2 ** 30


def _to_numpy(t: torch.Tensor) -> np.ndarray:
    return t.numpy()  # patched: detach + cpu + bf16->uint16


def _contains_tensor(obj):
    if isinstance(obj, torch.Tensor):
        return True
    if isinstance(obj, dict):
        return any(_contains_tensor(v) for v in obj.values())
    if isinstance(obj, (list, tuple)):
        return any(_contains_tensor(v) for v in obj)
    return False


def _is_state_dict(obj):
    """The reference's predicate (io.py:130-150): a dict whose values are Tensors, or dicts that hold no Tensor at
    any depth.  Anything else - an optimizer state with ``master_weights: {name: Tensor}``, a dict with a scalar at
    the top level - is NOT a state dict and goes through the generic pickler (Tensors become (name, ndarray))."""
    if not isinstance(obj, dict):
        return False
    for v in obj.values():
        if isinstance(v, dict):
            if any(_contains_tensor(x) for x in v.values()):
                return False
        elif not isinstance(v, torch.Tensor):
            return False
    return True


def _build_saved_state_dict(state_dict):
    save, names = {}, {}
    for k, v in state_dict.items():
        if isinstance(v, torch.Tensor):
            save[k] = _to_numpy(v)
            names[k] = v.name
        elif isinstance(v, dict):
            save[k] = {kk: (_to_numpy(vv) if isinstance(vv, torch.Tensor) else vv) for kk, vv in v.items()}
        else:
            save[k] = v
    save["StructuredToParameterName@@"] = names
    return save


def _unpack_big(saved, protocol):
    if protocol >= 4:
        return saved
    info = {}
    for k in list(saved.keys()):
        v = saved[k]
        if isinstance(v, np.ndarray) and v.nbytes > _MAX_BYTES:
            flat = v.reshape(-1)
            per = _MAX_BYTES // v.itemsize
            parts = []
            for i in range(0, flat.size, per):
                name = f"{k}@@.{len(parts)}"
                saved[name] = flat[i:i + per]
                parts.append(name)
            info[k] = {"OriginShape": v.shape, "slices": parts}
            del saved[k]
    if info:
        saved["UnpackBigParamInfor@@"] = info
    return saved


def _pack_big(loaded):
    info = loaded.pop("UnpackBigParamInfor@@", None)
    if info:
        for k, meta in info.items():
            loaded[k] = np.concatenate([loaded.pop(n) for n in meta["slices"]]).reshape(meta["OriginShape"])
    return loaded


def _reduce_tensor(t):
    return (tuple, ((t.name, _to_numpy(t)),))


class _TensorPickler(pickle.Pickler):
    dispatch_table = copyreg.dispatch_table.copy()

    def reducer_override(self, obj):
        if isinstance(obj, torch.Tensor):
            return _reduce_tensor(obj)
        return NotImplemented


def save(obj, path, protocol=4, **configs):
    if not isinstance(protocol, int) or protocol < 2 or protocol > 4:
        raise ValueError(f"Expected 1<'protocol'<5, but received protocol={protocol}")
    f = path if hasattr(path, "write") else None
    if f is None:
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
    if _is_state_dict(obj):
        saved = _unpack_big(_build_saved_state_dict(obj), protocol)
        data = pickle.dumps(saved, protocol=protocol)
    else:
        buf = _io.BytesIO()
        _TensorPickler(buf, protocol).dump(obj)
        data = buf.getvalue()
    if f is not None:
        f.write(data)
    else:
        tmp = path + ".tmp"
        with open(tmp, "wb") as fh:
            fh.write(data)
        os.replace(tmp, path)


def _ndarray_to_tensor(arr, name=None):
    if arr.dtype == np.uint16:
        t = torch.from_numpy(arr.copy()).view(torch.bfloat16)
    else:
        t = torch.from_numpy(np.ascontiguousarray(arr))
    if name:
        t.name = name
    return t


def _restore(obj, return_numpy):
    if isinstance(obj, tuple) and len(obj) == 2 and isinstance(obj[0], str) and isinstance(obj[1], np.ndarray):
        return obj[1] if return_numpy else _ndarray_to_tensor(obj[1], obj[0])
    if isinstance(obj, dict):
        return {k: _restore(v, return_numpy) for k, v in obj.items()}
    if isinstance(obj, list):
        return [_restore(v, return_numpy) for v in obj]
    if isinstance(obj, tuple):
        return tuple(_restore(v, return_numpy) for v in obj)
    return obj


def load(path, **configs):
    return_numpy = configs.get("return_numpy", False)
    keep_name_table = configs.get("keep_name_table", False)
    if hasattr(path, "read"):
        loaded = pickle.load(path, encoding="latin1")
    else:
        if not os.path.exists(path):
            raise ValueError(f"The ``path`` ({path}) to load model not exists.")
        with open(path, "rb") as f:
            loaded = pickle.load(f, encoding="latin1")
    if isinstance(loaded, dict) and "StructuredToParameterName@@" in loaded:
        loaded = _pack_big(loaded)
        names = loaded["StructuredToParameterName@@"]
        out = {}
        for k, v in loaded.items():
            if k == "StructuredToParameterName@@":
                if keep_name_table:
                    out[k] = v
                continue
            if isinstance(v, np.ndarray):
                out[k] = v if return_numpy else _ndarray_to_tensor(v, names.get(k))
            elif isinstance(v, dict):
                out[k] = {kk: (vv if return_numpy or not isinstance(vv, np.ndarray) else _ndarray_to_tensor(vv, kk))
                          for kk, vv in v.items()}
            else:
                out[k] = v
        return out
    if isinstance(loaded, dict):
        loaded = _pack_big(loaded)
    return _restore(loaded, return_numpy)


def async_save(obj, path, protocol=4, sync_other_task=False, **configs):
    """Background-thread save (reference: io.py:67-127): tensors are snapshotted to host first."""
    snap = _snapshot(obj)
    t = threading.Thread(target=save, args=(snap, path, protocol), kwargs=configs, daemon=False)
    t.start()
    _async_tasks.append(t)
    return t


_async_tasks = []


def clear_async_save_task_queue():
    while _async_tasks:
        _async_tasks.pop().join()


def _snapshot(obj):
    if isinstance(obj, torch.Tensor):
        c = obj.detach().cpu().clone()
        c.name = obj.name
        return c
    if isinstance(obj, dict):
        return {k: _snapshot(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_snapshot(v) for v in obj)
    return obj
