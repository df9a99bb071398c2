"""paddle.onnx (reference: python/paddle/onnx/export.py, which hands the traced program to the paddle2onnx package).

Export goes through ``torch.onnx`` on the same trace ``paddle.jit.save`` uses (framework kernels switched off while
tracing, so only standard operators appear).  It needs the ``onnx`` package, like the reference needs paddle2onnx; when
that is missing the call says so."""
from __future__ import annotations

import os

import torch

__all__ = ["export"]


def export(layer, path, input_spec=None, opset_version=9, **configs):
    """Write ``<path>.onnx``.  ``input_spec``: InputSpec objects or example tensors (unknown dims become dynamic axes)."""
    import importlib.util
    if importlib.util.find_spec("onnx") is None:
        raise ImportError("paddle.onnx.export needs the `onnx` package, which is not installed in this environment "
                          "(the reference needs `paddle2onnx` in the same way)")
    from ..framework.flags import get_flags, set_flags
    from ..jit import _example_inputs
    if input_spec is None:
        raise ValueError("input_spec is required to trace the model for export")
    params = list(layer.parameters())
    device = params[0].device if params else torch.device("cpu")
    examples = _example_inputs(input_spec, device)
    dynamic = {}
    names = []
    for i, spec in enumerate(input_spec):
        name = getattr(spec, "name", None) or f"x{i}"
        names.append(name)
        dims = {d: f"{name}_dim{d}" for d, s in enumerate(getattr(spec, "shape", ())) if s is None or (isinstance(s, int) and s < 0)}
        if dims:
            dynamic[name] = dims
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    old = get_flags(["FLAGS_use_native_kernels", "FLAGS_use_tcgen05_gemm", "FLAGS_use_fused_comm"])
    set_flags({"FLAGS_use_native_kernels": False, "FLAGS_use_tcgen05_gemm": False, "FLAGS_use_fused_comm": False})
    was_training = layer.training
    try:
        layer.eval()
        torch.onnx.export(layer, tuple(examples), path + ".onnx", input_names=names, dynamic_axes=dynamic or None,
                          opset_version=max(int(opset_version), 13), dynamo=False)
    finally:
        set_flags(old)
        layer.train(was_training)
    return path + ".onnx"
