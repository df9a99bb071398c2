"""Numerical debugging helpers (reference: python/paddle/amp/debugging.py, FLAGS_check_nan_inf +
paddle/phi/kernels/gpu/check_numerics_kernel.cu).  Implemented as module forward hooks / explicit checks."""
from __future__ import annotations

import contextlib
from enum import Enum

import torch

from ..framework import flags as _flags


class DebugMode(Enum):
    CHECK_NAN_INF_AND_ABORT = 0
    CHECK_NAN_INF = 1
    CHECK_ALL_FOR_OVERFLOW = 2
    CHECK_ALL = 4


class TensorCheckerConfig:
    def __init__(self, enable, debug_mode=DebugMode.CHECK_NAN_INF_AND_ABORT, output_dir=None, checked_op_list=None,
                 skipped_op_list=None, debug_step=None, stack_height_limit=1):
        self.enable, self.debug_mode, self.output_dir = enable, debug_mode, output_dir
        self.checked_op_list, self.skipped_op_list, self.debug_step = checked_op_list, skipped_op_list, debug_step


def check_numerics(tensor, op_type="", var_name="", debug_mode=DebugMode.CHECK_NAN_INF_AND_ABORT):
    """Returns (num_nan, num_inf, num_zero) tensors and raises in ABORT mode when a NaN/Inf is present."""
    t = tensor.detach().float()
    stats = torch.stack([torch.isnan(t).sum(), torch.isinf(t).sum(), (t == 0).sum()])
    if debug_mode == DebugMode.CHECK_NAN_INF_AND_ABORT and (stats[0] + stats[1]).item() > 0:
        raise RuntimeError(f"[check_numerics] op={op_type} var={var_name}: nan={int(stats[0])} inf={int(stats[1])}")
    return stats[0], stats[1], stats[2]


_handles = []


def enable_tensor_checker(checker_config):
    if not checker_config.enable:
        return
    _flags.set_flags({"check_nan_inf": True})

    def hook(mod, inp, out):
        outs = out if isinstance(out, (tuple, list)) else [out]
        for i, o in enumerate(outs):
            if isinstance(o, torch.Tensor) and o.is_floating_point():
                check_numerics(o, type(mod).__name__, f"out{i}", checker_config.debug_mode)

    _handles.append(torch.nn.modules.module.register_module_forward_hook(hook))


def disable_tensor_checker():
    _flags.set_flags({"check_nan_inf": False})
    while _handles:
        _handles.pop().remove()


def enable_operator_stats_collection():
    pass


def disable_operator_stats_collection():
    pass


@contextlib.contextmanager
def collect_operator_stats():
    yield


def compare_accuracy(dump_path, another_dump_path, output_filename, loss_scale=1, dump_all_tensors=False):
    raise NotImplementedError("accuracy_compare needs op-level dumps, which this runtime does not write")
