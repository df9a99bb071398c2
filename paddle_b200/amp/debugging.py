"""Numerical debugging for mixed precision (reference: python/paddle/amp/debugging.py:27-700,
python/paddle/amp/accuracy_compare.py, paddle/phi/kernels/gpu/check_numerics_kernel.cu, FLAGS_check_nan_inf).

* ``check_numerics`` - one pass over a tensor on the device (csrc/kernels/check_numerics.cu): NaN / Inf / zero counts
  and max / min / mean.  Nothing is read back unless the mode needs a decision on the host (abort, or a log line).
* ``TensorCheckerConfig`` + ``enable_tensor_checker`` - every layer output (and, in CHECK_ALL modes, every op output
  seen by a dispatch hook) is checked inside the configured step window; findings go to stdout or to
  ``output_dir/worker_<rank>.log`` as ``[PRECISION]`` lines.
* ``collect_operator_stats`` - counts op calls per compute dtype (fp16 / bf16 / fp32 / other) with a dispatch mode.
* ``compare_accuracy`` - joins the ``[PRECISION]`` logs of an fp32 run and a low-precision run per (op, tensor) and
  writes a table (csv; xlsx when xlsxwriter is importable) flagging tensors whose statistics drift apart.
"""
from __future__ import annotations

import contextlib
import csv
import os
import random
from enum import Enum

import torch

from ..framework import flags as _flags

__all__ = ["DebugMode", "TensorCheckerConfig", "check_numerics", "enable_operator_stats_collection",
           "disable_operator_stats_collection", "collect_operator_stats", "enable_tensor_checker",
           "disable_tensor_checker", "compare_accuracy", "check_layer_numerics", "set_checked_op_list",
           "set_skipped_op_list"]


class DebugMode(Enum):
    CHECK_NAN_INF_AND_ABORT = 0
    CHECK_NAN_INF = 1
    CHECK_ALL_FOR_OVERFLOW = 2
    CHECK_ALL = 4


_checked_ops: set = set()
_skipped_ops: set = set()


def set_checked_op_list(checked_op_list):
    _checked_ops.clear()
    _checked_ops.update(checked_op_list or [])


def set_skipped_op_list(skipped_op_list):
    _skipped_ops.clear()
    _skipped_ops.update(skipped_op_list or [])


class TensorCheckerConfig:
    """Which steps and ops are checked, and where the findings go (reference :156-336)."""
    current_step_id = 0

    def __init__(self, enable, debug_mode=DebugMode.CHECK_NAN_INF_AND_ABORT, output_dir=None, checked_op_list=None,
                 skipped_op_list=None, debug_step=None, stack_height_limit=1):
        self.enable = bool(enable)
        if not isinstance(debug_mode, DebugMode):
            raise ValueError(f"debug_mode must be a DebugMode, got {debug_mode!r}")
        self.debug_mode = debug_mode
        self.output_dir = output_dir
        self.checked_op_list = list(checked_op_list or [])
        self.skipped_op_list = list(skipped_op_list or [])
        self.debug_step = debug_step
        self.stack_height_limit = stack_height_limit
        self.start_step, self.end_step = None, None
        if debug_step is not None:
            if not (isinstance(debug_step, (tuple, list)) and len(debug_step) == 2 and debug_step[1] > debug_step[0]):
                raise ValueError("debug_step must be (start, end) with end > start")
            self.start_step, self.end_step = int(debug_step[0]), int(debug_step[1])
        self.seed = 123
        self.initial_seed = 123
        self._fh = None

    def keep_random(self):
        torch.manual_seed(self.seed)
        random.seed(self.seed)

    def update_and_check_step_id(self):
        """Advance the step counter; True while the current step is inside the debug window."""
        if not self.enable:
            return False
        if self.start_step is not None:
            TensorCheckerConfig.current_step_id += 1
            return self.start_step <= TensorCheckerConfig.current_step_id < self.end_step
        return True

    def start_check_nan_inf(self):
        if self.enable:
            self.keep_random()
            set_checked_op_list(self.checked_op_list)
            set_skipped_op_list(self.skipped_op_list)
            _flags.set_flags({"check_nan_inf": True})

    def stop_check_nan_inf(self):
        _flags.set_flags({"check_nan_inf": False})

    def _log(self, line):
        if self.output_dir:
            if self._fh is None:
                os.makedirs(self.output_dir, exist_ok=True)
                rank = os.environ.get("PADDLE_TRAINER_ID", os.environ.get("RANK", "0"))
                self._fh = open(os.path.join(self.output_dir, f"worker_{rank}.log"), "a")
            self._fh.write(line + "\n")
            self._fh.flush()
        else:
            print(line, flush=True)


def _stats(tensor):
    """-> (counts int64[3] = nan, inf, zero; values float[3] = max, min, mean) as tensors on the tensor's device."""
    t = tensor.detach()
    from ..ops._util import K, check, dtype_code, ptr, stream_ptr, use_native
    if t.is_cuda and t.dtype in (torch.float32, torch.bfloat16, torch.float16) and use_native(t) and hasattr(K(), "pdb200_check_numerics"):
        tc = t.contiguous()
        counts = torch.empty(3, dtype=torch.int64, device=t.device)
        values = torch.empty(3, dtype=torch.float32, device=t.device)
        check(K().pdb200_check_numerics(ptr(tc), tc.numel(), dtype_code(tc), ptr(counts), ptr(values), stream_ptr(t.device)),
              "check_numerics")
        finite = (tc.numel() - counts[0] - counts[1]).clamp(min=1).float()
        values = torch.stack([values[0], values[1], values[2] / finite])
        return counts, values
    f = t.float()
    nan, inf = torch.isnan(f), torch.isinf(f)
    fin = ~(nan | inf)
    counts = torch.stack([nan.sum(), inf.sum(), ((f == 0) & fin).sum()]).to(torch.int64)
    if bool(fin.any()):
        ff = f[fin]
        values = torch.stack([ff.max(), ff.min(), ff.mean()])
    else:
        values = torch.tensor([float("-inf"), float("inf"), 0.0], device=f.device)
    return counts, values


def _precision_line(level, op_type, var_name, tensor, counts, values):
    place = f"gpu:{tensor.device.index}" if tensor.is_cuda else "cpu"
    c, v = counts.tolist(), values.tolist()
    return (f"[PRECISION] [{level}] in [device={place}, op={op_type}, tensor={var_name}, dtype={str(tensor.dtype).replace('torch.', '')}], "
            f"numel={tensor.numel()}, num_nan={c[0]}, num_inf={c[1]}, num_zero={c[2]}, max={v[0]:e}, min={v[1]:e}, mean={v[2]:e}")


def check_numerics(tensor, op_type="", var_name="", debug_mode=DebugMode.CHECK_NAN_INF_AND_ABORT, _config=None):
    """(stats, values): ``stats`` = int64[3] (num_nan, num_inf, num_zero), ``values`` = float32[3] (max, min, mean),
    both on the tensor's device.  CHECK_NAN_INF_AND_ABORT raises when a NaN / Inf is present; CHECK_NAN_INF reports it;
    CHECK_ALL reports every tensor; CHECK_ALL_FOR_OVERFLOW reports fp32 tensors outside the fp16 range."""
    counts, values = _stats(tensor)
    log = _config._log if _config is not None else (lambda s: print(s, flush=True))
    if debug_mode == DebugMode.CHECK_ALL:
        log(_precision_line("INFO", op_type, var_name, tensor, counts, values))
        return counts, values
    bad = int(counts[0] + counts[1]) > 0                       # the one host read of the checking modes
    if debug_mode == DebugMode.CHECK_ALL_FOR_OVERFLOW:
        v = values.tolist()
        over = tensor.dtype == torch.float32 and (max(abs(v[0]), abs(v[1])) > 65504.0)
        if bad or over:
            log(_precision_line("ERROR" if bad else "WARNING", op_type, var_name, tensor, counts, values))
        return counts, values
    if bad:
        line = _precision_line("ERROR", op_type, var_name, tensor, counts, values)
        log(line)
        if debug_mode == DebugMode.CHECK_NAN_INF_AND_ABORT:
            raise RuntimeError(f"There are NAN or INF (num_nan={int(counts[0])}, num_inf={int(counts[1])}) in {line}")
    return counts, values


def check_layer_numerics(func):
    """Decorator for ``Layer.forward``: checks the first tensor argument and every tensor output (reference :63-134)."""
    def wrapper(self, *args, **kwargs):
        if args and isinstance(args[0], torch.Tensor):
            check_numerics(args[0], type(self).__name__, "input", DebugMode.CHECK_NAN_INF_AND_ABORT)
        elif args:
            raise RuntimeError("First input of this layer must be tensor.")
        out = func(self, *args, **kwargs)
        outs = out if isinstance(out, (tuple, list)) else [out]
        for i, o in enumerate(outs):
            if isinstance(o, torch.Tensor) and o.is_floating_point():
                check_numerics(o, type(self).__name__, f"output{i}", DebugMode.CHECK_NAN_INF_AND_ABORT)
        return out
    return wrapper


_handles = []
_active_config = [None]
_dispatch_mode = [None]


def _op_selected(name):
    if _skipped_ops and name in _skipped_ops:
        return False
    return not _checked_ops or name in _checked_ops


class _OpChecker(torch.utils._python_dispatch.TorchDispatchMode):
    """CHECK_ALL / CHECK_ALL_FOR_OVERFLOW: every floating-point op output, named by the aten op."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.busy = False

    def __torch_dispatch__(self, func, types, args=(), kwargs=None):
        out = func(*args, **(kwargs or {}))
        if self.busy:
            return out
        name = func.overloadpacket.__name__ if hasattr(func, "overloadpacket") else str(func)
        if not _op_selected(name):
            return out
        self.busy = True
        try:
            outs = out if isinstance(out, (tuple, list)) else [out]
            for i, o in enumerate(outs):
                if isinstance(o, torch.Tensor) and o.is_floating_point() and o.numel() > 0:
                    check_numerics(o, name, f"out{i}", self.cfg.debug_mode, _config=self.cfg)
        finally:
            self.busy = False
        return out


def enable_tensor_checker(checker_config):
    """Start checking (reference :628-667).  Call once per step: it advances the config's step window."""
    if not checker_config.update_and_check_step_id():
        disable_tensor_checker()
        return
    if _active_config[0] is checker_config:
        return
    disable_tensor_checker()
    checker_config.start_check_nan_inf()
    _active_config[0] = checker_config

    def hook(mod, inp, out):
        name = type(mod).__name__
        if not _op_selected(name):
            return
        outs = out if isinstance(out, (tuple, list)) else [out]
        for i, o in enumerate(outs):
            if isinstance(o, torch.Tensor) and o.is_floating_point():
                check_numerics(o, name, f"out{i}", checker_config.debug_mode, _config=checker_config)

    _handles.append(torch.nn.modules.module.register_module_forward_hook(hook))
    if checker_config.debug_mode in (DebugMode.CHECK_ALL, DebugMode.CHECK_ALL_FOR_OVERFLOW):
        m = _OpChecker(checker_config)
        m.__enter__()
        _dispatch_mode[0] = m


def disable_tensor_checker():
    if _active_config[0] is not None:
        _active_config[0].stop_check_nan_inf()
        _active_config[0] = None
    while _handles:
        _handles.pop().remove()
    if _dispatch_mode[0] is not None:
        _dispatch_mode[0].__exit__(None, None, None)
        _dispatch_mode[0] = None


# --------------------------------------------------------------------------------------------- operator statistics
class _OpStats(torch.utils._python_dispatch.TorchDispatchMode):
    def __init__(self):
        super().__init__()
        self.table = {}

    def __torch_dispatch__(self, func, types, args=(), kwargs=None):
        out = func(*args, **(kwargs or {}))
        first = out[0] if isinstance(out, (tuple, list)) and out else out
        if isinstance(first, torch.Tensor) and first.is_floating_point():
            name = func.overloadpacket.__name__ if hasattr(func, "overloadpacket") else str(func)
            row = self.table.setdefault(name, [0, 0, 0, 0])          # fp16, bf16, fp32, other
            col = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}.get(first.dtype, 3)
            row[col] += 1
        return out


_stats_mode = [None]


def _print_operator_stats(table):
    print("<{:-^120}>".format(" op list "))
    print("<{:-^40}".format(" Op Name "), "|", "{:-^17}".format(" FP16 Calls "), "|", "{:-^17}".format(" BF16 Calls "), "|",
          "{:-^17}".format(" FP32 Calls "), "|", "{:-^17}>".format(" Other Calls "))
    for name in sorted(table):
        a, b, c, d = table[name]
        print(f"  {name:<40}|  {a:<17}|  {b:<17}|  {c:<17}|  {d:<17}")
    print("<{:-^120}>\n".format(f" op count: {len(table)} "))


def enable_operator_stats_collection():
    if _stats_mode[0] is None:
        m = _OpStats()
        m.__enter__()
        _stats_mode[0] = m


def disable_operator_stats_collection():
    m = _stats_mode[0]
    if m is None:
        return {}
    m.__exit__(None, None, None)
    _stats_mode[0] = None
    _print_operator_stats(m.table)
    return m.table


@contextlib.contextmanager
def collect_operator_stats():
    enable_operator_stats_collection()
    try:
        yield
    finally:
        disable_operator_stats_collection()


# --------------------------------------------------------------------------------------------- accuracy comparison
def _parse_precision_logs(path):
    """{(op, tensor, occurrence): record} from every ``worker_*.log`` under ``path`` (or one file)."""
    files = []
    if os.path.isdir(path):
        files = sorted(os.path.join(path, f) for f in os.listdir(path) if f.endswith(".log"))
    elif os.path.isfile(path):
        files = [path]
    out, seen = {}, {}
    for f in files:
        with open(f, errors="replace") as fh:
            for line in fh:
                if "[PRECISION]" not in line:
                    continue
                try:
                    head, tail = line.split("],", 1)
                    meta = dict(kv.split("=", 1) for kv in head.split("[", 3)[-1].replace(" ", "").split(","))
                    nums = dict(kv.strip().split("=", 1) for kv in tail.strip().split(","))
                    key0 = (meta["op"], meta["tensor"])
                    k = seen.get(key0, 0)
                    seen[key0] = k + 1
                    out[(meta["op"], meta["tensor"], k)] = dict(dtype=meta["dtype"], numel=int(nums["numel"]),
                                                                num_nan=int(nums["num_nan"]), num_inf=int(nums["num_inf"]),
                                                                num_zero=int(nums["num_zero"]), max=float(nums["max"]),
                                                                min=float(nums["min"]), mean=float(nums["mean"]))
                except (KeyError, ValueError, IndexError):
                    continue
    return out


def compare_accuracy(dump_path, another_dump_path, output_filename, loss_scale=1, dump_all_tensors=False):
    """Join the CHECK_ALL logs of two runs (typically fp32 and fp16 / bf16) and write one row per tensor with both
    runs' statistics and a status: ``nan_inf`` (only one run is finite), ``diff`` (max / min / mean differ by more than
    1e-2 relative after dividing the second run by ``loss_scale``) or ``ok``.  Returns the rows."""
    a, b = _parse_precision_logs(dump_path), _parse_precision_logs(another_dump_path)
    rows = []
    for key in sorted(set(a) | set(b)):
        ra, rb = a.get(key), b.get(key)
        row = {"op": key[0], "tensor": key[1], "occurrence": key[2]}
        for tag, r in (("a", ra), ("b", rb)):
            for f in ("dtype", "numel", "num_nan", "num_inf", "num_zero", "max", "min", "mean"):
                row[f"{tag}_{f}"] = None if r is None else r[f]
        status = "missing" if ra is None or rb is None else "ok"
        if status == "ok":
            bad_a, bad_b = ra["num_nan"] + ra["num_inf"] > 0, rb["num_nan"] + rb["num_inf"] > 0
            if bad_a != bad_b:
                status = "nan_inf"
            else:
                for f in ("max", "min", "mean"):
                    x, y = ra[f], rb[f] / float(loss_scale)
                    if abs(x - y) > 1e-2 + 1e-2 * abs(x):
                        status = "diff"
        row["status"] = status
        if dump_all_tensors or status != "ok":
            rows.append(row)
    fields = list(rows[0].keys()) if rows else ["op", "tensor", "occurrence", "status"]
    try:
        import xlsxwriter  # noqa: F401
        wb = xlsxwriter.Workbook(output_filename)
        ws = wb.add_worksheet("compare")
        for j, f in enumerate(fields):
            ws.write(0, j, f)
        for i, r in enumerate(rows, 1):
            for j, f in enumerate(fields):
                ws.write(i, j, r[f])
        wb.close()
    except ImportError:
        path = output_filename if output_filename.endswith(".csv") else os.path.splitext(output_filename)[0] + ".csv"
        with open(path, "w", newline="") as fh:
            w = csv.DictWriter(fh, fieldnames=fields)
            w.writeheader()
            w.writerows(rows)
    return rows
