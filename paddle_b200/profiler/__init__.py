"""paddle.profiler (reference: python/paddle/profiler/{profiler,profiler_statistic,timer,utils}.py; native tracers
paddle/fluid/platform/profiler/*).  Kineto/CUPTI through torch.profiler does the tracing; this module keeps the
reference surface: ``Profiler(targets, scheduler, on_trace_ready, timer_only)``, the CLOSED/READY/RECORD/
RECORD_AND_RETURN state machine driven by ``step()``, ``make_scheduler``, ``export_chrome_tracing``,
``RecordEvent`` ranges (also emitted as NVTX), ``step_info()`` with reader_cost / batch_cost / ips."""
from __future__ import annotations

import contextlib
import os
import time
from enum import Enum

import torch


class ProfilerState(Enum):
    CLOSED = 0
    READY = 1
    RECORD = 2
    RECORD_AND_RETURN = 3


class ProfilerTarget(Enum):
    CPU = 0
    GPU = 1
    XPU = 2
    CUSTOM_DEVICE = 3


class SortedKeys(Enum):
    CPUTotal = 0
    CPUAvg = 1
    CPUMax = 2
    CPUMin = 3
    GPUTotal = 4
    GPUAvg = 5
    GPUMax = 6
    GPUMin = 7


class SummaryView(Enum):
    DeviceView = 0
    OverView = 1
    ModelView = 2
    DistributedView = 3
    KernelView = 4
    OperatorView = 5
    MemoryView = 6
    MemoryManipulationView = 7
    UDFView = 8


def make_scheduler(*, closed, ready, record, repeat=0, skip_first=0):
    def fn(step):
        assert step >= 0
        if step < skip_first:
            return ProfilerState.CLOSED
        step -= skip_first
        period = closed + ready + record
        if repeat > 0 and step // period >= repeat:
            return ProfilerState.CLOSED
        m = step % period
        if m < closed:
            return ProfilerState.CLOSED
        if m < closed + ready:
            return ProfilerState.READY
        if m < period - 1:
            return ProfilerState.RECORD
        return ProfilerState.RECORD_AND_RETURN
    assert closed >= 0 and ready >= 0 and record > 0 and repeat >= 0 and skip_first >= 0, "Invalid profiler scheduler arguments"
    if ready == 0:
        assert closed == 0 or True
    return fn


def _default_state_scheduler(step):
    return ProfilerState.RECORD


def export_chrome_tracing(dir_name, worker_name=None):
    os.makedirs(dir_name, exist_ok=True)

    def handle(prof):
        name = worker_name or f"host_{os.uname().nodename}pid_{os.getpid()}"
        path = os.path.join(dir_name, f"{name}_time_{time.strftime('%Y_%m_%d_%H_%M_%S')}.paddle_trace.json")
        prof.export(path, "json")
    return handle


def export_protobuf(dir_name, worker_name=None):
    return export_chrome_tracing(dir_name, worker_name)


class RecordEvent(contextlib.ContextDecorator):
    def __init__(self, name, event_type=None):
        self.name = name
        self._rf = None

    def begin(self):
        self._rf = torch.profiler.record_function(self.name)
        self._rf.__enter__()
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(self.name)

    def end(self):
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_pop()
        if self._rf is not None:
            self._rf.__exit__(None, None, None)
            self._rf = None

    def __enter__(self):
        self.begin()
        return self

    def __exit__(self, *a):
        self.end()
        return False


class _Benchmark:
    """reader_cost / batch_cost / ips bookkeeping (reference: profiler/timer.py)."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.step_start = None
        self.batch_costs, self.reader_costs, self.samples = [], [], []
        self._reader_t0 = None
        self._reader_acc = 0.0

    def begin(self):
        self.step_start = time.perf_counter()

    def before_reader(self):
        self._reader_t0 = time.perf_counter()

    def after_reader(self):
        if self._reader_t0 is not None:
            self._reader_acc += time.perf_counter() - self._reader_t0
            self._reader_t0 = None

    def step(self, num_samples=None):
        now = time.perf_counter()
        if self.step_start is not None:
            self.batch_costs.append(now - self.step_start)
            self.reader_costs.append(self._reader_acc)
            self.samples.append(num_samples)
        self._reader_acc = 0.0
        self.step_start = now

    def info(self, unit=None):
        if not self.batch_costs:
            return ""
        bc = sum(self.batch_costs) / len(self.batch_costs)
        rc = sum(self.reader_costs) / len(self.reader_costs)
        s = f"reader_cost: {rc:.5f} s batch_cost: {bc:.5f} s"
        ns = [x for x in self.samples if x]
        if ns:
            s += f" ips: {sum(ns) / sum(self.batch_costs):.3f} {unit or 'samples'}/s"
        else:
            s += f" ips: {1.0 / bc:.3f} steps/s"
        self.batch_costs, self.reader_costs, self.samples = [], [], []
        return s


def benchmark():
    return _bench


_bench = _Benchmark()


class Profiler:
    def __init__(self, *, targets=None, scheduler=None, on_trace_ready=None, record_shapes=False, profile_memory=False,
                 timer_only=False, emit_nvtx=False, custom_device_types=None, with_flops=False):
        self.targets = list(targets) if targets else [ProfilerTarget.CPU] + ([ProfilerTarget.GPU] if torch.cuda.is_available() else [])
        if scheduler is None:
            self.scheduler = _default_state_scheduler
        elif isinstance(scheduler, (tuple, list)):
            start, end = scheduler
            assert 0 <= start < end
            self.scheduler = make_scheduler(closed=max(start - 1, 0), ready=1 if start > 0 else 0, record=end - start, repeat=1)
        else:
            self.scheduler = scheduler
        self.on_trace_ready = on_trace_ready if on_trace_ready is not None else export_chrome_tracing("./profiler_log/")
        self.record_shapes, self.profile_memory, self.timer_only, self.with_flops = record_shapes, profile_memory, timer_only, with_flops
        self.step_num = 0
        self.previous_state = ProfilerState.CLOSED
        self.current_state = self.scheduler(self.step_num)
        self._prof = None
        self._last = None
        self._step_range = None

    def _torch_profiler(self):
        acts = [torch.profiler.ProfilerActivity.CPU]
        if ProfilerTarget.GPU in self.targets and torch.cuda.is_available():
            acts.append(torch.profiler.ProfilerActivity.CUDA)
        return torch.profiler.profile(activities=acts, record_shapes=self.record_shapes, profile_memory=self.profile_memory,
                                      with_flops=self.with_flops)

    def _start_record(self):
        if self._prof is None and not self.timer_only:
            self._prof = self._torch_profiler()
            self._prof.__enter__()

    def _stop_record(self, deliver):
        if self._prof is not None:
            self._prof.__exit__(None, None, None)
            self._last = self._prof
            self._prof = None
            if deliver and self.on_trace_ready:
                self.on_trace_ready(self)

    def start(self):
        _bench.reset()
        _bench.begin()
        self.current_state = self.scheduler(self.step_num)
        if self.current_state in (ProfilerState.RECORD, ProfilerState.RECORD_AND_RETURN):
            self._start_record()
        self._step_range = RecordEvent(f"ProfileStep#{self.step_num}")
        self._step_range.begin()

    def stop(self):
        if self._step_range:
            self._step_range.end()
            self._step_range = None
        if self.current_state in (ProfilerState.RECORD, ProfilerState.RECORD_AND_RETURN):
            self._stop_record(True)

    def step(self, num_samples=None):
        _bench.step(num_samples)
        if self._step_range:
            self._step_range.end()
        self.previous_state = self.current_state
        self.step_num += 1
        self.current_state = self.scheduler(self.step_num)
        prev, cur = self.previous_state, self.current_state
        recording = (ProfilerState.RECORD, ProfilerState.RECORD_AND_RETURN)
        if prev == ProfilerState.RECORD_AND_RETURN:
            self._stop_record(True)
            if cur in recording:
                self._start_record()
        elif prev in recording and cur not in recording:
            self._stop_record(True)
        elif prev not in recording and cur in recording:
            self._start_record()
        self._step_range = RecordEvent(f"ProfileStep#{self.step_num}")
        self._step_range.begin()

    def step_info(self, unit=None):
        return _bench.info(unit)

    def export(self, path="", format="json"):
        if self._last is not None:
            self._last.export_chrome_trace(path)

    def summary(self, sorted_by=SortedKeys.CPUTotal, op_detail=True, thread_sep=False, time_unit="ms", views=None):
        if self._last is None:
            return ""
        key = "cuda_time_total" if sorted_by.name.startswith("GPU") else "cpu_time_total"
        tbl = self._last.key_averages().table(sort_by=key, row_limit=50)
        print(tbl)
        return tbl

    def __enter__(self):
        self.start()
        return self

    def __exit__(self, *a):
        self.stop()
        return False


def load_profiler_result(filename):
    import json
    with open(filename) as f:
        return json.load(f)


@contextlib.contextmanager
def _nvprof_range(iter_id, start, end, exit_after_prof=True):
    if iter_id == start:
        torch.cuda.profiler.start()
    yield
    if iter_id == end - 1:
        torch.cuda.profiler.stop()
