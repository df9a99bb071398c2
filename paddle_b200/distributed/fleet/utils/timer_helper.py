"""Timers with device sync (reference: python/paddle/distributed/fleet/utils/timer_helper.py:34-113)."""
import time

import torch

_GLOBAL_TIMERS = None


def is_timer_initialized():
    return _GLOBAL_TIMERS is not None


def get_timers():
    assert _GLOBAL_TIMERS is not None, "timers are not initialized"
    return _GLOBAL_TIMERS


def set_timers():
    global _GLOBAL_TIMERS
    _GLOBAL_TIMERS = Timers()


class _Timer:
    def __init__(self, name):
        self.name, self.elapsed_, self.started_, self.start_time = name, 0.0, False, 0.0

    def _sync(self):
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def start(self):
        assert not self.started_, "timer has already started"
        self._sync()
        self.start_time = time.time()
        self.started_ = True

    def stop(self):
        assert self.started_, "timer is not started."
        self._sync()
        self.elapsed_ += time.time() - self.start_time
        self.started_ = False

    def reset(self):
        self.elapsed_, self.started_ = 0.0, False

    def elapsed(self, reset=True):
        started = self.started_
        if started:
            self.stop()
        e = self.elapsed_
        if reset:
            self.reset()
        if started:
            self.start()
        return e


class Timers:
    def __init__(self):
        self.timers = {}

    def __call__(self, name):
        if name not in self.timers:
            self.timers[name] = _Timer(name)
        return self.timers[name]

    def log(self, names, normalizer=1.0, reset=True):
        assert normalizer > 0.0
        s = "time (ms)"
        for n in names:
            s += f" | {n}: {self.timers[n].elapsed(reset=reset) * 1000.0 / normalizer:.2f}"
        print(s, flush=True)
        return s
