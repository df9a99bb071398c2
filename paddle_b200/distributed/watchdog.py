"""Communication watchdog / async trace (reference: paddle/phi/core/distributed/comm_task_manager.cc:65-280,
nccl_comm_task.cc:30-268; enabled by FLAGS_enable_async_trace).

Every collective registers a record (group id, global ranks, per-group sequence number, op, numel, sync flag, start
time).  The native manager in libpdb200_runtime.so runs the polling thread: it classifies records older than the
timeout as *not finished*, logs one trace line per offender plus the last finished task of the group, and can
abort the process so the launcher restarts the job.  When the native library is not built the same bookkeeping is
kept in Python (no background thread) so ``dump()`` still works for post-mortems.
"""
from __future__ import annotations

import threading
import time

from .. import lib
from ..framework.flags import flag

_lock = threading.Lock()
_seq = {}
_inflight = {}
_last_done = {}
_native = None
_groups = {}


def _rt():
    global _native
    if _native is None:
        _native = lib.runtime if lib.runtime.available() and hasattr(lib.runtime.require(), "pdb200_wd_start") else False
    return _native


def enabled():
    return bool(flag("enable_async_trace"))


def note_group(g):
    _groups[g.id] = list(g.ranks)


def start(rank):
    if not enabled():
        return
    rt = _rt()
    if rt:
        import atexit
        atexit.register(shutdown)
        poll = int(flag("comm_watchdog_poll_ms")) if "comm_watchdog_poll_ms" in _all_flags() else 10000
        rt.pdb200_wd_start(int(rank), int(flag("comm_watchdog_timeout_ms")), poll)
        rt.pdb200_wd_set_abort(int(bool(flag("comm_watchdog_abort"))) if "comm_watchdog_abort" in _all_flags() else 0)
        # the fused kernels' barriers publish the last epoch they passed in a pinned host word the trace prints
        if hasattr(rt.require(), "pdb200_wd_progress_ptr") and lib.have_kernels() and _cuda():
            p = rt.pdb200_wd_progress_ptr()
            if p:
                lib.kernels.pdb200_nvl_set_progress(p)


def _all_flags():
    from ..framework import flags as F
    return F._REGISTRY


def _cuda():
    import torch
    return torch.cuda.is_available()


def _stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def task_started(group, op, numel, sync_op):
    """Register a collective.  On a GPU the native manager records a CUDA start event on the current stream, so the
    task's state ("not started" / "started, not finished" / done) is read from the device."""
    if not enabled():
        return ""
    with _lock:
        s = _seq.get(group.id, 0) + 1
        _seq[group.id] = s
        tag = f"{group.id}:{s}"
        _inflight[tag] = [group.id, s, op, int(numel), bool(sync_op), time.time(), None, False]
    rt = _rt()
    if rt:
        if _cuda() and hasattr(rt.require(), "pdb200_wd_task_begin_ev"):
            ev = rt.pdb200_wd_task_begin_ev(int(group.id), int(s), op.encode(), int(numel), int(bool(sync_op)), _stream())
            _inflight[tag][7] = bool(ev)
        else:
            rt.pdb200_wd_task_begin(int(group.id), int(s), op.encode(), int(numel), int(bool(sync_op)))
    _sweep()
    return tag


def task_enqueued(tag, work=None):
    """The collective has been handed to the backend: remember its work handle so that an async task nobody waits on
    can still be retired (``_sweep``)."""
    rec = _inflight.get(tag) if tag else None
    if rec is not None:
        rec[6] = work


def _sweep():
    """Retire async tasks whose backend work handle reports completion (nobody has to call wait() on them)."""
    done = []
    with _lock:
        for tag, rec in _inflight.items():
            w = rec[6]
            if w is not None and not rec[4]:
                try:
                    if w.is_completed():
                        done.append(tag)
                except Exception:  # noqa: BLE001
                    pass
    for tag in done:
        task_finished(tag, device_done=True)


def task_finished(tag, device_done=False):
    """Called when ``wait()`` has made the current stream depend on the collective (or, ``device_done``, when the work
    handle already reports completion).  With device events the end event is recorded on the current stream now and
    the native manager retires the task once the GPU has passed it - so a collective that hangs on the device stays
    "started, not finished" although the host returned from wait()."""
    if not tag:
        return
    with _lock:
        rec = _inflight.pop(tag, None)
        if rec:
            _last_done[rec[0]] = rec
    rt = _rt()
    if rt and rec:
        if rec[7] and not device_done:
            rt.pdb200_wd_task_end_ev(int(rec[0]), int(rec[1]), _stream())
        else:
            rt.pdb200_wd_task_end(int(rec[0]), int(rec[1]))


def dump():
    """Trace lines for tasks still in flight (oldest first) and the last finished task per group."""
    _sweep()
    now = time.time()
    lines = []
    with _lock:
        for tag, (gid, s, op, numel, sync, t0, _w, _e) in sorted(_inflight.items(), key=lambda kv: kv[1][5]):
            lines.append(f"[in-flight] group={gid} ranks={_groups.get(gid)} seq={s} op={op} numel={numel} sync={sync} "
                         f"age={now - t0:.1f}s")
        for gid, (g, s, op, numel, sync, t0, _w, _e) in _last_done.items():
            lines.append(f"[last-done] group={gid} seq={s} op={op} numel={numel}")
    return lines


def shutdown():
    rt = _rt()
    if rt and enabled():
        rt.pdb200_wd_stop()
    with _lock:
        _inflight.clear()
        _seq.clear()
        _last_done.clear()
