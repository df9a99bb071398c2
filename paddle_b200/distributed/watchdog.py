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
        rt.pdb200_wd_start(int(rank), int(flag("comm_watchdog_timeout_ms")), 10000)


def task_started(group, op, numel, sync_op):
    if not enabled():
        return ""
    with _lock:
        s = _seq.get(group.id, 0) + 1
        _seq[group.id] = s
        tag = f"{group.id}:{s}"
        _inflight[tag] = (group.id, s, op, int(numel), bool(sync_op), time.time())
    rt = _rt()
    if rt:
        rt.pdb200_wd_task_begin(int(group.id), int(s), op.encode(), int(numel), int(bool(sync_op)))
    return tag


def task_finished(tag):
    if not tag:
        return
    with _lock:
        rec = _inflight.pop(tag, None)
        if rec:
            _last_done[rec[0]] = rec
    rt = _rt()
    if rt and rec:
        rt.pdb200_wd_task_end(int(rec[0]), int(rec[1]))


def dump():
    """Trace lines for tasks still in flight (oldest first) and the last finished task per group."""
    now = time.time()
    lines = []
    with _lock:
        for tag, (gid, s, op, numel, sync, t0) in sorted(_inflight.items(), key=lambda kv: kv[1][5]):
            lines.append(f"[in-flight] group={gid} ranks={_groups.get(gid)} seq={s} op={op} numel={numel} sync={sync} "
                         f"age={now - t0:.1f}s")
        for gid, (g, s, op, numel, sync, t0) in _last_done.items():
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
