"""Fault injection for failure-detection tests (the reference has no injector - its elastic tests kill
processes from the outside, test/collective/fleet/test_fleet_elastic_manager.py).

``FAULT_INJECT=<rank>:<step>:<kind>[:<restart>]`` arms one fault in the process whose global rank matches, when its
optimizer reaches ``step`` (1-based count of ``Optimizer.step`` calls), on launcher attempt ``restart`` (default 0, so
the restarted job runs clean):

* ``exit``  - the process dies at once with exit code 17 (a crashed worker): the launcher's watcher sees it and, with
  ``--max_restart``, relaunches the pod;
* ``hang``  - the process stops issuing work inside the step, holding a collective record open: the peers block in
  their next collective, the communication watchdog reports the timeout (and aborts when FLAGS_comm_watchdog_abort);
* ``raise`` - a RuntimeError is raised from ``step`` (a Python-level failure).
"""
from __future__ import annotations

import os
import sys
import time

_spec = [None, False]      # parsed spec, already fired
_count = [0]


def _parse():
    if _spec[0] is None:
        raw = os.environ.get("FAULT_INJECT", "")
        _spec[0] = ()
        if raw:
            parts = raw.split(":")
            try:
                rank, step, kind = int(parts[0]), int(parts[1]), parts[2]
                restart = int(parts[3]) if len(parts) > 3 else 0
                if kind not in ("exit", "hang", "raise"):
                    raise ValueError(kind)
                _spec[0] = (rank, step, kind, restart)
            except (ValueError, IndexError):
                sys.stderr.write(f"[fault] ignoring malformed FAULT_INJECT={raw!r} (want rank:step:kind[:restart])\n")
    return _spec[0]


def armed():
    return bool(_parse())


def tick(where="optimizer.step"):
    """Count one step; fire the configured fault when rank / step / launcher attempt match."""
    spec = _parse()
    if not spec or _spec[1]:
        return
    _count[0] += 1
    rank, step, kind, restart = spec
    my_rank = int(os.environ.get("PADDLE_TRAINER_ID", os.environ.get("RANK", "0")))
    attempt = int(os.environ.get("PADDLE_RESTART_COUNT", "0"))
    if my_rank != rank or _count[0] != step or attempt != restart:
        return
    _spec[1] = True
    sys.stderr.write(f"[fault] rank {my_rank}: injecting '{kind}' at {where} #{step} (launcher attempt {attempt})\n")
    sys.stderr.flush()
    if kind == "exit":
        os._exit(17)
    if kind == "raise":
        raise RuntimeError(f"injected fault at {where} #{step}")
    from . import watchdog
    from . import collective as C
    g = C._get_default_group() if C._state.default_group is not None else None
    if g is not None:
        watchdog.task_started(g, "injected_hang", 0, True)       # an open record the watchdog can time out
    while True:
        time.sleep(1.0)
