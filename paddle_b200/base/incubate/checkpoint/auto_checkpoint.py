"""Epoch-range auto checkpoint (reference: python/paddle/base/incubate/checkpoint/auto_checkpoint.py:71-687 and
checkpoint_saver.py).

The reference wraps a static-graph ``Executor`` loop: ``for epoch in train_epoch_range(max_epoch): ...`` saves the
program's persistables to HDFS every ``save_checkpoint_inter`` seconds and, after a restart of the same job id,
resumes the range at the first unfinished epoch.  The dygraph equivalent here keeps the contract and the
environment variables and replaces "persistables of the program" by objects registered with ``register`` (anything
with ``state_dict`` / ``set_state_dict``: layers, optimizers, LR schedulers, GradScaler):

    PADDLE_RUNNING_ENV=PADDLE_EDL_AUTO_CHECKPOINT    enables it (otherwise the range is a plain ``range``)
    PADDLE_JOB_ID                                    names the job (checkpoint directory key)
    PADDLE_EDL_HDFS_CHECKPOINT_PATH                  root directory (any path LocalFS can write; HDFS is out of scope)
    PADDLE_EDL_SAVE_CHECKPOINT_INTER                 seconds between saves (default 900; 0 = every epoch)
    PADDLE_TRAINER_ID                                only trainer 0 writes
Checkpoints are written to ``<root>/<job>/<range name>/epoch_<n>.tmp`` and renamed, so a crash mid-write never
leaves a half checkpoint as the latest one; the two most recent are kept."""
from __future__ import annotations

import json
import os
import shutil
import time

_registry = {}
g_train_epoch_range = None


def _enabled():
    return os.environ.get("PADDLE_RUNNING_ENV") == "PADDLE_EDL_AUTO_CHECKPOINT" and \
        bool(os.environ.get("PADDLE_EDL_HDFS_CHECKPOINT_PATH")) and bool(os.environ.get("PADDLE_JOB_ID"))


def register(name, obj):
    """Objects whose state travels with the checkpoint (must offer state_dict() and set_state_dict())."""
    _registry[name] = obj


def unregister_all():
    _registry.clear()


class TrainEpochRange:
    def __init__(self, max_epoch_num, name="default", save_checkpoint_inter=None):
        self._max = int(max_epoch_num)
        self.name = name
        self._root = os.path.join(os.environ["PADDLE_EDL_HDFS_CHECKPOINT_PATH"], os.environ["PADDLE_JOB_ID"], name)
        self._inter = float(os.environ.get("PADDLE_EDL_SAVE_CHECKPOINT_INTER", "900")) if save_checkpoint_inter is None \
            else float(save_checkpoint_inter)
        self._trainer = int(os.environ.get("PADDLE_TRAINER_ID", "0"))
        self._last_save = time.time()
        self._epoch_no = -1
        self.restored_from = None
        os.makedirs(self._root, exist_ok=True)
        self._restore()

    # ---- persistence
    def _ckpts(self):
        out = []
        for d in os.listdir(self._root):
            if d.startswith("epoch_") and not d.endswith(".tmp") and os.path.exists(os.path.join(self._root, d, "meta.json")):
                out.append(int(d.split("_")[1]))
        return sorted(out)

    def _restore(self):
        from paddle_b200.framework.io import load
        have = self._ckpts()
        if not have:
            return
        e = have[-1]
        d = os.path.join(self._root, f"epoch_{e}")
        for name, obj in _registry.items():
            f = os.path.join(d, name + ".pdstate")
            if os.path.exists(f):
                obj.set_state_dict(load(f))
        self._epoch_no = e
        self.restored_from = d

    def save_checkpoint(self):
        if self._trainer != 0:
            return
        from paddle_b200.framework.io import save
        d = os.path.join(self._root, f"epoch_{self._epoch_no}")
        tmp = d + ".tmp"
        shutil.rmtree(tmp, ignore_errors=True)
        os.makedirs(tmp)
        for name, obj in _registry.items():
            save(obj.state_dict(), os.path.join(tmp, name + ".pdstate"))
        with open(os.path.join(tmp, "meta.json"), "w") as f:
            json.dump({"epoch_no": self._epoch_no, "max_epoch_num": self._max, "name": self.name, "time": time.time()}, f)
        shutil.rmtree(d, ignore_errors=True)
        os.rename(tmp, d)
        for old in self._ckpts()[:-2]:
            shutil.rmtree(os.path.join(self._root, f"epoch_{old}"), ignore_errors=True)
        self._last_save = time.time()

    # ---- iteration
    def get(self):
        return self._epoch_no

    def next(self):
        start = self._epoch_no + 1
        for i in range(start, self._max):
            yield i
            self._epoch_no = i
            if self._inter <= 0 or time.time() - self._last_save >= self._inter or i == self._max - 1:
                self.save_checkpoint()


def train_epoch_range(max_epoch_num, save_checkpoint_inter=None, name="default"):
    """``for epoch in train_epoch_range(10): ...`` - resumes at the first unfinished epoch after a restart."""
    global g_train_epoch_range
    if not _enabled():
        yield from range(max_epoch_num)
        return
    g_train_epoch_range = TrainEpochRange(max_epoch_num, name, save_checkpoint_inter)
    try:
        yield from g_train_epoch_range.next()
    finally:
        g_train_epoch_range = None
