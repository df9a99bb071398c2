"""paddle.distributed.io (reference: python/paddle/distributed/io.py): persistables of a distributed job.

The reference functions take a static-graph ``Executor`` / ``Program``; in this dygraph-only framework the same names
accept a Layer (or a state dict) in place of the program and write / read one ``.pdparams`` file on rank 0."""
from __future__ import annotations

import os


def _state(obj):
    return obj.state_dict() if hasattr(obj, "state_dict") else obj


def save_persistables(executor, dirname, main_program=None, filename=None):
    from paddle_b200.distributed import collective as C
    from paddle_b200.framework.io import save
    if main_program is None:
        raise ValueError("pass the Layer (or its state dict) as main_program: there is no static Program here")
    if C.get_rank() == 0:
        os.makedirs(dirname, exist_ok=True)
        save(_state(main_program), os.path.join(dirname, filename or "persistables.pdparams"))


def load_persistables(executor, dirname, main_program=None, filename=None):
    from paddle_b200.framework.io import load
    sd = load(os.path.join(dirname, filename or "persistables.pdparams"))
    if main_program is not None and hasattr(main_program, "set_state_dict"):
        main_program.set_state_dict(sd)
    return sd


def is_persistable(var):
    return bool(getattr(var, "persistable", getattr(var, "requires_grad", False) is not None))
