"""paddle.distributed.spawn (reference: python/paddle/distributed/spawn.py:146-618): one process per rank with the
launcher's environment contract, results through the context, a child's traceback re-raised in the parent."""
import os
import sys

import pytest

import paddle.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import spawn_targets  # noqa: E402


def test_spawn_runs_a_collective_on_every_rank(monkeypatch):
    monkeypatch.setenv("PYTHONPATH", os.pathsep.join([os.path.dirname(os.path.abspath(__file__)),
                                                      os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                      os.environ.get("PYTHONPATH", "")]))
    ctx = dist.spawn(spawn_targets.allreduce_rank, args=(2.0,), nprocs=2, backend="gloo")
    results = sorted(q.get() for q in ctx.return_queues)
    assert [r[0] for r in results] == [0, 1]
    assert [r[1] for r in results] == [(0, 2, 6.0), (1, 2, 6.0)]
    assert all(p.exitcode == 0 for p in ctx.processes)


def test_spawn_propagates_the_childs_traceback(monkeypatch):
    monkeypatch.setenv("PYTHONPATH", os.pathsep.join([os.path.dirname(os.path.abspath(__file__)),
                                                      os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                      os.environ.get("PYTHONPATH", "")]))
    with pytest.raises(Exception) as e:
        dist.spawn(spawn_targets.fails_on_rank_one, nprocs=2, backend="gloo")
    assert "Process 1 terminated" in str(e.value) and "boom from rank 1" in str(e.value)
