"""Multi-process tests on the gloo backend (the reference's fake backend for no-GPU CI): the driver
only checks exit codes, assertions live in tests/dist_workers/*.py."""
import pytest

from util import run_dist

pytestmark = pytest.mark.dist


def test_collectives_gloo():
    run_dist("collectives.py", 2)


def test_data_parallel_reducer_parity():
    run_dist("data_parallel.py", 2)


def test_tp_layers_shard_vs_whole():
    run_dist("tp_layers.py", 2)
