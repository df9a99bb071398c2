"""Failure detection end to end: FAULT_INJECT kills or hangs one rank; the launcher's watcher notices,
restarts the pod (--max_restart) and the second attempt - where the injector stays quiet - finishes.  For the hang
the communication watchdog of the blocked rank must print its trace line and abort first."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(tmp_path, fault, max_restart=1, timeout=240):
    env = dict(os.environ)
    env.update({"PYTHONPATH": ROOT + os.pathsep + env.get("PYTHONPATH", ""), "PADDLE_DISTRI_BACKEND": "gloo",
                "CUDA_VISIBLE_DEVICES": "", "OMP_NUM_THREADS": "2", "FAULT_INJECT": fault})
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    log_dir = str(tmp_path / "log")
    cmd = [sys.executable, "-m", "paddle_b200.distributed.launch", "--nproc_per_node", "2", "--max_restart", str(max_restart),
           "--elastic_level", "0", "--log_dir", log_dir, os.path.join(ROOT, "tests", "dist_workers", "fault_worker.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    logs = ""
    if os.path.isdir(log_dir):
        for f in sorted(os.listdir(log_dir)):
            with open(os.path.join(log_dir, f), errors="replace") as fh:
                logs += f"\n===== {f} =====\n" + fh.read()
    return r, logs


@pytest.mark.dist
def test_killed_rank_is_detected_and_the_pod_restarts(tmp_path):
    r, logs = _launch(tmp_path, "1:2:exit")
    assert r.returncode == 0, r.stderr[-2000:] + logs[-3000:]
    assert "[fault] rank 1: injecting 'exit'" in logs
    assert "restarting pod (1/1)" in r.stderr
    assert "fault_worker done rank=0 attempt=1" in logs and "fault_worker done rank=1 attempt=1" in logs


@pytest.mark.dist
def test_hung_rank_trips_the_watchdog_then_the_pod_restarts(tmp_path):
    r, logs = _launch(tmp_path, "0:2:hang")
    assert r.returncode == 0, r.stderr[-2000:] + logs[-3000:]
    assert "[comm watchdog]" in logs and "timed out" in logs, logs[-3000:]
    assert "fault_worker done rank=0 attempt=1" in logs


@pytest.mark.dist
def test_without_restart_budget_the_failure_is_reported(tmp_path):
    r, logs = _launch(tmp_path, "1:2:exit", max_restart=0)
    assert r.returncode != 0
