"""Fused NVLink kernels need >= 2 GPUs of one node; skipped elsewhere (the driver's 1-GPU run included)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(script, nproc, timeout=600):
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "tests", "dist_workers", script)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, f"{script} failed\n{r.stdout[-3000:]}\n{r.stderr[-6000:]}"
    return r


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_nvl_fused_kernels_match_nccl():
    out = _run("nvl_fused.py", min(torch.cuda.device_count(), 8))
    assert "nvl_fused ok" in out.stdout
