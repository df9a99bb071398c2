import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_dist(script, nproc=2, timeout=480, extra_env=None, args=()):
    """Driver half of the worker-script pattern: launch ``script`` on ``nproc`` CPU/gloo ranks through
    paddle.distributed.launch; assertions live in the worker; only the exit code is checked here."""
    env = dict(os.environ)
    env.update({"PYTHONPATH": ROOT + os.pathsep + env.get("PYTHONPATH", ""), "PADDLE_DISTRI_BACKEND": "gloo",
                "CUDA_VISIBLE_DEVICES": "", "OMP_NUM_THREADS": "2"})
    env.update(extra_env or {})
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    log_dir = os.path.join("/tmp", f"pdb200_test_logs_{os.getpid()}_{os.path.basename(script)}")
    cmd = [sys.executable, "-m", "paddle_b200.distributed.launch", "--nproc_per_node", str(nproc), "--log_dir", log_dir,
           os.path.join(ROOT, "tests", "dist_workers", script), *args]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    if r.returncode != 0:
        logs = ""
        if os.path.isdir(log_dir):
            for f in sorted(os.listdir(log_dir)):
                with open(os.path.join(log_dir, f), errors="replace") as fh:
                    logs += f"\n===== {f} =====\n" + fh.read()[-3000:]
        raise AssertionError(f"{script} failed rc={r.returncode}\nstdout:{r.stdout[-2000:]}\nstderr:{r.stderr[-3000:]}{logs}")
    return r
