"""AutoTuner (reference: python/paddle/distributed/auto_tuner/tuner.py:21-70 + the trial loop of
python/paddle/distributed/launch/main.py --auto_tuner_json).

``AutoTuner(cfg).search_once()`` hands out the next un-pruned configuration; ``tune(run_trial)`` drives the loop:
run_trial(cfg) -> dict with the metric (or ``{"error": "oom"}``), every result is recorded and feeds the pruning of
later candidates.  ``launch_trial`` runs a training script under ``paddle.distributed.launch`` with the candidate
exported as ``PADDLE_AUTO_TUNER_CFG`` (JSON) and reads the metric from the last JSON line of the worker log."""
from __future__ import annotations

import json
import os
import subprocess
import sys

from .recorder import HistoryRecorder
from .search import ALGOS, GridSearch  # noqa: F401
from .utils import default_candidates


class AutoTuner:
    def __init__(self, tuner_cfg):
        self.cur_task_id = 1
        self.task_limit = tuner_cfg.get("task_limit", 100)
        tuner_cfg = dict(tuner_cfg)
        tuner_cfg["candidates"] = default_candidates(tuner_cfg)
        self.tuner_cfg = tuner_cfg
        algo = tuner_cfg.get("search_algo", "grid")
        algo = algo.get("name", "grid") if isinstance(algo, dict) else algo
        if algo not in ALGOS:
            raise ValueError(f"search_algo {algo!r}: choose from {sorted(ALGOS)}")
        self.algo = ALGOS[algo](tuner_cfg)
        self.recorder = HistoryRecorder(tuner_cfg.get("metric_cfg", {}).get("name", "tokens_per_second"),
                                        tuner_cfg.get("metric_cfg", {}).get("OptimizationDirection", "Maximize"))
        self.history_cfgs = self.recorder.history

    def search_once(self):
        if self.cur_task_id > self.task_limit:
            return None
        cfg = self.algo.search_once(self.history_cfgs)
        if cfg is not None:
            self.cur_task_id += 1
        return cfg

    def add_cfg(self, cfg):
        self.recorder.add_cfg(**cfg)

    def tune(self, run_trial):
        while True:
            cfg = self.search_once()
            if cfg is None:
                break
            res = run_trial(cfg) or {}
            cfg.update(res)
            self.add_cfg(cfg)
        return self.recorder.get_best()[0]


def launch_trial(script, cfg, nproc, log_dir="./auto_tuner_log", extra_args=(), timeout=3600):
    env = dict(os.environ, PADDLE_AUTO_TUNER_CFG=json.dumps(cfg))
    d = os.path.join(log_dir, "_".join(f"{k[:2]}{int(v) if not isinstance(v, bool) else int(v)}" for k, v in sorted(cfg.items())))
    cmd = [sys.executable, "-m", "paddle_b200.distributed.launch", "--nproc_per_node", str(nproc), "--log_dir", d, script,
           *extra_args]
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    except subprocess.TimeoutExpired:
        return {"error": "timeout"}
    text = r.stdout + "\n" + r.stderr
    log0 = os.path.join(d, "workerlog.0")
    if os.path.exists(log0):
        text += open(log0, errors="replace").read()
    if "out of memory" in text.lower():
        return {"error": "oom"}
    if r.returncode != 0:
        return {"error": f"rc={r.returncode}"}
    for line in reversed(text.splitlines()):
        line = line.strip()
        if line.startswith("{") and line.endswith("}"):
            try:
                return json.loads(line)
            except ValueError:
                continue
    return {"error": "no metric line"}
