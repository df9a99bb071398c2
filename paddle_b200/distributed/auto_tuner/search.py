"""Search strategies (reference: python/paddle/distributed/auto_tuner/search.py: GridSearch :48, DpEstimationSearch :96,
GBSSearch :123, CustomizeSearch :144).  ``CostModelSearch`` is this framework's addition: the grid ordered by the
analytic step-time estimate, so that the first few trials already are the likely winners."""
from __future__ import annotations

from .prune import should_prune
from .utils import search_all


class SearchAlgo:
    def __init__(self, tuner_cfg):
        self.tuner_cfg = tuner_cfg
        self.all_tasks = self.candidates(tuner_cfg)
        self.idx = 0

    def candidates(self, tuner_cfg):
        return search_all(tuner_cfg)

    def search_once(self, history):
        while self.idx < len(self.all_tasks):
            cur = self.all_tasks[self.idx]
            self.idx += 1
            if not should_prune(self.tuner_cfg, cur, history):
                return dict(cur)
        return None


class GridSearch(SearchAlgo):
    pass


class CostModelSearch(SearchAlgo):
    def candidates(self, tuner_cfg):
        from .cost_model import estimate_step_ms
        m = tuner_cfg.get("model_cfg", {})
        tasks = search_all(tuner_cfg)
        if all(k in m for k in ("hidden_size", "num_layers", "vocab_size")):
            tasks.sort(key=lambda c: estimate_step_ms(c, m))
        return tasks


class DpEstimationSearch(SearchAlgo):
    """Tune on a slice of the machine: the data-parallel degree is pinned to 1 and the device count divided accordingly, the
    winner is then replicated dp times (data parallelism does not change the per-GPU picture)."""

    def candidates(self, tuner_cfg):
        cfg = dict(tuner_cfg)
        dp = int(tuner_cfg.get("dp_estimation_degree", 0)) or max(d for d in tuner_cfg["candidates"]["dp_degree"])
        n = int(tuner_cfg["num_gpus"])
        assert n % dp == 0
        cfg["num_gpus"] = n // dp
        cand = dict(tuner_cfg["candidates"])
        cand["dp_degree"] = [1]
        cand = {k: ([v for v in vs if (n // dp) % v == 0] if k.endswith("_degree") else vs) for k, vs in cand.items()}
        cfg["candidates"] = cand
        m = dict(tuner_cfg.get("model_cfg", {}))
        if m.get("global_batch_size"):
            m["global_batch_size"] = max(m["global_batch_size"] // dp, 1)
        cfg["model_cfg"] = m
        self.tuner_cfg = cfg
        self.dp = dp
        return search_all(cfg)


class GBSSearch(SearchAlgo):
    """The parallel layout is fixed; the global batch size doubles from ``gbs_start`` until a trial fails."""

    def candidates(self, tuner_cfg):
        base = search_all(tuner_cfg)[:1]
        if not base:
            return []
        start = int(tuner_cfg.get("gbs_start", tuner_cfg.get("model_cfg", {}).get("global_batch_size", 8)))
        out = []
        g = start
        for _ in range(int(tuner_cfg.get("gbs_trials", 8))):
            out.append(dict(base[0], global_batch_size=g))
            g *= 2
        return out

    def search_once(self, history):
        if history and history[-1].get("error"):
            return None                       # the previous size no longer fits: stop growing
        if self.idx < len(self.all_tasks):
            self.idx += 1
            return dict(self.all_tasks[self.idx - 1])
        return None


class CustomizeSearch(SearchAlgo):
    """The user lists the configurations (``tuner_cfg["configs"]``); pruning still applies."""

    def candidates(self, tuner_cfg):
        return [dict(c) for c in tuner_cfg.get("configs", [])]


ALGOS = {"grid": GridSearch, "cost_model": CostModelSearch, "dp_estimation": DpEstimationSearch, "gbs": GBSSearch,
         "customize": CustomizeSearch}
