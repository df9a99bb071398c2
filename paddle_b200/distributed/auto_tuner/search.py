"""Search strategies (reference: python/paddle/distributed/auto_tuner/search.py GridSearch)."""
from __future__ import annotations

from .prune import should_prune
from .utils import search_all


class GridSearch:
    def __init__(self, tuner_cfg):
        self.tuner_cfg = tuner_cfg
        self.all_tasks = search_all(tuner_cfg)
        self.idx = 0

    def search_once(self, history):
        while self.idx < len(self.all_tasks):
            cur = self.all_tasks[self.idx]
            self.idx += 1
            if not should_prune(self.tuner_cfg, cur, history):
                return dict(cur)
        return None
