"""History of finished trials (reference: python/paddle/distributed/auto_tuner/recorder.py)."""
from __future__ import annotations

import csv


class HistoryRecorder:
    def __init__(self, metric="tokens_per_second", direction="Maximize"):
        self.history = []
        self.metric, self.direction = metric, direction

    def add_cfg(self, **cfg):
        self.history.append(dict(cfg))

    def sort_metric(self):
        sign = -1.0 if self.direction == "Maximize" else 1.0
        ok = [h for h in self.history if h.get(self.metric) is not None and not h.get("error")]
        ok.sort(key=lambda h: sign * float(h[self.metric]))
        return ok

    def get_best(self):
        ranked = self.sort_metric()
        return (ranked[0], False) if ranked else (None, True)

    def store_history(self, path="./history.csv"):
        keys = sorted({k for h in self.history for k in h})
        with open(path, "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=keys)
            w.writeheader()
            for h in self.history:
                w.writerow(h)

    def load_history(self, path="./history.csv"):
        with open(path, newline="") as f:
            self.history = [dict(r) for r in csv.DictReader(f)]
        return self.history
