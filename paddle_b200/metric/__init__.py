"""paddle.metric (reference: python/paddle/metric/metrics.py)."""
from __future__ import annotations

import numpy as np
import torch


class Metric:
    def reset(self):
        raise NotImplementedError

    def update(self, *args):
        raise NotImplementedError

    def accumulate(self):
        raise NotImplementedError

    def name(self):
        return self._name

    def compute(self, *args):
        return args


class Accuracy(Metric):
    def __init__(self, topk=(1,), name=None, *args, **kwargs):
        self.topk = tuple(topk)
        self.maxk = max(self.topk)
        base = name or "acc"
        # one name per k, as the reference reports them (metrics.py Accuracy._init_name)
        self._name = [f"{base}_top{k}" for k in self.topk] if self.maxk != 1 else [base]
        self.reset()

    def name(self):
        return self._name

    def compute(self, pred, label, *args):
        idx = torch.argsort(pred, dim=-1, descending=True)[..., : self.maxk]
        if label.dim() == pred.dim() and label.shape[-1] != 1:
            label = label.argmax(-1, keepdim=True)
        elif label.dim() < pred.dim():
            label = label.unsqueeze(-1)
        return (idx == label).float()

    def update(self, correct, *args):
        correct = correct.detach().cpu().numpy() if isinstance(correct, torch.Tensor) else np.asarray(correct)
        accs = []
        for i, k in enumerate(self.topk):
            num = correct[..., :k].sum()
            self.total[i] += num
            self.count[i] += correct.shape[0] if correct.ndim > 1 else 1
            accs.append(float(num) / max(correct.shape[0], 1))
        return accs[0] if len(accs) == 1 else accs

    def reset(self):
        self.total = [0.0] * len(self.topk)
        self.count = [0] * len(self.topk)

    def accumulate(self):
        res = [t / c if c else 0.0 for t, c in zip(self.total, self.count)]
        return res[0] if len(res) == 1 else res


class Precision(Metric):
    def __init__(self, name="precision", *a, **k):
        self._name = name
        self.reset()

    def update(self, preds, labels):
        p = (np.asarray(preds.cpu() if isinstance(preds, torch.Tensor) else preds) > 0.5).reshape(-1)
        l = np.asarray(labels.cpu() if isinstance(labels, torch.Tensor) else labels).reshape(-1)
        self.tp += int(((p == 1) & (l == 1)).sum())
        self.fp += int(((p == 1) & (l == 0)).sum())

    def reset(self):
        self.tp = self.fp = 0

    def accumulate(self):
        return self.tp / (self.tp + self.fp) if self.tp + self.fp else 0.0


class Recall(Metric):
    def __init__(self, name="recall", *a, **k):
        self._name = name
        self.reset()

    def update(self, preds, labels):
        p = (np.asarray(preds.cpu() if isinstance(preds, torch.Tensor) else preds) > 0.5).reshape(-1)
        l = np.asarray(labels.cpu() if isinstance(labels, torch.Tensor) else labels).reshape(-1)
        self.tp += int(((p == 1) & (l == 1)).sum())
        self.fn += int(((p == 0) & (l == 1)).sum())

    def reset(self):
        self.tp = self.fn = 0

    def accumulate(self):
        return self.tp / (self.tp + self.fn) if self.tp + self.fn else 0.0


def accuracy(input, label, k=1, correct=None, total=None, name=None):
    idx = torch.topk(input, k, dim=-1).indices
    lab = label.reshape(-1, 1)
    return (idx == lab).any(-1).float().mean()


class Auc(Metric):
    """Area under the ROC / PR curve from thresholded histograms (reference: metric/metrics.py Auc)."""

    def __init__(self, curve="ROC", num_thresholds=4095, name="auc", *args, **kwargs):
        self._curve, self._n, self._name = curve, num_thresholds, name
        self.reset()

    def reset(self):
        import numpy as np
        self._pos = np.zeros(self._n + 1, dtype=np.int64)
        self._neg = np.zeros(self._n + 1, dtype=np.int64)

    def update(self, preds, labels):
        import numpy as np
        preds = preds.detach().cpu().numpy() if hasattr(preds, "detach") else np.asarray(preds)
        labels = labels.detach().cpu().numpy() if hasattr(labels, "detach") else np.asarray(labels)
        p1 = preds[:, 1] if preds.ndim == 2 and preds.shape[1] == 2 else preds.reshape(-1)
        bins = np.clip((p1 * self._n).astype(np.int64), 0, self._n)
        lab = labels.reshape(-1).astype(bool)
        np.add.at(self._pos, bins[lab], 1)
        np.add.at(self._neg, bins[~lab], 1)

    def accumulate(self):
        tot_pos = tot_neg = 0.0
        auc = 0.0
        for i in range(self._n, -1, -1):
            p_prev, n_prev = tot_pos, tot_neg
            tot_pos += self._pos[i]
            tot_neg += self._neg[i]
            auc += (tot_neg - n_prev) * (tot_pos + p_prev) / 2.0
        return auc / tot_pos / tot_neg if tot_pos > 0 and tot_neg > 0 else 0.0

    def name(self):
        return self._name
