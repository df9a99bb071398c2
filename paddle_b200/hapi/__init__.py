"""paddle.Model high-level API (reference: python/paddle/hapi/model.py, callbacks.py, model_summary.py) - the
dygraph subset: prepare / fit / evaluate / predict / save / load / train_batch / eval_batch with callbacks."""
from __future__ import annotations

import os
import time

import numpy as np
import torch

from .callbacks import Callback, EarlyStopping, LRScheduler, ModelCheckpoint, ProgBarLogger  # noqa: F401


def summary(net, input_size=None, dtypes=None, input=None):
    total = sum(p.numel() for p in net.parameters())
    trainable = sum(p.numel() for p in net.parameters() if p.requires_grad)
    rows = [f"{n:<50}{type(m).__name__:<24}{sum(p.numel() for p in m.parameters(include_sublayers=False) if True):>14,}"
            for n, m in net.named_sublayers() if len(list(m.children())) == 0]
    print("\n".join(rows))
    print(f"Total params: {total:,}\nTrainable params: {trainable:,}\nNon-trainable params: {total - trainable:,}")
    return {"total_params": total, "trainable_params": trainable}


class Model:
    def __init__(self, network, inputs=None, labels=None):
        self.network = network
        self._inputs, self._labels = inputs, labels
        self._optimizer = self._loss = None
        self._metrics = []
        self._scaler = None
        self._amp_level = "O0"
        self.stop_training = False

    def prepare(self, optimizer=None, loss=None, metrics=None, amp_configs=None):
        self._optimizer, self._loss = optimizer, loss
        self._metrics = metrics if isinstance(metrics, (list, tuple)) else ([metrics] if metrics else [])
        if amp_configs:
            from paddle_b200 import amp
            level = amp_configs if isinstance(amp_configs, str) else amp_configs.get("level", "O1")
            self._amp_level = level
            if level != "O0":
                self._scaler = amp.GradScaler(init_loss_scaling=(amp_configs.get("init_loss_scaling", 2.0 ** 15)
                                                                 if isinstance(amp_configs, dict) else 2.0 ** 15))

    def _to_list(self, x):
        return list(x) if isinstance(x, (list, tuple)) else [x]

    def train_batch(self, inputs, labels=None, update=True):
        self.network.train()
        inputs, labels = self._to_list(inputs), self._to_list(labels) if labels is not None else []
        from paddle_b200 import amp
        with amp.auto_cast(enable=self._amp_level != "O0", level=self._amp_level if self._amp_level != "O0" else "O1",
                           dtype="bfloat16"):
            outs = self._to_list(self.network(*inputs))
            loss = self._loss(*outs, *labels) if self._loss else outs[0]
        if self._scaler:
            self._scaler.scale(loss).backward()
            if update:
                self._scaler.step(self._optimizer)
                self._scaler.update()
                self._optimizer.clear_grad()
        else:
            loss.backward()
            if update:
                self._optimizer.step()
                self._optimizer.clear_grad()
        res = [float(loss.detach())]
        ms = []
        for m in self._metrics:
            ms.append(m.update(*self._to_list(m.compute(*outs, *labels))))
        return (res, ms) if ms else res

    @torch.no_grad()
    def eval_batch(self, inputs, labels=None):
        self.network.eval()
        inputs, labels = self._to_list(inputs), self._to_list(labels) if labels is not None else []
        outs = self._to_list(self.network(*inputs))
        res = [float(self._loss(*outs, *labels))] if self._loss else []
        ms = [m.update(*self._to_list(m.compute(*outs, *labels))) for m in self._metrics]
        return (res, ms) if ms else res

    @torch.no_grad()
    def predict_batch(self, inputs):
        self.network.eval()
        return [o.cpu().numpy() for o in self._to_list(self.network(*self._to_list(inputs)))]

    def _loader(self, data, batch_size, shuffle, num_workers, drop_last=False):
        from paddle_b200.io import DataLoader
        if data is None or isinstance(data, DataLoader):
            return data
        return DataLoader(data, batch_size=batch_size, shuffle=shuffle, num_workers=num_workers, drop_last=drop_last)

    def _split(self, batch):
        batch = self._to_list(batch)
        n_in = len(self._to_list(self._inputs)) if self._inputs is not None else max(len(batch) - 1, 1)
        return batch[:n_in], batch[n_in:]

    def fit(self, train_data=None, eval_data=None, batch_size=1, epochs=1, eval_freq=1, log_freq=10, save_dir=None,
            save_freq=1, verbose=2, drop_last=False, shuffle=True, num_workers=0, callbacks=None, accumulate_grad_batches=1,
            num_iters=None):
        loader = self._loader(train_data, batch_size, shuffle, num_workers, drop_last)
        eval_loader = self._loader(eval_data, batch_size, False, num_workers)
        cbs = list(callbacks or [])
        if verbose:
            cbs.append(ProgBarLogger(log_freq, verbose))
        if save_dir:
            cbs.append(ModelCheckpoint(save_freq, save_dir))
        try:
            steps = len(loader)
        except TypeError:
            steps = None
        metric_names = ["loss"] + [n for m in self._metrics for n in (m.name() if isinstance(m.name(), (list, tuple)) else [m.name()])]
        self._fit_has_eval = eval_loader is not None
        for cb in cbs:
            cb.set_model(self)
            cb.set_params({"batch_size": batch_size, "epochs": epochs, "steps": steps, "verbose": verbose, "metrics": metric_names})
            cb.on_train_begin()
        it = 0
        for epoch in range(epochs):
            for m in self._metrics:
                m.reset()
            for cb in cbs:
                cb.on_epoch_begin(epoch)
            logs = {}
            for step, batch in enumerate(loader):
                for cb in cbs:
                    cb.on_train_batch_begin(step, logs)
                ins, labs = self._split(batch)
                out = self.train_batch(ins, labs, update=(step + 1) % accumulate_grad_batches == 0)
                logs = {"loss": out[0] if isinstance(out, tuple) else out, "step": step}
                for cb in cbs:
                    cb.on_train_batch_end(step, logs)
                it += 1
                if num_iters is not None and it >= num_iters:
                    break
            for cb in cbs:
                cb.on_epoch_end(epoch, logs)
            if eval_loader is not None and (epoch + 1) % eval_freq == 0:
                logs.update(self.evaluate(eval_loader, verbose=0, callbacks=cbs, _from_fit=True))
            if self.stop_training or (num_iters is not None and it >= num_iters):
                break
        for cb in cbs:
            cb.on_train_end()

    def evaluate(self, eval_data, batch_size=1, log_freq=10, verbose=2, num_workers=0, callbacks=None, num_iters=None,
                 _from_fit=False):
        loader = self._loader(eval_data, batch_size, False, num_workers)
        cbs = list(callbacks or [])
        for cb in cbs:
            if not _from_fit:
                cb.set_model(self)
            cb.on_eval_begin({})
        for m in self._metrics:
            m.reset()
        losses = []
        for i, batch in enumerate(loader):
            ins, labs = self._split(batch)
            out = self.eval_batch(ins, labs)
            l = out[0] if isinstance(out, tuple) else out
            if l:
                losses.append(l[0])
            if num_iters is not None and i + 1 >= num_iters:
                break
        res = {"loss": [float(np.mean(losses))] if losses else []}
        for m in self._metrics:
            names, vals = m.name(), m.accumulate()
            names = names if isinstance(names, (list, tuple)) else [names]
            vals = vals if isinstance(vals, (list, tuple)) else [vals]
            if len(names) == 1:
                res[names[0]] = vals[0] if len(vals) == 1 else list(vals)
            else:
                res.update(dict(zip(names, vals)))
        for cb in cbs:
            cb.on_eval_end(res)
        return res

    def predict(self, test_data, batch_size=1, num_workers=0, stack_outputs=False, verbose=1, callbacks=None):
        loader = self._loader(test_data, batch_size, False, num_workers)
        outs = []
        for batch in loader:
            ins, _ = self._split(batch)
            outs.append(self.predict_batch(ins))
        outs = list(zip(*outs))
        return [np.vstack(o) for o in outs] if stack_outputs else [list(o) for o in outs]

    def save(self, path, training=True):
        from paddle_b200.framework.io import save
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        save(self.network.state_dict(), path + ".pdparams")
        if training and self._optimizer is not None:
            save(self._optimizer.state_dict(), path + ".pdopt")

    def load(self, path, skip_mismatch=False, reset_optimizer=False):
        from paddle_b200.framework.io import load
        self.network.set_state_dict(load(path + ".pdparams"))
        if not reset_optimizer and self._optimizer is not None and os.path.exists(path + ".pdopt"):
            self._optimizer.set_state_dict(load(path + ".pdopt"))

    def parameters(self, *a, **k):
        return self.network.parameters(*a, **k)

    def summary(self, input_size=None, dtype=None):
        return summary(self.network, input_size, dtype)
