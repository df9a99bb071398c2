"""hapi callbacks (reference: python/paddle/hapi/callbacks.py)."""
from __future__ import annotations

import os
import time


class Callback:
    def __init__(self):
        self.model = None
        self.params = {}

    def set_params(self, params):
        self.params = params

    def set_model(self, model):
        self.model = model

    def on_train_begin(self, logs=None):
        pass

    def on_train_end(self, logs=None):
        pass

    def on_epoch_begin(self, epoch, logs=None):
        pass

    def on_epoch_end(self, epoch, logs=None):
        pass

    def on_train_batch_begin(self, step, logs=None):
        pass

    def on_train_batch_end(self, step, logs=None):
        pass

    def on_eval_begin(self, logs=None):
        pass

    def on_eval_end(self, logs=None):
        pass

    def on_eval_batch_begin(self, step, logs=None):
        pass

    def on_eval_batch_end(self, step, logs=None):
        pass

    def on_predict_begin(self, logs=None):
        pass

    def on_predict_end(self, logs=None):
        pass

    def on_predict_batch_begin(self, step, logs=None):
        pass

    def on_predict_batch_end(self, step, logs=None):
        pass


class ProgBarLogger(Callback):
    def __init__(self, log_freq=1, verbose=2):
        super().__init__()
        self.log_freq, self.verbose = log_freq, verbose
        self._t0 = time.time()

    def on_epoch_begin(self, epoch, logs=None):
        self.epoch = epoch

    def on_train_batch_end(self, step, logs=None):
        if self.verbose and step % self.log_freq == 0:
            print(f"Epoch {self.epoch} step {step}: loss {logs.get('loss')}", flush=True)


class ModelCheckpoint(Callback):
    def __init__(self, save_freq=1, save_dir=None):
        super().__init__()
        self.save_freq, self.save_dir = save_freq, save_dir

    def on_epoch_end(self, epoch, logs=None):
        if self.save_dir and (epoch + 1) % self.save_freq == 0:
            self.model.save(os.path.join(self.save_dir, str(epoch)))

    def on_train_end(self, logs=None):
        if self.save_dir:
            self.model.save(os.path.join(self.save_dir, "final"))


class LRScheduler(Callback):
    def __init__(self, by_step=True, by_epoch=False):
        super().__init__()
        self.by_step, self.by_epoch = by_step, by_epoch

    def _step(self):
        lr = getattr(self.model._optimizer, "_learning_rate", None)
        if hasattr(lr, "step"):
            lr.step()

    def on_train_batch_end(self, step, logs=None):
        if self.by_step:
            self._step()

    def on_epoch_end(self, epoch, logs=None):
        if self.by_epoch:
            self._step()


class EarlyStopping(Callback):
    def __init__(self, monitor="loss", mode="auto", patience=0, verbose=1, min_delta=0, baseline=None, save_best_model=True):
        super().__init__()
        self.monitor, self.patience, self.min_delta = monitor, patience, abs(min_delta)
        self.mode = "min" if mode == "auto" and "loss" in monitor else ("max" if mode == "auto" else mode)
        self.best, self.wait = baseline, 0

    def on_epoch_end(self, epoch, logs=None):
        # without evaluation data the training logs are all there is to monitor
        if not getattr(self.model, "_fit_has_eval", False):
            self._check(logs)

    def on_eval_end(self, logs=None):
        self._check(logs)

    def _check(self, logs):
        v = (logs or {}).get(self.monitor)
        if v is None or (isinstance(v, (list, tuple)) and not v):
            return
        v = v[0] if isinstance(v, (list, tuple)) else v
        better = self.best is None or (v < self.best - self.min_delta if self.mode == "min" else v > self.best + self.min_delta)
        if better:
            self.best, self.wait = v, 0
        else:
            self.wait += 1
            if self.wait > self.patience:
                self.model.stop_training = True


class ReduceLROnPlateau(Callback):
    """Multiply the learning rate by ``factor`` when ``monitor`` has not improved for ``patience`` evaluations
    (reference: hapi/callbacks.py ReduceLROnPlateau).  The optimizer's learning rate must be a float."""

    def __init__(self, monitor="loss", factor=0.1, patience=10, verbose=1, mode="auto", min_delta=1e-4, cooldown=0, min_lr=0):
        super().__init__()
        if factor >= 1.0:
            raise ValueError("ReduceLROnPlateau does not support a factor >= 1.0.")
        self.monitor, self.factor, self.patience, self.verbose = monitor, factor, patience, verbose
        self.min_delta, self.cooldown, self.min_lr = min_delta, cooldown, min_lr
        self.mode = mode if mode in ("min", "max") else ("max" if "acc" in monitor else "min")
        self.cooldown_counter, self.wait, self.best, self.epoch = 0, 0, None, 0

    def _better(self, value):
        if self.best is None:
            return True
        return value < self.best - self.min_delta if self.mode == "min" else value > self.best + self.min_delta

    def on_train_begin(self, logs=None):
        self.cooldown_counter, self.wait, self.best = 0, 0, None

    def on_eval_end(self, logs=None):
        value = (logs or {}).get(self.monitor)
        if value is None:
            import warnings
            warnings.warn(f"Monitor of ReduceLROnPlateau should be loss or metric name, {self.monitor} was not found")
            return
        value = float(value[0] if isinstance(value, (list, tuple)) else value)
        opt = self.model._optimizer
        if not isinstance(getattr(opt, "_learning_rate", None), (int, float)):
            import warnings
            warnings.warn("Expected learning_rate be float, ReduceLROnPlateau is skipped")
            return
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.wait = 0
        if self._better(value):
            self.best, self.wait = value, 0
        elif self.cooldown_counter <= 0:
            self.wait += 1
            if self.wait >= self.patience:
                old = float(opt.get_lr())
                if old > float(self.min_lr):
                    new = max(old * self.factor, self.min_lr)
                    opt.set_lr(new)
                    if self.verbose:
                        print(f"\nEpoch {self.epoch + 1}: ReduceLROnPlateau reducing learning rate to {new}.")
                    self.cooldown_counter, self.wait = self.cooldown, 0
        self.epoch += 1



class VisualDL(Callback):
    """Scalar logger.  The visualdl package is not installed here, so the scalars go to ``<log_dir>/scalars.jsonl``
    (one JSON object per record: tag, step, value) - the same tags the reference writes (train/loss, eval/acc ...)."""

    def __init__(self, log_dir):
        super().__init__()
        self.log_dir, self.epochs, self.steps, self.epoch = log_dir, None, None, 0
        self._train_step = 0

    def _write(self, mode, step, logs):
        import json
        os.makedirs(self.log_dir, exist_ok=True)
        with open(os.path.join(self.log_dir, "scalars.jsonl"), "a") as f:
            for k, v in (logs or {}).items():
                if k in ("step", "batch_size"):
                    continue
                v = v[0] if isinstance(v, (list, tuple)) else v
                try:
                    f.write(json.dumps({"tag": f"{mode}/{k}", "step": int(step), "value": float(v)}) + "\n")
                except (TypeError, ValueError):
                    continue

    def on_train_begin(self, logs=None):
        self.epochs, self.steps = self.params.get("epochs"), self.params.get("steps")
        self._train_step = 0

    def on_epoch_begin(self, epoch=None, logs=None):
        self.epoch = epoch

    def on_train_batch_end(self, step, logs=None):
        self._train_step += 1
        self._write("train", self._train_step, logs)

    def on_eval_end(self, logs=None):
        self._write("eval", self.epoch or 0, logs)


class WandbCallback(Callback):
    """Weights & Biases logger; needs the ``wandb`` package (not installed here: constructing it raises ImportError,
    exactly as the reference does without the package)."""

    def __init__(self, project=None, entity=None, name=None, dir=None, mode=None, job_type=None, **kwargs):
        super().__init__()
        try:
            import wandb
        except ImportError as e:
            raise ImportError("You want to use `wandb` which is not installed yet install it with `pip install wandb`") from e
        self.wandb = wandb
        self.wandb_args = dict(project=project, name=name, entity=entity, dir=dir, mode=mode, job_type=job_type, **kwargs)
        self._run = None

    @property
    def run(self):
        if self._run is None:
            self._run = self.wandb.run if self.wandb.run is not None else self.wandb.init(**self.wandb_args)
        return self._run

    def on_train_batch_end(self, step, logs=None):
        self.run.log({f"train/{k}": (v[0] if isinstance(v, (list, tuple)) else v) for k, v in (logs or {}).items()})

    def on_eval_end(self, logs=None):
        self.run.log({f"eval/{k}": (v[0] if isinstance(v, (list, tuple)) else v) for k, v in (logs or {}).items()})

    def on_train_end(self, logs=None):
        self.run.finish()
