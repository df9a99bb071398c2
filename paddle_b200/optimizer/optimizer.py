"""Optimizers (reference: python/paddle/optimizer/{optimizer,sgd,momentum,adam,adamw}.py).

Kept from the reference: constructor signatures, parameter groups, accumulator naming
``f"{param.name}_{acc}_0"`` (optimizer.py:936-937) with ``moment1/moment2/beta1_pow_acc/beta2_pow_acc/velocity``,
``master_weights`` and ``LR_Scheduler`` entries in ``state_dict`` (optimizer.py:335-355), ``multi_precision`` fp32
master weights, ``main_grad`` as the gradient source when present, ``_apply_optimize``/``_grad_clip`` hooks that the
hybrid-parallel and sharding optimizers wrap.

B200 design: on a GPU the whole step is ONE multi-tensor kernel launch (K12, csrc/kernels/optim.cu) fed by a
descriptor table; clip coefficient, loss-scale and found_inf are device scalars read by that kernel, so ``step()``
never synchronises with the host.  On CPU the same math runs as torch ops (oracle / gloo tests).
"""
from __future__ import annotations

import os
from collections import OrderedDict, defaultdict

import numpy as np
import torch

from .. import ops
from ..framework.flags import flag
from ..ops._util import dtype_code
from .lr import LRScheduler


def _grad_of(p):
    mg = getattr(p, "main_grad", None)
    return mg if mg is not None else p.grad


class Optimizer:
    _acc_names: tuple = ()
    # who adds the regulariser (the ``weight_decay`` argument, or a parameter's own ``regularizer``) to the gradient:
    #   "grad"      the base class does, after clipping (reference: optimizer.py _apply_optimize -> append_regularization_ops)
    #   "kernel"    the update rule adds ``coeff * p`` itself (L2 only; the reference fuses it the same way in momentum)
    #   "decoupled" ``weight_decay`` is not a regulariser (AdamW, Lamb); only a parameter's own regulariser reaches the gradient
    _decay_mode = "grad"
    # True when the update masks itself on the GradScaler's device flag; otherwise the base class reads the flag and
    # skips the whole step, as the reference's scaler does (python/paddle/amp/grad_scaler.py: step -> _found_inf)
    _masks_found_inf = False

    def __init__(self, learning_rate, parameters=None, weight_decay=None, grad_clip=None, name=None,
                 multi_precision=False):
        if parameters is None:
            raise AttributeError("parameters argument given to the Optimizer should not be None in dygraph mode.")
        if isinstance(parameters, torch.Tensor):
            raise TypeError("`parameters` argument given to the optimizer should be an iterable of Tensors or dicts")
        parameters = list(parameters)
        self._learning_rate = learning_rate
        self._grad_clip = grad_clip
        self._name = name
        self._multi_precision = bool(multi_precision)
        self.regularization = weight_decay
        self._default_dict = {"weight_decay": weight_decay, "grad_clip": grad_clip}
        self._param_groups = []
        if parameters and isinstance(parameters[0], dict):
            self._parameter_list = []
            for g in parameters:
                self._add_param_group(dict(g))
        else:
            self._parameter_list = parameters
            self._param_groups = [{"params": parameters}]
        self._accumulators = defaultdict(dict)       # acc name -> {param.name: tensor}
        self._master_weights = {}                     # param.name -> fp32 tensor
        self._accumulators_holder = {}
        self._step_count = 0
        self._found_inf = None                        # device fp32[1] set by GradScaler
        self._grad_scale_dev = None                   # device fp32[1] multiplied into grads inside the kernel
        self._tables = {}

    # ------------------------------------------------------------- groups / lr
    def _add_param_group(self, group):
        params = group["params"]
        group["params"] = [params] if isinstance(params, torch.Tensor) else list(params)
        for k, v in self._default_dict.items():
            group.setdefault(k, v)
        self._param_groups.append(group)
        self._parameter_list.extend(group["params"])

    def _all_params(self):
        for g in self._param_groups:
            for p in g["params"]:
                yield p, g

    def get_lr(self):
        if isinstance(self._learning_rate, LRScheduler):
            return float(self._learning_rate())
        return float(self._learning_rate)

    def set_lr(self, value):
        if isinstance(self._learning_rate, LRScheduler):
            raise RuntimeError("optimizer's learning rate can't be LRScheduler when invoke this API")
        self._learning_rate = float(value)

    def set_lr_scheduler(self, scheduler):
        self._learning_rate = scheduler

    # ------------------------------------------------------------- accumulators
    def _get_accumulator(self, name, p):
        return self._accumulators[name][p.name]

    def _add_accumulator(self, name, p, dtype=torch.float32, fill=0.0, shape=None):
        if p.name in self._accumulators[name]:
            return self._accumulators[name][p.name]
        t = torch.full(tuple(shape) if shape is not None else tuple(p.shape), fill, dtype=dtype, device=p.device)
        key = f"{p.name}_{name}_0"
        if key in self._accumulators_holder:
            t.copy_(torch.as_tensor(self._accumulators_holder.pop(key)).to(t.device, t.dtype).reshape(t.shape))
        t.name = key
        self._accumulators[name][p.name] = t
        return t

    def _use_master(self, p):
        return self._multi_precision and p.dtype in (torch.float16, torch.bfloat16)

    def _create_master_weight(self, p):
        if p.name not in self._master_weights:
            mw = p.detach().float().clone()
            holder = self._accumulators_holder.get("master_weights", {})
            if p.name in holder:
                mw.copy_(torch.as_tensor(holder[p.name]).to(mw.device).reshape(mw.shape))
            mw.name = p.name + "_fp32_master_0"
            self._master_weights[p.name] = mw
        return self._master_weights[p.name]

    def _create_accumulators(self, params):
        for p in params:
            if self._use_master(p):
                self._create_master_weight(p)

    # ------------------------------------------------------------- step
    def _collect_params_grads(self):
        out = []
        for p, _ in self._all_params():
            if not p.requires_grad and getattr(p, "main_grad", None) is None:
                continue
            g = _grad_of(p)
            if g is not None:
                out.append((p, g))
        return out

    @torch.no_grad()
    def step(self):
        if os.environ.get("FAULT_INJECT"):
            from ..distributed import fault
            fault.tick()
        params_grads = self._collect_params_grads()
        if any(hasattr(p, "placements") for p, _ in params_grads):
            # auto_parallel DistParameters: the update is shard-local arithmetic, not an op on distributed tensors
            with torch._C.DisableTorchFunctionSubclass():
                return self._apply_optimize(None, None, params_grads)
        self._apply_optimize(None, None, params_grads)

    @torch.no_grad()
    def _apply_optimize(self, loss, startup_program, params_grads, param_group_idx=0):
        folded = False
        if self._grad_clip is not None:
            if self._can_fold_clip(params_grads):
                # global-norm clip folded into the update kernel: one multi-tensor L2-norm launch, the coefficient
                # stays on the device and multiplies the gradients inside the AdamW kernel (no pass over the grads)
                clip = self._grad_clip
                if hasattr(clip, "fold_coef"):          # sharded clip: its own (collective) coefficient
                    coef = clip.fold_coef(params_grads)
                else:
                    grads = [g for p, g in params_grads if getattr(p, "need_clip", True)]
                    coef = clip.clip_coef(clip.global_norm_sq(grads))
                prev = self._grad_scale_dev
                self._grad_scale_dev = coef if prev is None else prev * coef
                folded = True
            else:
                params_grads = self._grad_clip(params_grads)
        if not params_grads:
            return
        if self._found_inf is not None and not self._masks_found_inf and bool((self._found_inf > 0).any()):
            return
        try:
            params_grads = self._regularize(params_grads)
            self._create_accumulators([p for p, _ in params_grads])
            self._update(params_grads)
        finally:
            if folded:
                self._grad_scale_dev = prev
        self._step_count += 1

    def _regularize(self, params_grads):
        """Gradients with the regulariser added, according to ``_decay_mode``.  A device-side gradient scale that the
        update applies later (folded clip coefficient, deferred unscale) is divided out of the added term, so the
        update sees ``scale * g + reg(p)`` as if the gradient had been scaled first."""
        from ..regularizer import L1Decay
        mode = self._decay_mode
        own_any = any(getattr(p, "regularizer", None) is not None for p, _ in params_grads)
        if mode == "grad":
            if not own_any and self.regularization is None and not any("weight_decay" in g for g in self._param_groups):
                return params_grads
        elif not own_any and not (mode == "kernel" and any(
                isinstance(g.get("weight_decay", self.regularization), L1Decay) for g in self._param_groups)):
            return params_grads
        scale = self._grad_scale_dev
        out = []
        for p, g in params_grads:
            grp = self._group_of(p)
            default = grp.get("weight_decay", self.regularization)
            own = getattr(p, "regularizer", None)
            if mode == "decoupled":
                default = None
            if own is None and (default is None or (mode == "kernel" and not isinstance(default, L1Decay))):
                out.append((p, g))
                continue
            zero = torch.zeros_like(g)
            (_, term), = self.append_regularization_ops([(p, zero)], default if default is not None else 0.0)
            if mode == "kernel":
                c = self._coeff(grp)           # the update will add c * p on its own: take it back out here
                if c:
                    term = term - c * p.detach().to(g.dtype)
            if scale is not None:
                term = term / scale.to(term.dtype)
            out.append((p, g + term))
        return out

    def _can_fold_clip(self, params_grads):
        from ..nn.clip import ClipGradByGlobalNorm
        return ((type(self._grad_clip) is ClipGradByGlobalNorm or hasattr(self._grad_clip, "fold_coef"))
                and getattr(self, "_fold_clip_ok", False) and bool(params_grads)
                and all(self._native(p) and getattr(p, "need_clip", True) for p, _ in params_grads))

    def _update(self, params_grads):
        raise NotImplementedError

    def minimize(self, loss, startup_program=None, parameters=None, no_grad_set=None):
        """Apply the gradients that ``loss.backward()`` left on the parameters (dygraph contract of the reference:
        minimize does not run the backward pass itself).  Returns (None, [(param, grad)])."""
        params_grads = self._collect_params_grads()
        self.step()
        return None, params_grads

    def backward(self, loss, startup_program=None, parameters=None, no_grad_set=None, callbacks=None):
        """[(param, grad)] of the parameters that already hold a gradient (reference: optimizer.py backward, dygraph
        branch); ``parameters`` restricts the list."""
        pairs = self._collect_params_grads()
        if parameters is not None:
            wanted = {id(p) for p in parameters}
            pairs = [(p, g) for p, g in pairs if id(p) in wanted]
        return pairs

    def append_regularization_ops(self, parameters_and_grads, regularization=None):
        """grad + regulariser(param) for every pair: the parameter's own ``regularizer`` attribute wins over
        ``regularization`` (or the optimizer's ``weight_decay``)."""
        from ..regularizer import L1Decay, L2Decay
        default = regularization if regularization is not None else self.regularization
        out = []
        for p, g in parameters_and_grads:
            reg = getattr(p, "regularizer", None) or default
            if g is None or reg is None:
                out.append((p, g))
                continue
            if isinstance(reg, (int, float)):
                reg = L2Decay(float(reg))
            coeff = float(getattr(reg, "_coeff", getattr(reg, "coeff", 0.0)))
            if isinstance(reg, L1Decay):
                out.append((p, g + coeff * torch.sign(p.detach()).to(g.dtype)))
            elif isinstance(reg, L2Decay):
                out.append((p, g + coeff * p.detach().to(g.dtype)))
            else:
                out.append((p, g))
        return out

    @torch.no_grad()
    def apply_gradients(self, params_grads):
        """Clip and apply explicit [(param, grad)] pairs."""
        return self._apply_optimize(None, None, [(p, g) for p, g in params_grads if g is not None])

    def get_opti_var_name_list(self):
        return [f"{p}_{acc}" for acc, table in self._accumulators.items() for p in table]

    def clear_grad(self, set_to_zero=True):
        for p, _ in self._all_params():
            if p.grad is not None:
                if set_to_zero:
                    p.grad.zero_()
                else:
                    p.grad = None
            mg = getattr(p, "main_grad", None)
            if mg is not None:
                mg.zero_()

    clear_gradients = clear_grad

    def _group_of(self, p):
        for g in self._param_groups:
            for q in g["params"]:
                if q is p:
                    return g
        return self._param_groups[0]

    # ------------------------------------------------------------- state dict
    def state_dict(self):
        sd = OrderedDict()
        for name, per_param in self._accumulators.items():
            for pname, t in per_param.items():
                sd[f"{pname}_{name}_0"] = t
        if self._master_weights:
            sd["master_weights"] = dict(self._master_weights)
        if isinstance(self._learning_rate, LRScheduler):
            sd["LR_Scheduler"] = self._learning_rate.state_dict()
        return sd

    def set_state_dict(self, state_dict):
        state_dict = dict(state_dict)
        if isinstance(self._learning_rate, LRScheduler) and "LR_Scheduler" in state_dict:
            self._learning_rate.set_state_dict(state_dict.pop("LR_Scheduler"))
        state_dict.pop("LR_Scheduler", None)
        mw = state_dict.pop("master_weights", None)
        if mw is not None:
            for k, v in mw.items():
                if k in self._master_weights:
                    self._master_weights[k].copy_(torch.as_tensor(_np_to_t(v)).to(self._master_weights[k].device))
            self._accumulators_holder["master_weights"] = {k: _np_to_t(v) for k, v in mw.items()}
        for k, v in state_dict.items():
            v = _np_to_t(v)
            placed = False
            for name, per_param in self._accumulators.items():
                suffix = f"_{name}_0"
                if k.endswith(suffix) and k[: -len(suffix)] in per_param:
                    tgt = per_param[k[: -len(suffix)]]
                    tgt.copy_(v.to(tgt.device, tgt.dtype).reshape(tgt.shape))
                    placed = True
                    break
            if not placed:
                self._accumulators_holder[k] = v

    set_dict = set_state_dict
    load_state_dict = set_state_dict

    # ------------------------------------------------------------- helpers
    def _native(self, p):
        return p.is_cuda and flag("use_native_kernels") and p.dtype in (torch.float32, torch.bfloat16, torch.float16)

    def _decay_enabled(self, p, group):
        return True

    def _lr_ratio(self, p):
        return float(getattr(p, "optimize_attr", {}).get("learning_rate", 1.0))


def _np_to_t(v):
    if isinstance(v, torch.Tensor):
        return v
    if isinstance(v, np.ndarray):
        if v.dtype == np.uint16:
            return torch.from_numpy(v.copy()).view(torch.bfloat16)
        return torch.from_numpy(np.ascontiguousarray(v))
    return torch.as_tensor(v)


# ======================================================================================= SGD / Momentum
class SGD(Optimizer):
    _decay_mode = "kernel"
    _masks_found_inf = True
    def __init__(self, learning_rate=0.001, parameters=None, weight_decay=None, grad_clip=None, multi_precision=False,
                 name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name, multi_precision)

    def _coeff(self, group):
        wd = group.get("weight_decay", self.regularization)
        if wd is None:
            return 0.0
        return float(getattr(wd, "_coeff", getattr(wd, "coeff", wd)))

    def _update(self, params_grads):
        lr = self.get_lr()
        for p, g in params_grads:
            grp = self._group_of(p)
            coeff = self._coeff(grp)
            master = self._master_weights.get(p.name) if self._use_master(p) else None
            tgt = master if master is not None else p
            gf = g.to(tgt.dtype)
            if self._grad_scale_dev is not None:
                gf = gf * self._grad_scale_dev.to(gf.dtype)
            if coeff:
                gf = gf + coeff * tgt
            upd = lr * self._lr_ratio(p) * grp.get("learning_rate", 1.0) * gf
            if self._found_inf is not None:
                upd = torch.where(self._found_inf.bool(), torch.zeros_like(upd), upd)
            tgt.sub_(upd)
            if master is not None:
                p.copy_(master)


class Momentum(SGD):
    def __init__(self, learning_rate=0.001, momentum=0.9, parameters=None, use_nesterov=False, weight_decay=None,
                 grad_clip=None, multi_precision=False, rescale_grad=1.0, use_multi_tensor=False, name=None):
        Optimizer.__init__(self, learning_rate, parameters, weight_decay, grad_clip, name, multi_precision)
        self._momentum, self._use_nesterov, self._rescale_grad = momentum, use_nesterov, rescale_grad

    def _create_accumulators(self, params):
        super()._create_accumulators(params)
        for p in params:
            self._add_accumulator("velocity", p)

    def _update(self, params_grads):
        lr = self.get_lr()
        if all(self._native(p) for p, _ in params_grads):
            by_group = defaultdict(list)
            for p, g in params_grads:
                by_group[id(self._group_of(p))].append((p, g))
            for gid, pgs in by_group.items():
                grp = self._group_of(pgs[0][0])
                tab = self._tables.setdefault(("mom", gid), ops.TensorTable(pgs[0][0].device))
                rows = []
                for p, g in pgs:
                    master = self._master_weights.get(p.name) if self._use_master(p) else None
                    rows.append(dict(p=p.data_ptr(), g=g.data_ptr(), m=self._get_accumulator("velocity", p).data_ptr(),
                                     master=master.data_ptr() if master is not None else 0, numel=p.numel(),
                                     lr_ratio=self._lr_ratio(p), decay=1.0, p_dtype=dtype_code(p), g_dtype=dtype_code(g)))
                tab.update(rows)
                ops.momentum_step(tab, lr * grp.get("learning_rate", 1.0), self._momentum, self._coeff(grp),
                                  self._use_nesterov, self._grad_scale_dev, self._found_inf)
            return
        for p, g in params_grads:
            grp = self._group_of(p)
            coeff = self._coeff(grp)
            master = self._master_weights.get(p.name) if self._use_master(p) else None
            tgt = master if master is not None else p
            vel = self._get_accumulator("velocity", p)
            gf = g.to(torch.float32) * self._rescale_grad
            if self._grad_scale_dev is not None:
                gf = gf * self._grad_scale_dev
            if coeff:
                gf = gf + coeff * tgt.float()
            new_vel = self._momentum * vel + gf
            lrr = lr * self._lr_ratio(p) * grp.get("learning_rate", 1.0)
            upd = (gf + self._momentum * new_vel) * lrr if self._use_nesterov else lrr * new_vel
            if self._found_inf is not None:
                skip = self._found_inf.bool()
                upd = torch.where(skip, torch.zeros_like(upd), upd)
                new_vel = torch.where(skip, vel, new_vel)
            vel.copy_(new_vel)
            tgt.sub_(upd.to(tgt.dtype))
            if master is not None:
                p.copy_(master)


# ======================================================================================= Adam / AdamW
class Adam(Optimizer):
    _decoupled = False
    _masks_found_inf = True
    _fold_clip_ok = True      # the native kernel multiplies grads by a device scalar

    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-08, parameters=None, weight_decay=None,
                 grad_clip=None, lazy_mode=False, multi_precision=False, use_multi_tensor=False, amsgrad=False, name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name, multi_precision)
        self._beta1, self._beta2, self._epsilon = float(beta1), float(beta2), float(epsilon)
        self._beta_pows = {}   # device -> fp32[2] (beta1^t, beta2^t)

    def _coeff(self, group):
        """Decoupled decay coefficient; plain Adam has none (its regulariser is added to the gradients by the base)."""
        if not self._decoupled:
            return 0.0
        wd = group.get("weight_decay", self.regularization)
        if wd is None:
            return 0.0
        return float(getattr(wd, "_coeff", getattr(wd, "coeff", wd)))

    def _create_accumulators(self, params):
        super()._create_accumulators(params)
        for p in params:
            self._add_accumulator("moment1", p)
            self._add_accumulator("moment2", p)
            self._add_accumulator("beta1_pow_acc", p, fill=self._beta1, shape=[1])
            self._add_accumulator("beta2_pow_acc", p, fill=self._beta2, shape=[1])

    def _pows(self, p):
        """Shared per-device (beta1^t, beta2^t) for the step being applied, advanced on device unless found_inf."""
        dev = p.device
        if dev not in self._beta_pows:
            b1 = self._get_accumulator("beta1_pow_acc", p)
            b2 = self._get_accumulator("beta2_pow_acc", p)
            self._beta_pows[dev] = torch.cat([b1.float().reshape(1), b2.float().reshape(1)]).to(dev)
        return self._beta_pows[dev]

    def _advance_pows(self, pows):
        betas = torch.tensor([self._beta1, self._beta2], dtype=torch.float32, device=pows.device)
        if self._found_inf is not None:
            betas = torch.where(self._found_inf.to(pows.device).bool(), torch.ones_like(betas), betas)
        pows.mul_(betas)

    def _update(self, params_grads):
        lr = self.get_lr()
        by_group = defaultdict(list)
        for p, g in params_grads:
            by_group[id(self._group_of(p))].append((p, g))
        advanced = set()
        for gid, pgs in by_group.items():
            grp = self._group_of(pgs[0][0])
            glr = lr * grp.get("learning_rate", 1.0)
            coeff = self._coeff(grp)
            b1, b2 = float(grp.get("beta1", self._beta1)), float(grp.get("beta2", self._beta2))
            pows = self._pows(pgs[0][0])
            if all(self._native(p) for p, _ in pgs):
                tab = self._tables.setdefault(("adam", gid), ops.TensorTable(pgs[0][0].device))
                rows = []
                for p, g in pgs:
                    master = self._master_weights.get(p.name) if self._use_master(p) else None
                    rows.append(dict(p=p.data_ptr(), g=g.data_ptr(), m=self._get_accumulator("moment1", p).data_ptr(),
                                     v=self._get_accumulator("moment2", p).data_ptr(),
                                     master=master.data_ptr() if master is not None else 0, numel=p.numel(),
                                     lr_ratio=self._lr_ratio(p) * self._param_lr_ratio(p),
                                     decay=1.0 if self._decay_enabled(p, grp) else 0.0,
                                     p_dtype=dtype_code(p), g_dtype=dtype_code(g)))
                tab.update(rows)
                if not self._decoupled and coeff:
                    raise NotImplementedError("Adam with L2 regularization on GPU: use AdamW (decoupled) or Momentum")
                ops.adamw_step(tab, glr, b1, b2, self._epsilon, coeff if self._decoupled else 0.0,
                               beta_pows=pows, grad_scale=self._grad_scale_dev, found_inf=self._found_inf)
            else:
                for p, g in pgs:
                    self._update_one_torch(p, g, glr, coeff, b1, b2, pows, grp)
            if id(pows) not in advanced:
                advanced.add(id(pows))
        for dev, pows in self._beta_pows.items():
            self._advance_pows(pows)

    def _param_lr_ratio(self, p):
        return 1.0

    def _update_one_torch(self, p, g, lr, coeff, b1, b2, pows, grp):
        master = self._master_weights.get(p.name) if self._use_master(p) else None
        tgt = master if master is not None else p
        m, v = self._get_accumulator("moment1", p), self._get_accumulator("moment2", p)
        gf = g.float()
        if self._grad_scale_dev is not None:
            gf = gf * self._grad_scale_dev
        lrr = lr * self._lr_ratio(p) * self._param_lr_ratio(p)
        pf = tgt.float()
        if self._decoupled:
            if self._decay_enabled(p, grp):
                pf = pf * (1.0 - lrr * coeff)
        elif coeff:
            gf = gf + coeff * pf
        m_new = b1 * m + (1 - b1) * gf
        v_new = b2 * v + (1 - b2) * gf * gf
        pw = pows.to(p.device)
        denom = v_new.sqrt() / torch.sqrt(1 - pw[1]) + self._epsilon
        pf = pf - (lrr / (1 - pw[0])) * (m_new / denom)
        if self._found_inf is not None:
            skip = self._found_inf.to(p.device).bool()
            pf = torch.where(skip, tgt.float(), pf)
            m_new = torch.where(skip, m, m_new)
            v_new = torch.where(skip, v, v_new)
        m.copy_(m_new)
        v.copy_(v_new)
        tgt.copy_(pf.to(tgt.dtype))
        if master is not None:
            p.copy_(master)

    def state_dict(self):
        # materialise the shared device pows into the per-parameter accumulators the reference format carries
        for dev, pows in self._beta_pows.items():
            for pname, t in self._accumulators["beta1_pow_acc"].items():
                if t.device == dev:
                    t.copy_(pows[0:1])
            for pname, t in self._accumulators["beta2_pow_acc"].items():
                if t.device == dev:
                    t.copy_(pows[1:2])
        return super().state_dict()

    def set_state_dict(self, state_dict):
        super().set_state_dict(state_dict)
        self._beta_pows.clear()


class AdamW(Adam):
    """Decoupled weight decay: ``p *= 1 - lr * coeff`` before the Adam update (adamw_kernel.cu:35-130)."""
    _decoupled = True
    _decay_mode = "decoupled"

    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-08, parameters=None, weight_decay=0.01,
                 lr_ratio=None, apply_decay_param_fun=None, grad_clip=None, lazy_mode=False, multi_precision=False,
                 amsgrad=False, name=None):
        super().__init__(learning_rate, beta1, beta2, epsilon, parameters, weight_decay, grad_clip, lazy_mode,
                         multi_precision, name=name)
        self._lr_ratio_fn = lr_ratio
        self._apply_decay_param_fun = apply_decay_param_fun

    def _decay_enabled(self, p, group):
        if self._apply_decay_param_fun is not None:
            return bool(self._apply_decay_param_fun(p.name))
        return True

    def _param_lr_ratio(self, p):
        return float(self._lr_ratio_fn(p)) if self._lr_ratio_fn is not None else 1.0


class Adagrad(Optimizer):
    def __init__(self, learning_rate, epsilon=1e-06, parameters=None, weight_decay=None, grad_clip=None, name=None,
                 initial_accumulator_value=0.0):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._epsilon, self._init = epsilon, initial_accumulator_value

    def _create_accumulators(self, params):
        for p in params:
            self._add_accumulator("moment", p, fill=self._init)

    def _update(self, params_grads):
        lr = self.get_lr()
        for p, g in params_grads:
            mom = self._get_accumulator("moment", p)
            gf = g.float()
            mom.add_(gf * gf)
            p.sub_((lr * gf / (mom.sqrt() + self._epsilon)).to(p.dtype))


class RMSProp(Optimizer):
    def __init__(self, learning_rate, rho=0.95, epsilon=1e-06, momentum=0.0, centered=False, parameters=None,
                 weight_decay=None, grad_clip=None, name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._rho, self._epsilon, self._momentum, self._centered = rho, epsilon, momentum, centered

    def _create_accumulators(self, params):
        for p in params:
            self._add_accumulator("mean_square", p)
            self._add_accumulator("momentum", p)
            self._add_accumulator("mean_grad", p)

    def _update(self, params_grads):
        lr = self.get_lr()
        for p, g in params_grads:
            ms, mom, mg = (self._get_accumulator(n, p) for n in ("mean_square", "momentum", "mean_grad"))
            gf = g.float()
            ms.mul_(self._rho).add_((1 - self._rho) * gf * gf)
            if self._centered:
                mg.mul_(self._rho).add_((1 - self._rho) * gf)
                denom = (ms - mg * mg + self._epsilon).sqrt()
            else:
                denom = (ms + self._epsilon).sqrt()
            mom.mul_(self._momentum).add_(lr * gf / denom)
            p.sub_(mom.to(p.dtype))


class Lamb(Adam):
    _fold_clip_ok = False
    _decay_mode = "decoupled"  # lamb_weight_decay enters the trust-ratio update, not the gradient
    _masks_found_inf = False

    def _coeff(self, group):
        wd = group.get("weight_decay", self.regularization)
        return 0.0 if wd is None else float(getattr(wd, "_coeff", getattr(wd, "coeff", wd)))

    def __init__(self, learning_rate=0.001, lamb_weight_decay=0.01, beta1=0.9, beta2=0.999, epsilon=1e-06,
                 parameters=None, grad_clip=None, exclude_from_weight_decay_fn=None, multi_precision=False,
                 always_adapt=False, name=None):
        super().__init__(learning_rate, beta1, beta2, epsilon, parameters, lamb_weight_decay, grad_clip,
                         multi_precision=multi_precision, name=name)
        self._exclude = exclude_from_weight_decay_fn

    def _update(self, params_grads):
        lr = self.get_lr()
        for p, g in params_grads:
            pows = self._pows(p)
            master = self._master_weights.get(p.name) if self._use_master(p) else None
            tgt = master if master is not None else p
            m, v = self._get_accumulator("moment1", p), self._get_accumulator("moment2", p)
            gf = g.float()
            m.mul_(self._beta1).add_((1 - self._beta1) * gf)
            v.mul_(self._beta2).add_((1 - self._beta2) * gf * gf)
            mh, vh = m / (1 - pows[0]), v / (1 - pows[1])
            wd = 0.0 if (self._exclude is not None and self._exclude(p)) else self._coeff(self._group_of(p))
            r = mh / (vh.sqrt() + self._epsilon) + wd * tgt.float()
            pn, rn = tgt.float().norm(), r.norm()
            trust = torch.where((pn > 0) & (rn > 0), pn / rn, torch.ones_like(pn))
            tgt.sub_((lr * trust * r).to(tgt.dtype))
            if master is not None:
                p.copy_(master)
        for dev, pows in self._beta_pows.items():
            self._advance_pows(pows)


# ---------------------------------------------------------------------------------------------- further optimizers
# (reference: python/paddle/optimizer/{adadelta,adamax,nadam,radam,asgd,rprop}.py; update rules of the phi kernels)
class Adadelta(Optimizer):
    def __init__(self, learning_rate=0.001, epsilon=1e-06, rho=0.95, parameters=None, weight_decay=None, grad_clip=None,
                 name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._epsilon, self._rho = epsilon, rho

    def _create_accumulators(self, params):
        for p in params:
            self._add_accumulator("_avg_squared_grad", p)
            self._add_accumulator("_avg_squared_update", p)

    def _update(self, params_grads):
        lr = self.get_lr()
        for p, g in params_grads:
            eg, eu = self._get_accumulator("_avg_squared_grad", p), self._get_accumulator("_avg_squared_update", p)
            gf = g.float()
            eg.mul_(self._rho).add_((1 - self._rho) * gf * gf)
            upd = -((eu + self._epsilon) / (eg + self._epsilon)).sqrt() * gf
            eu.mul_(self._rho).add_((1 - self._rho) * upd * upd)
            p.add_((lr * upd).to(p.dtype))


class Adamax(Optimizer):
    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-08, parameters=None, weight_decay=None,
                 grad_clip=None, name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._beta1, self._beta2, self._epsilon = beta1, beta2, epsilon

    def _create_accumulators(self, params):
        for p in params:
            self._add_accumulator("moment", p)
            self._add_accumulator("inf_norm", p)
            self._add_accumulator("beta1_pow_acc", p, shape=[1], fill=self._beta1)

    def _update(self, params_grads):
        lr = self.get_lr()
        for p, g in params_grads:
            m, u, b1p = (self._get_accumulator(n, p) for n in ("moment", "inf_norm", "beta1_pow_acc"))
            gf = g.float()
            m.mul_(self._beta1).add_((1 - self._beta1) * gf)
            torch.maximum(u * self._beta2, gf.abs() + self._epsilon, out=u)
            p.sub_((lr / (1 - b1p) * m / u).to(p.dtype))
            b1p.mul_(self._beta1)


class NAdam(Optimizer):
    def __init__(self, learning_rate=0.002, beta1=0.9, beta2=0.999, epsilon=1e-08, momentum_decay=0.004, parameters=None,
                 weight_decay=None, grad_clip=None, name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._beta1, self._beta2, self._epsilon, self._psi = beta1, beta2, epsilon, momentum_decay
        self._t = 0
        self._mu_product = 1.0

    def _create_accumulators(self, params):
        for p in params:
            self._add_accumulator("moment1", p)
            self._add_accumulator("moment2", p)

    def _update(self, params_grads):
        lr = self.get_lr()
        self._t += 1
        t = self._t
        mu_t = self._beta1 * (1 - 0.5 * 0.96 ** (t * self._psi))
        mu_next = self._beta1 * (1 - 0.5 * 0.96 ** ((t + 1) * self._psi))
        self._mu_product *= mu_t
        for p, g in params_grads:
            m, v = self._get_accumulator("moment1", p), self._get_accumulator("moment2", p)
            gf = g.float()
            m.mul_(self._beta1).add_((1 - self._beta1) * gf)
            v.mul_(self._beta2).add_((1 - self._beta2) * gf * gf)
            m_hat = mu_next * m / (1 - self._mu_product * mu_next) + (1 - mu_t) * gf / (1 - self._mu_product)
            v_hat = v / (1 - self._beta2 ** t)
            p.sub_((lr * m_hat / (v_hat.sqrt() + self._epsilon)).to(p.dtype))


class RAdam(Optimizer):
    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-08, parameters=None, weight_decay=None,
                 grad_clip=None, name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._beta1, self._beta2, self._epsilon = beta1, beta2, epsilon
        self._t = 0

    def _create_accumulators(self, params):
        for p in params:
            self._add_accumulator("moment1", p)
            self._add_accumulator("moment2", p)

    def _update(self, params_grads):
        lr = self.get_lr()
        self._t += 1
        t, b1, b2 = self._t, self._beta1, self._beta2
        rho_inf = 2.0 / (1 - b2) - 1
        rho_t = rho_inf - 2.0 * t * b2 ** t / (1 - b2 ** t)
        for p, g in params_grads:
            m, v = self._get_accumulator("moment1", p), self._get_accumulator("moment2", p)
            gf = g.float()
            m.mul_(b1).add_((1 - b1) * gf)
            v.mul_(b2).add_((1 - b2) * gf * gf)
            m_hat = m / (1 - b1 ** t)
            if rho_t > 5.0:
                l_t = ((1 - b2 ** t) ** 0.5) / (v.sqrt() + self._epsilon)
                r_t = (((rho_t - 4) * (rho_t - 2) * rho_inf) / ((rho_inf - 4) * (rho_inf - 2) * rho_t)) ** 0.5
                p.sub_((lr * m_hat * r_t * l_t).to(p.dtype))
            else:
                p.sub_((lr * m_hat).to(p.dtype))


class ASGD(Optimizer):
    """Stochastic average gradient over the last ``batch_num`` gradients (reference: python/paddle/optimizer/asgd.py:
    30-45 for the rule, 146-168 for the state).  Per parameter: ``d`` the running sum of the kept gradients, ``y`` the
    ``[batch_num, *shape]`` ring of them, ``m`` the step count; all three are accumulators, so they are saved."""

    def __init__(self, learning_rate=0.001, batch_num=1, parameters=None, weight_decay=None, grad_clip=None,
                 multi_precision=False, name=None):
        if batch_num is None or batch_num <= 0:
            raise ValueError("batch_num should be greater than 0")
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._n = int(batch_num)
        self._multi_precision = multi_precision

    def _create_accumulators(self, params):
        for p in params:
            if self._use_master(p):
                self._create_master_weight(p)
            self._add_accumulator("d", p)
            self._add_accumulator("y", p, shape=[self._n] + list(p.shape))
            self._add_accumulator("m", p, dtype=torch.int64, shape=[1])

    def _update(self, params_grads):
        lr = self.get_lr()
        for p, g in params_grads:
            d, ys, m = self._get_accumulator("d", p), self._get_accumulator("y", p), self._get_accumulator("m", p)
            step = int(m.item())
            slot = step % self._n
            gf = g.float()
            d.add_(gf - ys[slot])
            ys[slot].copy_(gf)
            m.add_(1)
            target = self._master_weights[p.name] if self._use_master(p) else p
            target.sub_((lr * d / min(step + 1, self._n)).to(target.dtype))
            if target is not p:
                p.copy_(target.to(p.dtype))


class Rprop(Optimizer):
    def __init__(self, learning_rate=0.001, learning_rate_range=(1e-5, 50), parameters=None, etas=(0.5, 1.2), grad_clip=None,
                 multi_precision=False, name=None):
        super().__init__(learning_rate, parameters, None, grad_clip, name)
        self._range, self._etas = learning_rate_range, etas

    def _create_accumulators(self, params):
        for p in params:
            self._add_accumulator("prev", p)
            self._add_accumulator("learning_rate", p, fill=float(self.get_lr()))

    def _update(self, params_grads):
        eta_n, eta_p = self._etas
        lo, hi = self._range
        for p, g in params_grads:
            prev, lrs = self._get_accumulator("prev", p), self._get_accumulator("learning_rate", p)
            gf = g.float().clone()
            sign = (gf * prev).sign()
            lrs.mul_(torch.where(sign > 0, eta_p, torch.where(sign < 0, eta_n, 1.0))).clamp_(lo, hi)
            gf[sign < 0] = 0
            p.sub_((lrs * gf.sign()).to(p.dtype))
            prev.copy_(gf)


class LBFGS(Optimizer):
    """L-BFGS with optional strong-Wolfe line search (reference: python/paddle/optimizer/lbfgs.py); the two-loop
    recursion and line search are torch.optim.LBFGS, driven through Paddle's ``step(closure)`` contract."""

    def __init__(self, learning_rate=1.0, max_iter=20, max_eval=None, tolerance_grad=1e-07, tolerance_change=1e-09,
                 history_size=100, line_search_fn=None, parameters=None, weight_decay=None, grad_clip=None, name=None):
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name)
        self._impl = torch.optim.LBFGS([p for p, _ in self._all_params()], lr=float(self.get_lr()), max_iter=max_iter,
                                       max_eval=max_eval, tolerance_grad=tolerance_grad, tolerance_change=tolerance_change,
                                       history_size=history_size, line_search_fn=line_search_fn)

    def step(self, closure):
        def wrapped():
            self.clear_grad(set_to_zero=False)
            with torch.enable_grad():
                loss = closure()
            return loss
        for g in self._impl.param_groups:
            g["lr"] = float(self.get_lr())
        return self._impl.step(wrapped)

    def state_dict(self):
        return {"lbfgs": self._impl.state_dict()}

    def set_state_dict(self, state_dict):
        if "lbfgs" in state_dict:
            self._impl.load_state_dict(state_dict["lbfgs"])
