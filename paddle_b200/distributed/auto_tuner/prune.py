"""Pruning rules (reference: python/paddle/distributed/auto_tuner/prune.py - registered functions that reject a
candidate before it is launched, from model divisibility and from the history of finished trials)."""
from __future__ import annotations

_PRUNES = []


def register_prune(fn):
    _PRUNES.append(fn)
    return fn


@register_prune
def prune_by_divisibility(tuner_cfg, cur, history):
    m = tuner_cfg.get("model_cfg", {})
    heads, layers, vocab = m.get("num_attention_heads"), m.get("num_layers"), m.get("vocab_size")
    hidden, gbs = m.get("hidden_size"), m.get("global_batch_size")
    if heads and heads % cur["mp_degree"]:
        return True
    if hidden and hidden % cur["mp_degree"]:
        return True
    if vocab and vocab % cur["mp_degree"]:
        return True
    if layers and layers % cur["pp_degree"]:
        return True
    if gbs:
        data_ranks = cur["dp_degree"] * cur["sharding_degree"]
        if gbs % data_ranks or (gbs // data_ranks) % cur["micro_batch_size"]:
            return True
        acc = gbs // data_ranks // cur["micro_batch_size"]
        if cur["pp_degree"] > 1 and acc < cur["pp_degree"]:
            return True         # 1F1B needs at least pp micro-batches in flight
    return False


@register_prune
def prune_by_memory(tuner_cfg, cur, history):
    """A configuration that is a strictly heavier version of one that already ran out of memory cannot fit either."""
    for h in history:
        if h.get("error") != "oom":
            continue
        same_split = all(h[k] == cur[k] for k in ("dp_degree", "mp_degree", "pp_degree", "sharding_degree", "sharding_stage"))
        if same_split and cur["micro_batch_size"] >= h["micro_batch_size"] and (h["use_recompute"] or not cur["use_recompute"]):
            return True
    return False


@register_prune
def prune_by_memory_model(tuner_cfg, cur, history):
    """The analytic estimate (cost_model.estimate_memory_gb) says the candidate cannot fit in HBM.  Needs the model
    shape in ``model_cfg``; off unless ``tuner_cfg["use_memory_model"]`` (default on when the shape is known)."""
    m = tuner_cfg.get("model_cfg", {})
    if not all(k in m for k in ("hidden_size", "num_layers", "vocab_size")) or tuner_cfg.get("use_memory_model") is False:
        return False
    from .cost_model import HBM_GB, estimate_memory_gb
    return estimate_memory_gb(cur, m) > float(tuner_cfg.get("max_mem_usage_gb", HBM_GB * 0.92))


def should_prune(tuner_cfg, cur, history):
    return any(fn(tuner_cfg, cur, history) for fn in _PRUNES)
