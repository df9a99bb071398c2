"""Candidate generation for the parallel-strategy search (reference: python/paddle/distributed/auto_tuner/utils.py)."""
from __future__ import annotations

import itertools


def _divisors(n):
    return [d for d in range(1, n + 1) if n % d == 0]


def default_candidates(tuner_cfg):
    """Fill every dimension the user left on "auto" with its divisors of the device count / natural choices."""
    n = int(tuner_cfg["num_gpus"])
    c = {}
    for key in ("dp_degree", "mp_degree", "pp_degree", "sharding_degree"):
        v = tuner_cfg.get(key, "auto")
        c[key] = _divisors(n) if v == "auto" else list(v) if isinstance(v, (list, tuple)) else [int(v)]
    v = tuner_cfg.get("sharding_stage", "auto")
    c["sharding_stage"] = [1, 2, 3] if v == "auto" else list(v) if isinstance(v, (list, tuple)) else [int(v)]
    gbs = int(tuner_cfg.get("model_cfg", {}).get("global_batch_size", 8))
    v = tuner_cfg.get("micro_batch_size", "auto")
    c["micro_batch_size"] = _divisors(gbs) if v == "auto" else list(v) if isinstance(v, (list, tuple)) else [int(v)]
    v = tuner_cfg.get("use_recompute", "auto")
    c["use_recompute"] = [False, True] if v == "auto" else list(v) if isinstance(v, (list, tuple)) else [bool(v)]
    return c


def search_all(tuner_cfg):
    """Cartesian product of the candidates whose degrees multiply to the device count, cheapest-first."""
    c = tuner_cfg["candidates"]
    n = int(tuner_cfg["num_gpus"])
    keys = ["dp_degree", "mp_degree", "pp_degree", "sharding_degree", "sharding_stage", "micro_batch_size", "use_recompute"]
    out = []
    for combo in itertools.product(*(c[k] for k in keys)):
        cfg = dict(zip(keys, combo))
        if cfg["dp_degree"] * cfg["mp_degree"] * cfg["pp_degree"] * cfg["sharding_degree"] != n:
            continue
        if cfg["sharding_degree"] == 1 and cfg["sharding_stage"] != c["sharding_stage"][0]:
            continue            # the stage is irrelevant without sharding: keep one representative
        out.append(cfg)
    # larger micro batches and no recompute first (faster if they fit), fewer model splits first
    out.sort(key=lambda k: (k["use_recompute"], -k["micro_batch_size"], k["mp_degree"] * k["pp_degree"], k["sharding_stage"]))
    return out
