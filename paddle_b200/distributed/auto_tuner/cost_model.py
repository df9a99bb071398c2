"""Memory and step-time estimates of a hybrid-parallel transformer configuration on B200 (reference:
python/paddle/distributed/auto_tuner/cost_model.py ``get_mem`` / ``get_not_oom_cfgs`` and memory_cost_model.py; the
reference's constants describe 80 GB parts and an un-fused attention, these describe this framework on 180 GB parts).

Per GPU, bf16 training with fp32 master weights and AdamW:
  parameters   2 B / param (bf16), divided by mp, pp and - at sharding stage 3 - the sharding degree
  gradients    2 B / param (bf16) (stage >= 2: / sharding)
  optimizer    12 B / param (fp32 master + two moments) (stage >= 1: / sharding)
  activations  per layer and micro-batch ~ s b h (10 + 24 / mp) bytes with the flash-attention kernels (no s^2 term is
               ever stored) - 2 s b h when the layer is recomputed; a pipeline stage holds up to ``pp`` micro-batches
               of its layers in flight (1F1B).
Step time: matmul FLOPs at the measured sustained bf16 rate, plus what the collectives that cannot be hidden cost at
NVLink 5 bandwidth, plus the 1F1B bubble."""
from __future__ import annotations

GB = float(1 << 30)
HBM_GB = 180.0
SUSTAINED_TFLOPS = 1380.0        # cuBLAS bf16 sustained under the power cap (MEASURED_PEAKS.json)
NVLINK_GBS = 900.0               # per direction per GPU


def param_count(m):
    h, L, V = m["hidden_size"], m["num_layers"], m["vocab_size"]
    inter = m.get("intermediate_size", 4 * h)
    kv = m.get("num_key_value_heads", m.get("num_attention_heads", 1)) / max(m.get("num_attention_heads", 1), 1)
    per_layer = h * h * (2 + 2 * kv) + 3 * h * inter + 2 * h
    return L * per_layer + 2 * V * h + h


def estimate_memory_gb(cfg, m):
    """Peak bytes per GPU (GB) of one candidate."""
    mp, pp, sh = cfg["mp_degree"], cfg["pp_degree"], cfg["sharding_degree"]
    stage = cfg.get("sharding_stage", 1) if sh > 1 else 0
    P = param_count(m) / (mp * pp)
    params = 2.0 * P / (sh if stage >= 3 else 1)
    grads = 2.0 * P / (sh if stage >= 2 else 1)
    optim = 12.0 * P / (sh if stage >= 1 else 1)
    s, b, h = m.get("seq_length", m.get("max_position_embeddings", 4096)), cfg["micro_batch_size"], m["hidden_size"]
    layers_here = m["num_layers"] / pp
    per_layer = 2.0 * s * b * h if cfg.get("use_recompute") else s * b * h * (10.0 + 24.0 / mp)
    in_flight = min(pp, max(_acc_steps(cfg, m), 1)) if pp > 1 else 1
    acts = per_layer * layers_here * in_flight
    logits = 6.0 * s * b * m["vocab_size"] / mp if pp == 1 or True else 0.0     # bf16 logits + fp32 softmax statistics
    workspace = 2.0 * GB
    return (params + grads + optim + acts + logits + workspace) / GB


def _acc_steps(cfg, m):
    gbs = m.get("global_batch_size")
    if not gbs:
        return 1
    return max(gbs // (cfg["dp_degree"] * cfg["sharding_degree"]) // cfg["micro_batch_size"], 1)


def estimate_step_ms(cfg, m):
    """Rough step time of one candidate (ms): ranks candidates, not a prediction."""
    mp, pp, sh, dp = cfg["mp_degree"], cfg["pp_degree"], cfg["sharding_degree"], cfg["dp_degree"]
    s, b, h = m.get("seq_length", m.get("max_position_embeddings", 4096)), cfg["micro_batch_size"], m["hidden_size"]
    acc = _acc_steps(cfg, m)
    tokens = s * b * acc
    P = param_count(m)
    flops = (6.0 + (2.0 if cfg.get("use_recompute") else 0.0)) * P * tokens / (mp * pp)
    flops += 12.0 * m["num_layers"] / pp * s * s * h / mp * b * acc * (1.0 if not cfg.get("use_recompute") else 4.0 / 3.0) / 2
    # small per-GPU matmuls run below the sustained rate: one efficiency knee at ~1k rows x 1k columns per GPU
    eff = min(1.0, (s * b) / 2048.0) * min(1.0, (h / mp) / 1024.0) ** 0.5
    compute = flops / (SUSTAINED_TFLOPS * 1e12 * max(eff, 0.05)) * 1e3
    comm = 0.0
    if mp > 1:        # 4 all-reduce-equivalents of [s b h] per layer and micro-batch (fwd + bwd), 2 (n-1)/n wire factor
        comm += 4 * m["num_layers"] / pp * acc * 2.0 * s * b * h * 2 * (mp - 1) / mp / (NVLINK_GBS * 1e9) * 1e3 * 0.5
    data_ranks = dp * sh
    if data_ranks > 1:  # gradient reduce-scatter + parameter all-gather (stage 3: twice), mostly overlapped: count 25 %
        vol = 2.0 * P / (mp * pp) * (3 if cfg.get("sharding_stage", 1) >= 3 and sh > 1 else 2) * (data_ranks - 1) / data_ranks
        comm += 0.25 * vol / (NVLINK_GBS * 1e9) * 1e3
    bubble = (pp - 1) / max(acc, 1) if pp > 1 else 0.0
    return (compute + comm) * (1.0 + bubble)


def get_not_oom_cfgs(cfgs, tuner_cfg):
    m = tuner_cfg.get("model_cfg", {})
    limit = float(tuner_cfg.get("max_mem_usage_gb", HBM_GB * 0.92))
    return [c for c in cfgs if estimate_memory_gb(c, m) <= limit]


def hardware_peaks():
    """Measured peaks of the box (``MEASURED_PEAKS.json`` at the repo root, written by the driver) with the constants
    above as fall-back: {"hbm_gbs", "tflops_bf16", "tflops_bf16_burst", "nvlink_gbs", "hbm_gb"}."""
    import json
    import os
    peaks = {"hbm_gbs": 6550.0, "tflops_bf16": SUSTAINED_TFLOPS, "tflops_bf16_burst": 1614.0, "nvlink_gbs": NVLINK_GBS, "hbm_gb": HBM_GB}
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    path = os.path.join(root, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                m = json.load(f)
            peaks["hbm_gbs"] = float(m.get("hbm_gbs", peaks["hbm_gbs"]))
            peaks["tflops_bf16"] = float(m.get("bf16_tflops_sustained", peaks["tflops_bf16"]))
            peaks["tflops_bf16_burst"] = float(m.get("bf16_tflops", peaks["tflops_bf16_burst"]))
        except (OSError, ValueError):
            pass
    return peaks
