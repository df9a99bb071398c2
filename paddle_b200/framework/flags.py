"""Exported-flag registry (reference: paddle/common/flags.cc PHI_DEFINE_EXPORTED_*, python/paddle/base/framework.py
set_flags/get_flags).  Flags are typed, pick up ``FLAGS_<name>`` from the environment at import, and are readable
from native code through ``pdb200_flag_get`` in the runtime library (mirrored on every set)."""
from __future__ import annotations

import os
from typing import Any, Callable

_REGISTRY: dict[str, dict[str, Any]] = {}
_listeners: list[Callable[[str, Any], None]] = []


def _parse(kind, raw):
    if kind is bool:
        return str(raw).lower() in ("1", "true", "yes", "on")
    return kind(raw)


def define_flag(name: str, default, doc: str = "", kind=None):
    kind = kind or type(default)
    val = default
    env = os.environ.get("FLAGS_" + name)
    if env is not None:
        try:
            val = _parse(kind, env)
        except ValueError:
            val = default
    _REGISTRY[name] = {"value": val, "default": default, "kind": kind, "doc": doc}
    return val


def _key(name: str) -> str:
    return name[6:] if name.startswith("FLAGS_") else name


def set_flags(flags: dict):
    for k, v in flags.items():
        k = _key(k)
        if k not in _REGISTRY:
            raise ValueError(f"Flag {k} cannot set its value through this function.")
        ent = _REGISTRY[k]
        ent["value"] = _parse(ent["kind"], v) if not isinstance(v, ent["kind"]) else v
        for fn in _listeners:
            fn(k, ent["value"])


def get_flags(flags):
    if isinstance(flags, str):
        flags = [flags]
    out = {}
    for k in flags:
        kk = _key(k)
        if kk not in _REGISTRY:
            raise ValueError(f"Flag {kk} cannot get its value through this function.")
        out["FLAGS_" + kk] = _REGISTRY[kk]["value"]
    return out


def flag(name: str):
    return _REGISTRY[_key(name)]["value"]


def modified_flags():
    return {k: v["value"] for k, v in _REGISTRY.items() if v["value"] != v["default"]}


def add_listener(fn):
    _listeners.append(fn)


# ---- the subset of the reference's 156 exported flags that the collective-training product reads ----
define_flag("sync_nccl_allreduce", True, "host-sync after all-reduce (flags.cc:320)")
define_flag("nccl_blocking_wait", False, "host spin until collective done (flags.cc:1242)")
define_flag("benchmark_nccl", False, "device sync after every collective (flags.cc:1246)")
define_flag("benchmark", False, "device sync after every op")
define_flag("dynamic_static_unified_comm", True, "(flags.cc:1571)")
define_flag("enable_async_trace", False, "comm watchdog (flags.cc:1586)")
define_flag("async_trace_count", 5, "(flags.cc:1590)")
define_flag("use_stream_safe_cuda_allocator", True, "record_stream on comm buffers")
define_flag("allocator_strategy", "auto_growth", "")
define_flag("fraction_of_gpu_memory_to_use", 0.92, "")
define_flag("cudnn_deterministic", False, "")
define_flag("embedding_deterministic", 0, "")
define_flag("check_nan_inf", False, "per-op output scan")
define_flag("check_nan_inf_level", 0, "")
define_flag("eager_communication_connection", False, "create communicators at init")
define_flag("nccl_nrings", 1, "")
define_flag("selected_gpus", "", "device ordinal(s) for this process")
define_flag("stop_check_timeout", 900, "")
define_flag("sharding_sort_parameters", True, "")
define_flag("p2p_device_synchronize", False, "")
define_flag("pipeline_nccl_comm_init_option", 0, "")
define_flag("enable_nccl_dynamic_check", False, "")
# product-specific switches
define_flag("use_fused_comm", True, "use in-kernel NVLink collectives where available; 0 = NCCL-only baseline path")
define_flag("use_native_kernels", True, "use hand-written sm_100a kernels; 0 = torch composition (oracle)")
define_flag("gemm_macro_tile", True, "large bf16 GEMMs use the 256x512 macro-tile CTA-pair kernel with the TMA-store epilogue (gemm_tc3.cu)")
define_flag("gemm_kernel", "", "force one GEMM kernel: macro | pair | single (empty: pick per shape)")
define_flag("comm_watchdog_poll_ms", 10000, "polling interval of the communication watchdog thread")
define_flag("comm_watchdog_abort", False, "the watchdog ends the process (exit code 113) when a collective times out, so the launcher can restart the pod")
define_flag("use_native_process_group", False, "route paddle.distributed collectives on CUDA tensors through the native NCCL process group of libpdb200_runtime.so (own communicators, comm stream, event-based tasks) instead of torch.distributed")
define_flag("ag_gemm_overlap", True, "K5: chunked copy-engine all-gather with per-chunk flags, consumed by the gated CTA-pair GEMM while later chunks are still in flight (False: gather, then GEMM)")
define_flag("gemm_rs_overlap", True, "K6: the GEMM counts finished owner blocks, a side stream ships each block to its owner with the copy engines while the GEMM continues, a local kernel sums the n partial blocks (False: GEMM, then one NVLS pull-reduce kernel)")
define_flag("stage3_nvl_transport", True, "GroupSharded stage 3: move parameters/gradients with the NVLink symmetric-memory kernels (K3 push all-gather, pull reduce-scatter) instead of NCCL")
define_flag("gemm_cta_pair", True, "use the CTA-pair (cta_group::2, 256x256 tile) GEMM kernel for large shapes")
define_flag("use_tcgen05_gemm", True, "route Linear GEMMs through the tcgen05/TMA kernel when shapes allow")
define_flag("comm_watchdog_timeout_ms", 1800000, "")
