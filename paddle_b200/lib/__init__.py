"""ctypes binding of the in-tree native libraries.

Signatures are derived from the C headers under ``include/`` (one ``PDB_API`` declaration per statement), so the
header is the single source of truth for the ABI.  On a GPU box a missing library is a hard error: ops must not
silently fall back to a torch composition there (the driver records which ``.so`` files the process loaded).
"""
from __future__ import annotations

import ctypes
import os
import re
from pathlib import Path

_LIB_DIR = Path(__file__).resolve().parent
_ROOT = _LIB_DIR.parents[1]

_CTYPES = {
    "void*": ctypes.c_void_p, "const void*": ctypes.c_void_p,
    "float*": ctypes.c_void_p, "const float*": ctypes.c_void_p,
    "int64_t*": ctypes.c_void_p, "const int64_t*": ctypes.c_void_p,
    "int*": ctypes.c_void_p, "const int*": ctypes.c_void_p,
    "int32_t*": ctypes.c_void_p, "const int32_t*": ctypes.c_void_p,
    "uint64_t*": ctypes.c_void_p, "const uint64_t*": ctypes.c_void_p,
    "uint8_t*": ctypes.c_void_p, "const uint8_t*": ctypes.c_void_p,
    "char*": ctypes.c_char_p, "const char*": ctypes.c_char_p,
    "int": ctypes.c_int, "int64_t": ctypes.c_int64, "uint64_t": ctypes.c_uint64, "float": ctypes.c_float,
    "double": ctypes.c_double, "void": None,
}
_DECL = re.compile(r"PDB_API\s+([\w\s\*]+?)\s*(\w+)\s*\(([^)]*)\)\s*;")


def _norm_type(t: str) -> str:
    t = t.strip()
    t = re.sub(r"\s*\*\s*", "*", t)
    return re.sub(r"\s+", " ", t)


def parse_header(path: Path):
    """-> {name: (restype, [argtypes], [argnames])}"""
    out = {}
    text = re.sub(r"/\*.*?\*/", "", path.read_text(), flags=re.S)
    text = "\n".join(l for l in text.splitlines() if not l.lstrip().startswith("#"))
    for m in _DECL.finditer(text):
        ret, name, args = _norm_type(m.group(1)), m.group(2), m.group(3).strip()
        argtypes, argnames = [], []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                mm = re.match(r"(.+?)(\w+)$", a)
                argtypes.append(_CTYPES[_norm_type(mm.group(1))])
                argnames.append(mm.group(2))
        out[name] = (_CTYPES[ret], argtypes, argnames)
    return out


class NativeLib:
    def __init__(self, so_name: str, header: str, env_override: str = ""):
        self.so_path = _LIB_DIR / so_name
        if env_override and os.environ.get(env_override):      # sanitizer builds (tools/sanitize.sh)
            self.so_path = Path(os.environ[env_override])
        self.header = _ROOT / "include" / header
        self._dll = None
        self._err = None

    def available(self) -> bool:
        return self._load() is not None

    def _load(self):
        if self._dll is not None or self._err is not None:
            return self._dll
        if not self.so_path.exists():
            self._err = f"{self.so_path} not built (run `python -m paddle_b200.lib.build`)"
            return None
        try:
            dll = ctypes.CDLL(str(self.so_path), mode=ctypes.RTLD_GLOBAL)
            for name, (ret, argtypes, _) in parse_header(self.header).items():
                fn = getattr(dll, name, None)
                if fn is None:
                    raise OSError(f"symbol {name} declared in {self.header.name} missing from {self.so_path.name}")
                fn.restype = ret
                fn.argtypes = argtypes
            self._dll = dll
        except OSError as e:  # noqa: PERF203
            self._err = str(e)
        return self._dll

    def require(self):
        dll = self._load()
        if dll is None:
            raise RuntimeError(f"native library unavailable: {self._err}")
        return dll

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        fn = getattr(self.require(), name)
        extra = _LAUNCHES.get(name)
        if extra is None:
            return fn

        def counted(*args):
            _count[0] += extra
            return fn(*args)
        counted.__name__ = name
        self.__dict__[name] = counted     # cache: later lookups bypass __getattr__
        return counted


# kernel launches per entry point (for bench.py's "gpu_launches" claim); entries not listed launch nothing
_LAUNCHES = {
    "pdb200_norm_fwd": 1, "pdb200_norm_bwd": 2, "pdb200_swiglu_fwd": 1, "pdb200_swiglu_bwd": 1, "pdb200_rope": 1, "pdb200_qkv_rope": 1,
    "pdb200_bias_act": 1, "pdb200_adamw_multi": 1, "pdb200_momentum_multi": 1, "pdb200_l2norm_multi": 2,
    "pdb200_unscale_check_multi": 1, "pdb200_ce_fwd": 1, "pdb200_ce_combine": 1, "pdb200_ce_bwd": 1,
    "pdb200_number_count": 1, "pdb200_assign_pos": 1, "pdb200_limit_by_capacity": 1,
    "pdb200_prune_gate_by_capacity": 2, "pdb200_random_routing": 1, "pdb200_topk_softmax": 1,
    "pdb200_gemm_bf16": 1, "pdb200_gemm_bf16_pair": 1, "pdb200_gemm_bf16_macro": 1, "pdb200_gemm_bf16_gated": 1, "pdb200_gemm_bf16_counted": 1, "pdb200_nvl_wait_counter": 1, "pdb200_nvl_wait_flags": 1, "pdb200_nvl_signal": 1, "pdb200_check_numerics": 2, "pdb200_gemm_grouped": 1, "pdb200_gemm_grouped_wgrad": 1, "pdb200_flash_attn_fwd": 1, "pdb200_flash_attn_bwd": 3, "pdb200_flash_attn_fwd_dropout": 1, "pdb200_flash_attn_bwd_dropout": 3, "pdb200_flash_attn_varlen_fwd": 1, "pdb200_flash_attn_varlen_bwd": 3,
    "pdb200_nvl_barrier": 1, "pdb200_nvl_allreduce": 3, "pdb200_nvl_rs_adamw_ag": 3, "pdb200_nvl_reduce_scatter": 3,
    "pdb200_nvl_ce_combine": 3, "pdb200_nvl_vocab_embed": 3, "pdb200_nvl_allgather": 3, "pdb200_nvl_allgather_multi": 1, "pdb200_nvl_reduce_scatter_multi": 1, "pdb200_reduce_slots_multi": 1, "pdb200_reduce_slots_multi2": 1, "pdb200_nvl_signal_one": 1, "pdb200_nvl_allgather_from": 3, "pdb200_nvl_rs_sum_bf16": 3, "pdb200_moe_dispatch": 5, "pdb200_moe_pull": 3,
}
_count = [0]


def launch_count() -> int:
    return _count[0]


def reset_launch_count():
    _count[0] = 0


kernels = NativeLib("libpdb200_kernels.so", "pdb200_kernels.h", env_override="PDB200_KERNELS_LIB")
runtime = NativeLib("libpdb200_runtime.so", "pdb200_runtime.h", env_override="PDB200_RUNTIME_LIB")


class KernelError(RuntimeError):
    pass


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = kernels.require().pdb200_last_error()
        raise KernelError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")


def have_kernels() -> bool:
    return kernels.available()


def ensure_built(verbose=False):
    from . import build
    return build.build_all(verbose)
