"""tcgen05 / TMA GEMM entry (K23/K13): dispatch between the CTA-pair kernel (csrc/kernels/gemm_tc2.cu, 256x256 tiles)
and the single-CTA kernel (gemm_tc.cu, 128xBN tiles).  ``supported`` is the shape / layout gate Linear consults;
unsupported shapes go to cuBLAS through torch.matmul."""
from __future__ import annotations

import torch

from .. import lib
from ..framework.flags import flag
from ._util import check, ptr, stream_ptr

_MIN_DIM = 256


def _have():
    if not lib.have_kernels():
        return False
    return hasattr(lib.kernels.require(), "pdb200_gemm_bf16")


def supported(a, b, trans_a=False, trans_b=False, accum_into=None) -> bool:
    if not _have() or a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        return False
    if a.dim() != 2 or b.dim() != 2 or not a.is_contiguous() or not b.is_contiguous():
        return False
    M, K = (a.shape[1], a.shape[0]) if trans_a else a.shape
    N = b.shape[0] if trans_b else b.shape[1]
    if accum_into is not None and (accum_into.dtype != torch.float32 or not accum_into.is_contiguous()):
        return False
    return M >= _MIN_DIM and N >= _MIN_DIM and K >= 64 and M % 8 == 0 and N % 8 == 0 and K % 8 == 0 and \
        a.data_ptr() % 16 == 0 and b.data_ptr() % 16 == 0


def _pick(M, N, K, beta):
    """Kernel choice: "macro" (gemm_tc3: 256x512 blocks per CTA pair, TMA-store epilogue) for the large shapes,
    "pair" (gemm_tc2: 256x256) and "single" (gemm_tc: 128xBN) below that."""
    forced = flag("gemm_kernel")
    if forced in ("macro", "pair", "single"):
        return forced
    if not flag("gemm_cta_pair") or M < 512 or N < 512:
        return "single"
    if flag("gemm_macro_tile") and beta in (0.0, 1.0) and hasattr(lib.kernels.require(), "pdb200_gemm_bf16_macro"):
        return "macro"
    return "pair"


def gemm(a, b, trans_a=False, trans_b=False, out_dtype=None, accum_into=None, bias=None, act=None, out=None):
    """C[M,N] = op(A) @ op(B) with A,B bf16 row-major storage.  accum_into: fp32 C += (K13).  out: preallocated
    contiguous [M,N] bf16 / fp32 result buffer."""
    M, K = (a.shape[1], a.shape[0]) if trans_a else a.shape
    N = b.shape[0] if trans_b else b.shape[1]
    if accum_into is not None:
        out = accum_into
        out_code, beta = 0, 1.0
    elif out is not None:
        assert out.is_contiguous() and out.numel() == M * N and out.dtype in (torch.bfloat16, torch.float32)
        out_code, beta = (0 if out.dtype == torch.float32 else 1), 0.0
    else:
        od = out_dtype or torch.bfloat16
        out = torch.empty(M, N, dtype=od, device=a.device)
        out_code, beta = (0 if od == torch.float32 else 1), 0.0
    kind = _pick(M, N, K, beta)
    fn = {"macro": lib.kernels.pdb200_gemm_bf16_macro, "pair": lib.kernels.pdb200_gemm_bf16_pair,
          "single": lib.kernels.pdb200_gemm_bf16}[kind] if kind != "macro" or hasattr(lib.kernels.require(), "pdb200_gemm_bf16_macro") \
        else lib.kernels.pdb200_gemm_bf16_pair
    check(fn(ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, int(trans_a), int(trans_b), out_code, float(beta),
             {"none": 0, None: 0, "gelu": 1, "relu": 2}[act], stream_ptr(a.device)), "gemm_bf16")
    return out
