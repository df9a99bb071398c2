"""Per-kernel roofline harness for the tcgen05 GEMM vs cuBLAS (torch.matmul) at Llama-2 7B/13B linear shapes.
Times with CUDA events after warm-up; operands are rotated through > 126 MB of buffers so L2 never holds them."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200.framework import flags  # noqa: E402
from paddle_b200.ops import gemm as G  # noqa: E402

PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["bf16_tflops"] \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 1590.0


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    T = 4096
    shapes = [("qkv fwd", T, 12288, 4096, False, False), ("o fwd", T, 4096, 4096, False, False),
              ("gate_up fwd", T, 22016, 4096, False, False), ("down fwd", T, 4096, 11008, False, False),
              ("lm_head fwd", T, 32000, 4096, False, False),
              ("gate_up dgrad", T, 4096, 22016, False, True), ("gate_up wgrad", 4096, 22016, T, True, False),
              ("down wgrad", 11008, 4096, T, True, False), ("13b qkv fwd", T, 15360, 5120, False, False)]
    rows = []
    for name, M, N, K, ta, tb in shapes:
        nbuf = max(2, int(200e6 // ((M * K + K * N) * 2)) + 1)
        As = [torch.randn((K, M) if ta else (M, K), device="cuda").bfloat16() for _ in range(nbuf)]
        Bs = [torch.randn((N, K) if tb else (K, N), device="cuda").bfloat16() for _ in range(nbuf)]
        i = [0]

        def ours():
            i[0] = (i[0] + 1) % nbuf
            return G.gemm(As[i[0]], Bs[i[0]], ta, tb)

        def cublas():
            i[0] = (i[0] + 1) % nbuf
            a = As[i[0]].t() if ta else As[i[0]]
            b = Bs[i[0]].t() if tb else Bs[i[0]]
            return torch.matmul(a, b)
        i[0] = 0
        ref = cublas().float()
        i[0] = 0
        got = ours().float()
        err = (got - ref).abs().max().item() / max(ref.abs().max().item(), 1e-6)
        t_o, t_c = timeit(ours), timeit(cublas)
        flags.set_flags({"gemm_cta_pair": False})
        t_s = timeit(ours)
        flags.set_flags({"gemm_cta_pair": True})
        fl = 2.0 * M * N * K
        rows.append(dict(name=name, M=M, N=N, K=K, ours_ms=round(t_o, 4), cublas_ms=round(t_c, 4),
                         ours_tflops=round(fl / t_o / 1e9, 1), single_cta_tflops=round(fl / t_s / 1e9, 1),
                         cublas_tflops=round(fl / t_c / 1e9, 1),
                         frac_of_measured_peak=round(fl / t_o / 1e9 / PEAK, 3), rel_err=err))
        print(rows[-1], flush=True)
    out = os.path.join(os.path.dirname(__file__), "..", "gpurun_out", "gemm_bench.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(rows, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
