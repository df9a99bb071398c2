"""Flash attention: tcgen05 kernel vs torch SDPA (cuDNN / FA2 library path) at the Llama-2 7B shape."""
import json, math, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200 import ops
from paddle_b200.framework import flags


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


rows = []
for (b, s, h, d) in [(1, 4096, 32, 128), (2, 4096, 32, 128), (4, 2048, 32, 128), (1, 8192, 32, 128)]:
    q, k, v = (torch.randn(b, s, h, d, device="cuda").bfloat16() for _ in range(3))
    fl = 4.0 * b * h * s * s * d / 2
    flags.set_flags({"use_native_kernels": True})
    t_o = timeit(lambda: ops.flash_attention(q, k, v, causal=True))
    flags.set_flags({"use_native_kernels": False})
    t_l = timeit(lambda: ops.flash_attention(q, k, v, causal=True))
    flags.set_flags({"use_native_kernels": True})
    # backward only (forward graph built once, grads recomputed each call); 2.5x the forward FLOPs
    def bwd_timer(native):
        flags.set_flags({"use_native_kernels": native})
        qq, kk, vv = (t.detach().requires_grad_(True) for t in (q, k, v))
        out = ops.flash_attention(qq, kk, vv, causal=True)
        g = torch.randn_like(out)
        return timeit(lambda: torch.autograd.grad(out, (qq, kk, vv), g, retain_graph=True))
    tb_o, tb_l = bwd_timer(True), bwd_timer(False)
    flags.set_flags({"use_native_kernels": True})
    rows.append(dict(shape=[b, s, h, d], ours_ms=round(t_o, 4), lib_ms=round(t_l, 4), ours_tflops=round(fl / t_o / 1e9, 1),
                     lib_tflops=round(fl / t_l / 1e9, 1), bwd_ours_ms=round(tb_o, 4), bwd_lib_ms=round(tb_l, 4),
                     bwd_ours_tflops=round(2.5 * fl / tb_o / 1e9, 1), bwd_lib_tflops=round(2.5 * fl / tb_l / 1e9, 1)))
    print(rows[-1], flush=True)
json.dump(rows, open(os.path.join(os.path.dirname(__file__), "..", "gpurun_out", "attn_bench.json"), "w"), indent=1)
