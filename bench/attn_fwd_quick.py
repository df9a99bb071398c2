"""Forward (and optionally backward) flash attention timing at the 7B shapes; one JSON line per shape.
PDB200_KERNELS_LIB selects a variant library built by tools/build_variant.py."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200 import ops

bwd = "--bwd" in sys.argv


def timeit(fn, iters=20):
    for _ in range(4):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


for (b, s, h, d) in ((2, 4096, 32, 128), (1, 8192, 32, 128)):
    torch.manual_seed(0)
    q, k, v = (torch.randn(b, s, h, d, device="cuda").bfloat16().requires_grad_(bwd) for _ in range(3))
    o = ops.flash_attention(q, k, v, causal=True)
    ref = torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True).transpose(1, 2)
    fl = 4.0 * b * h * s * s * d / 2
    ms = timeit(lambda: ops.flash_attention(q, k, v, causal=True))
    rec = dict(lib=os.path.basename(os.environ.get("PDB200_KERNELS_LIB", "default")), shape=[b, s, h, d], fwd_ms=round(ms, 4),
               fwd_tflops=round(fl / ms / 1e9, 1), max_err_vs_sdpa=round((o.float() - ref.float()).abs().max().item(), 4))
    if bwd:
        g = torch.randn_like(o)
        def fb():
            oo = ops.flash_attention(q, k, v, causal=True)
            oo.backward(g)
        t = timeit(fb) - ms
        rec.update(bwd_ms=round(t, 4), bwd_tflops=round(2.5 * fl / t / 1e9, 1))
    print(json.dumps(rec), flush=True)
