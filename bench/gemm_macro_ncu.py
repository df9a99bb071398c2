"""ncu target: the macro-tile GEMM (and the pair kernel for comparison) once each inside a profiler range."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200.framework import flags
from paddle_b200.ops import gemm as G

M, N, K = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (8192, 22016, 4096)
a = torch.randn(M, K, device="cuda").bfloat16()
b = torch.randn(K, N, device="cuda").bfloat16()
out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
for kind in ("macro", "pair"):
    flags.set_flags({"gemm_kernel": kind})
    G.gemm(a, b, out=out)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for kind in ("macro", "pair"):
    flags.set_flags({"gemm_kernel": kind})
    G.gemm(a, b, out=out)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
