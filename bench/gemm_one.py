"""A few launches of the CTA-pair tcgen05 GEMM at the Llama-2 7B gate_up shape (target of ncu captures)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200.ops import gemm as G

M, N, K = 8192, 22016, 4096
a = torch.randn(M, K, device="cuda").bfloat16()
b = torch.randn(K, N, device="cuda").bfloat16()
for _ in range(4):
    c = G.gemm(a, b)
torch.cuda.synchronize()
print("ok", float(c[0, 0]))
