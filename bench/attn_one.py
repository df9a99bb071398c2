"""One forward + one backward of the tcgen05 flash attention at the 7B shape (target of ncu captures)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200 import ops

b, s, h, d = 1, 4096, 32, 128
q, k, v = (torch.randn(b, s, h, d, device="cuda").bfloat16().requires_grad_(True) for _ in range(3))
for _ in range(2):
    o = ops.flash_attention(q, k, v, causal=True)
    o.backward(torch.randn_like(o))
torch.cuda.synchronize()
print("ok")
