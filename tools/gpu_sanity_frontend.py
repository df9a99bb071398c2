"""Short CUDA sanity pass over front-end layers whose Python changed late in the round (batch norm running
statistics, conv padding algebra, transformer caches, weight norm, amp lists).  Prints one line per check."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import paddle
import paddle.nn as nn
import paddle.nn.functional as F

dev = "cuda" if torch.cuda.is_available() else "cpu"
torch.manual_seed(0)
x = torch.randn(8, 16, 12, 12, device=dev)
bn = nn.BatchNorm2D(16).to(dev)
bn.train()
y = bn(x)
y.sum().backward()
ref_var = 0.9 * torch.ones(16, device=dev) + 0.1 * x.var(axis=[0, 2, 3], unbiased=False)
print("bn_running_var_biased", float((bn._variance - ref_var).abs().max()))
conv = nn.Conv2D(16, 8, 3, stride=2, padding="SAME").to(dev)
print("conv_same", tuple(conv(x).shape))
enc = nn.TransformerEncoderLayer(32, 4, 64).to(dev)
s = torch.randn(2, 5, 32, device=dev)
cache = enc.gen_cache(s)
o, c2 = enc(s, None, cache)
print("encoder_cache", tuple(o.shape), tuple(c2.k.shape))
lin = nn.Linear(32, 32).to(dev)
nn.utils.weight_norm(lin, dim=1)
lin(s).sum().backward()
print("weight_norm_grad", lin.weight_g.grad is not None)
with paddle.amp.auto_cast(custom_black_list={"matmul", "linear"}, dtype="bfloat16"):
    z = nn.Linear(32, 32).to(dev)(s)
print("amp_black_linear", z.dtype)
with paddle.amp.auto_cast(dtype="bfloat16"):
    z = nn.Linear(32, 32).to(dev)(s)
print("amp_default_linear", z.dtype)
print("SANITY_OK")
