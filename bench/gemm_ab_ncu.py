"""ncu target: one launch each of cuBLAS (torch.matmul) and the CTA-pair tcgen05 GEMM on the same operands, inside a
cudaProfilerStart/Stop range (run ncu with --profile-from-start off).  Shape from argv: M N K (default gate_up fwd)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_b200 import lib
from paddle_b200.ops import gemm as G

M, N, K = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (8192, 22016, 4096)
a = torch.randn(M, K, device="cuda").bfloat16()
b = torch.randn(K, N, device="cuda").bfloat16()
out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
for _ in range(2):
    torch.matmul(a, b, out=out)
    G.gemm(a, b, out=out)
torch.cuda.synchronize()
torch.cuda.profiler.start()
torch.matmul(a, b, out=out)
for g in (4, 16):
    lib.kernels.pdb200_gemm_set_group_m(g)
    G.gemm(a, b, out=out)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
