"""Worker: Ulysses / all-gather-KV / ring attention over sequence shards == single-device attention (fwd + bwd)."""
import math

import torch

import paddle_b200 as paddle  # noqa: F401
import paddle_b200.distributed as dist
from paddle_b200.distributed.fleet.utils import context_parallel as cp

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
g = dist.new_group(list(range(n)))
B, S, H, HK, D = 2, 16 * n, 4, 2, 8
torch.manual_seed(0)
q = torch.randn(B, S, H, D)
k = torch.randn(B, S, HK, D)
v = torch.randn(B, S, HK, D)
go = torch.randn(B, S, H, D)


def ref(q, k, v, causal):
    kf, vf = k.repeat_interleave(H // HK, 2), v.repeat_interleave(H // HK, 2)
    s = torch.einsum("bqhd,bkhd->bhqk", q, kf) / math.sqrt(D)
    if causal:
        s = s.masked_fill(~torch.ones(S, S, dtype=torch.bool).tril(), float("-inf"))
    return torch.einsum("bhqk,bkhd->bqhd", torch.softmax(s, -1), vf)


for causal in (True, False):
    qr, kr, vr = (t.clone().requires_grad_(True) for t in (q, k, v))
    ref(qr, kr, vr, causal).backward(go)
    sl = slice(r * S // n, (r + 1) * S // n)
    for name, fn in (("ulysses", cp.ulysses_attention), ("allgather_kv", cp.allgather_kv_attention),
                     ("ring", cp.ring_flash_attention)):
        ql, kl, vl = (t[:, sl].clone().requires_grad_(True) for t in (q, k, v))
        out = fn(ql, kl, vl, g, causal=causal)
        out.backward(go[:, sl].contiguous())
        want = ref(q, k, v, causal)[:, sl]
        torch.testing.assert_close(out, want, rtol=1e-4, atol=1e-5, msg=lambda m: f"{name} causal={causal} fwd: {m}")
        for a, b_, nm in ((ql, qr, "dq"), (kl, kr, "dk"), (vl, vr, "dv")):
            torch.testing.assert_close(a.grad, b_.grad[:, sl], rtol=1e-4, atol=1e-5,
                                       msg=lambda m: f"{name} causal={causal} {nm}: {m}")
# split / gather helpers
x = torch.arange(S * 3, dtype=torch.float32).reshape(1, S, 3).requires_grad_(True)
y = cp.gather_sequence(cp.split_sequence(x, g, 1), g, 1)
torch.testing.assert_close(y, x)
print("context_parallel ok", r)
