"""Worker: expert-parallel MoELayer (ep=2, global_scatter/global_gather) == dense per-token expert evaluation."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200 import nn
from paddle_b200.incubate.distributed.models.moe import MoELayer

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
g = dist.new_group(list(range(n)))
D, E_LOCAL = 8, 2


class Expert(nn.Layer):
    def __init__(self):
        super().__init__()
        self.fc = nn.Linear(D, D)

    def forward(self, x):
        return torch.relu(self.fc(x))


paddle.seed(5)
with paddle.utils.unique_name.guard():
    all_experts = [Expert() for _ in range(E_LOCAL * n)]      # identical on every rank (same seed)
    local = nn.LayerList([all_experts[r * E_LOCAL + i] for i in range(E_LOCAL)])
    moe = MoELayer(D, local, gate={"type": "naive", "top_k": 2}, moe_group=g)
dist.broadcast(moe.gate.gate.weight.detach(), 0)
dist.broadcast(moe.gate.gate.bias.detach(), 0)
torch.manual_seed(40 + r)
x = torch.randn(2, 5, D, requires_grad=True)
y = moe(x)
y.sum().backward()
# dense oracle on this rank's tokens with the full expert set
xr = x.detach().clone().requires_grad_(True)
flat = xr.reshape(-1, D)
logits = flat @ moe.gate.gate.weight.detach() + moe.gate.gate.bias.detach()
val, idx = torch.topk(logits, 2, dim=-1)
out = torch.zeros_like(flat)
for t in range(flat.shape[0]):
    for k in range(2):
        out[t] = out[t] + val[t, k] * all_experts[int(idx[t, k])](flat[t:t + 1])[0]
out.sum().backward()
torch.testing.assert_close(y.reshape(-1, D), out, rtol=1e-4, atol=1e-5)
torch.testing.assert_close(x.grad, xr.grad, rtol=1e-4, atol=1e-5)
assert all(getattr(p, "no_sync", False) for p in local.parameters())
print("moe ok", r)

# ---- grouped experts (stacked weights, one module for all local experts) == the same experts as a LayerList
from paddle_b200.incubate.distributed.models.moe import GroupedExpertsFFN
from paddle_b200.models.moe import ExpertFFN, MoEConfig

cfg = MoEConfig.tiny(hidden_size=D, intermediate_size=16)
paddle.seed(9)
with paddle.utils.unique_name.guard():
    ffns = [ExpertFFN(cfg) for _ in range(E_LOCAL * n)]
    mine = [ffns[r * E_LOCAL + i] for i in range(E_LOCAL)]
    moe_list = MoELayer(D, nn.LayerList(mine), gate={"type": "naive", "top_k": 2}, moe_group=g)
    moe_grp = MoELayer(D, GroupedExpertsFFN.from_experts(mine), gate={"type": "naive", "top_k": 2}, moe_group=g)
for a, b in ((moe_grp.gate.gate.weight, moe_list.gate.gate.weight), (moe_grp.gate.gate.bias, moe_list.gate.gate.bias)):
    a.detach().copy_(b.detach())
    dist.broadcast(a.detach(), 0)
    dist.broadcast(b.detach(), 0)
xa = x.detach().clone().requires_grad_(True)
xb = x.detach().clone().requires_grad_(True)
ya, yb = moe_list(xa), moe_grp(xb)
ya.sum().backward()
yb.sum().backward()
torch.testing.assert_close(yb, ya, rtol=1e-4, atol=1e-5)
torch.testing.assert_close(xb.grad, xa.grad, rtol=1e-4, atol=1e-5)
for i, e in enumerate(mine):
    torch.testing.assert_close(moe_grp.experts.gate_up_weight.grad[i], e.gate_up_proj.weight.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(moe_grp.experts.down_weight.grad[i], e.down_proj.weight.grad, rtol=1e-4, atol=1e-5)
print("moe grouped ok", r)
