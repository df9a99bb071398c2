"""Worker (>= 2 GPUs, NCCL): MoELayer through the fused NVLink dispatch / combine kernels (K9 / K10) == the same
layer through NCCL all-to-all (FLAGS_use_fused_comm=0), forward and backward, fp32 and bf16."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200 import nn
from paddle_b200.incubate.distributed.models.moe import MoELayer
from paddle_b200.parallel import moe_fused, symm

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", torch.cuda.current_device())
g = dist.collective._get_default_group()
D, E_LOCAL, TOPK = 64, 2, 2


class Expert(nn.Layer):
    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(D, 2 * D)
        self.fc2 = nn.Linear(2 * D, D)

    def forward(self, x):
        return self.fc2(torch.nn.functional.gelu(self.fc1(x)))


for dtype, tol in ((torch.float32, dict(rtol=1e-4, atol=1e-5)), (torch.bfloat16, dict(rtol=3e-2, atol=3e-2))):
    paddle.seed(5 + r)
    with paddle.utils.unique_name.guard():
        local = nn.LayerList([Expert() for _ in range(E_LOCAL)])
        moe = MoELayer(D, local, gate={"type": "naive", "top_k": TOPK}, moe_group=g)
    moe.to(dev)
    if dtype == torch.bfloat16:
        moe.to(dtype="bfloat16")
    dist.broadcast(moe.gate.gate.weight.detach(), 0)
    dist.broadcast(moe.gate.gate.bias.detach(), 0)
    torch.manual_seed(40 + r)
    x0 = torch.randn(3, 50, D, device=dev).to(dtype)
    gout = torch.randn(3, 50, D, device=dev).to(dtype)
    res = {}
    for fused in (False, True):
        paddle.set_flags({"FLAGS_use_fused_comm": fused})
        x = x0.clone().requires_grad_(True)
        for p in moe.parameters():
            p.grad = None
        assert moe_fused.available(g, x.reshape(-1, D)) == fused
        y = moe(x)
        y.backward(gout)
        torch.cuda.synchronize()
        res[fused] = (y.detach().float(), x.grad.float(), [p.grad.float().clone() for p in moe.parameters() if p.grad is not None])
    a, b = res[False], res[True]
    torch.testing.assert_close(b[0], a[0], **tol)
    torch.testing.assert_close(b[1], a[1], **tol)
    assert len(a[2]) == len(b[2]) and len(a[2]) > 0
    for ga, gb in zip(a[2], b[2]):
        torch.testing.assert_close(gb, ga, rtol=tol["rtol"], atol=tol["atol"] * 4)

# ---- grouped experts on the fused path: device-scheduled grouped GEMMs, no host synchronisation in the layer
from paddle_b200.incubate.distributed.models.moe import GroupedExpertsFFN
from paddle_b200.models.moe import ExpertFFN, MoEConfig

DG = 256
cfg = MoEConfig.tiny(hidden_size=DG, intermediate_size=512)
paddle.seed(50 + r)
paddle.set_default_dtype("bfloat16")
with paddle.utils.unique_name.guard():
    ffns = [ExpertFFN(cfg).to(dev) for _ in range(E_LOCAL)]
    moe_list = MoELayer(DG, nn.LayerList(ffns), gate={"type": "naive", "top_k": TOPK}, moe_group=g).to(dev)
    moe_grp = MoELayer(DG, GroupedExpertsFFN.from_experts(ffns).to(dev), gate={"type": "naive", "top_k": TOPK}, moe_group=g).to(dev)
paddle.set_default_dtype("float32")
for m in (moe_list, moe_grp):
    m.gate.gate.weight.detach().copy_(moe_list.gate.gate.weight.detach())
    m.gate.gate.bias.detach().copy_(moe_list.gate.gate.bias.detach())
    dist.broadcast(m.gate.gate.weight.detach(), 0)
    dist.broadcast(m.gate.gate.bias.detach(), 0)
paddle.set_flags({"FLAGS_use_fused_comm": True})
torch.manual_seed(70 + r)
x0 = (torch.randn(4, 96, DG, device=dev) * 0.5).bfloat16()
gout = (torch.randn(4, 96, DG, device=dev) * 0.5).bfloat16()
xa = x0.clone().requires_grad_(True)
ya = moe_list(xa)
ya.backward(gout)
xb = x0.clone().requires_grad_(True)
moe_grp(xb.detach())                                # warm-up: arenas / tensor maps are created outside the checked region
torch.cuda.synchronize()
torch.cuda.set_sync_debug_mode("error")            # any host synchronisation inside the layer raises
try:
    yb = moe_grp(xb)
    yb.backward(gout)
finally:
    torch.cuda.set_sync_debug_mode("default")
torch.cuda.synchronize()
tol = dict(rtol=3e-2, atol=3e-2)
torch.testing.assert_close(yb.float(), ya.float(), **tol)
torch.testing.assert_close(xb.grad.float(), xa.grad.float(), **tol)
for i, e in enumerate(ffns):
    torch.testing.assert_close(moe_grp.experts.gate_up_weight.grad[i].float(), e.gate_up_proj.weight.grad.float(), rtol=3e-2, atol=6e-2)
    torch.testing.assert_close(moe_grp.experts.down_weight.grad[i].float(), e.down_proj.weight.grad.float(), rtol=3e-2, atol=6e-2)
print("moe grouped fused ok", r, flush=True)
print("moe_fused ok", r, flush=True)
moe_fused.release_all()
symm.release_all()
dist.destroy_process_group()
