"""Worker (4 ranks, gloo): a 2 x 2 mesh (dp, mp).  Batch sharded over dp, Megatron-style weights over mp: the loss and
every parameter gradient equal the single-process result, with no collective written by the user."""
import torch
import torch.nn.functional as F

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import auto_parallel as ap

dist.init_parallel_env()
r = dist.get_rank()
mesh = ap.ProcessMesh([[0, 1], [2, 3]], dim_names=["dp", "mp"])
dp, mp = mesh.coord_of(r)
R_, S = ap.Replicate, ap.Shard
torch.manual_seed(0)
x, y = torch.randn(8, 16), torch.randn(8, 4)
w1, w2 = torch.randn(16, 32) * 0.3, torch.randn(32, 4) * 0.3
rw1, rw2 = w1.clone().requires_grad_(True), w2.clone().requires_grad_(True)
ref = ((F.silu(x @ rw1) @ rw2 - y) ** 2).mean()
ref.backward()

dx = ap.shard_tensor(x, mesh, [S(0), R_()])
dy = ap.shard_tensor(y, mesh, [S(0), R_()])
dw1 = ap.shard_tensor(torch.nn.Parameter(w1.clone()), mesh, [R_(), S(1)])
dw2 = ap.shard_tensor(torch.nn.Parameter(w2.clone()), mesh, [R_(), S(0)])
h = F.silu(dx @ dw1)
assert h.placements == [S(0), S(1)] and tuple(h.shape) == (4, 16)
out = h @ dw2
assert out.placements[0] == S(0) and out.placements[1].is_partial()
loss = ((out - dy) ** 2).mean()
torch.testing.assert_close(torch.tensor(float(loss)), ref.detach(), rtol=1e-5, atol=1e-5)
loss.backward()
torch.testing.assert_close(dw1.grad, rw1.grad[:, 16 * mp:16 * mp + 16], rtol=1e-5, atol=1e-5)
torch.testing.assert_close(dw2.grad, rw2.grad[16 * mp:16 * mp + 16], rtol=1e-5, atol=1e-5)
# reshard across both axes at once and back
full = ap.unshard_dtensor(ap.reshard(h.detach(), mesh, [S(1), S(0)]))
torch.testing.assert_close(full, F.silu(x @ w1), rtol=1e-5, atol=1e-5)
print("auto_parallel_2d ok", r)
