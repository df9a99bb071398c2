"""Worker (2 ranks, gloo): pipeline stages as meshes.  Layer 1 lives on mesh [0], layer 2 on mesh [1]; the activation is
resharded from one mesh to the other.  Every rank runs the same script; the loss and all gradients equal the
single-process model, and two optimizer steps leave the same parameters."""
import torch
import torch.nn.functional as F

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import auto_parallel as ap

dist.init_parallel_env()
r = dist.get_rank()
m0, m1 = ap.ProcessMesh([0], dim_names=["x"]), ap.ProcessMesh([1], dim_names=["x"])
R_ = ap.Replicate


class Net(paddle.nn.Layer):
    def __init__(self):
        super().__init__()
        self.fc1 = paddle.nn.Linear(16, 32)
        self.fc2 = paddle.nn.Linear(32, 4)

    def forward(self, a):
        h = F.gelu(self.fc1(a))
        if ap.is_dist(h):
            h = ap.reshard(h, m1, [R_()])            # stage boundary
        return self.fc2(h)


def build():
    paddle.seed(5)
    return Net()


def shard_fn(name, layer, _):
    mesh = {"fc1": m0, "fc2": m1}.get(name)
    if mesh is not None:
        layer.weight = ap.shard_tensor(layer.weight, mesh, [R_()])
        layer.bias = ap.shard_tensor(layer.bias, mesh, [R_()])


ref, net = build(), ap.shard_layer(build(), m0, shard_fn)
assert (net.fc1.weight.numel() > 0) == (r == 0) and (net.fc2.weight.numel() > 0) == (r == 1)
ropt = paddle.optimizer.AdamW(1e-2, parameters=ref.parameters())
opt = ap.shard_optimizer(paddle.optimizer.AdamW(1e-2, parameters=net.parameters()))
torch.manual_seed(0)
for step in range(2):
    x, y = torch.randn(8, 16), torch.randn(8, 4)
    rl = ((ref(x) - y) ** 2).mean()
    rl.backward()
    out = net(ap.shard_tensor(x, m0, [R_()]))
    loss = ((out - ap.shard_tensor(y, m1, [R_()])) ** 2).mean()
    if r == 1:
        torch.testing.assert_close(torch.tensor(float(loss)), rl.detach(), rtol=1e-5, atol=1e-5)
    loss.backward()
    mine, theirs = (net.fc1, ref.fc1) if r == 0 else (net.fc2, ref.fc2)
    torch.testing.assert_close(mine.weight.grad, theirs.weight.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(mine.bias.grad, theirs.bias.grad, rtol=1e-5, atol=1e-5)
    ropt.step(); ropt.clear_grad()
    opt.step(); opt.clear_grad()
mine, theirs = (net.fc1, ref.fc1) if r == 0 else (net.fc2, ref.fc2)
torch.testing.assert_close(ap.dtensor_to_local(mine.weight), theirs.weight.detach(), rtol=1e-4, atol=1e-5)
print("auto_parallel_pp ok", r)
