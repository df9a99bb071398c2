"""Worker (2 ranks, gloo): DistTensors propagate their sharding through ordinary ops and autograd gives the right shard of
the global gradient.  Each scenario is compared with the same computation done whole, in one process, by plain torch:
  tp    column-parallel then row-parallel MLP (replicated input, Shard(1) / Shard(0) weights)
  dp    batch-sharded input, replicated parameters (the gradient all-reduce is implied by the placements)
  vocab vocab-parallel embedding + class-sharded cross entropy
  layer paddle layers with shard_layer / shard_optimizer / shard_dataloader: two AdamW steps with a global-norm clip and a
        ShardingStage1 partition of the update
"""
import torch
import torch.nn.functional as F

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import auto_parallel as ap

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
mesh = ap.ProcessMesh([0, 1], dim_names=["x"])
R_, S = ap.Replicate, ap.Shard
loc = ap.dtensor_to_local


def close(a, b, tol=1e-5, what=""):
    torch.testing.assert_close(a, b, rtol=tol, atol=tol, msg=lambda m: f"{what}: {m}")


torch.manual_seed(0)
x, y = torch.randn(8, 16), torch.randn(8, 4)
w1, b1, w2 = torch.randn(16, 32) * 0.3, torch.randn(32) * 0.1, torch.randn(32, 4) * 0.3


def ref_mlp():
    ws = [t.clone().requires_grad_(True) for t in (w1, b1, w2)]
    xx = x.clone().requires_grad_(True)
    loss = ((F.gelu(xx @ ws[0] + ws[1]) @ ws[2] - y) ** 2).mean()
    loss.backward()
    return loss.detach(), xx.grad, [w.grad for w in ws]


ref_loss, ref_dx, (g1, gb1, g2) = ref_mlp()

# ---- tp
dx = ap.shard_tensor(x.clone().requires_grad_(True), mesh, [R_()])
dw1 = ap.shard_tensor(torch.nn.Parameter(w1.clone()), mesh, [S(1)])
db1 = ap.shard_tensor(torch.nn.Parameter(b1.clone()), mesh, [S(0)])
dw2 = ap.shard_tensor(torch.nn.Parameter(w2.clone()), mesh, [S(0)])
h = F.gelu(dx @ dw1 + db1)
assert h.placements == [S(1)] and tuple(h.shape) == (8, 16)
out = h @ dw2
assert out.placements[0].is_partial()
loss = ((out - y) ** 2).mean()
assert loss.placements == [R_()], loss.placements
close(loc(loss), ref_loss, what="tp loss")
loss.backward()
close(dw1.grad, g1[:, 16 * r:16 * r + 16], what="tp dw1")
close(db1.grad, gb1[16 * r:16 * r + 16], what="tp db1")
close(dw2.grad, g2[16 * r:16 * r + 16], what="tp dw2")
close(dx.grad, ref_dx, what="tp dx (replicated input in front of a column-parallel matmul)")

# ---- dp
dx = ap.shard_tensor(x, mesh, [S(0)])
dy = ap.shard_tensor(y, mesh, [S(0)])
ps = [ap.shard_tensor(torch.nn.Parameter(t.clone()), mesh, [R_()]) for t in (w1, b1, w2)]
loss = ((F.gelu(dx @ ps[0] + ps[1]) @ ps[2] - dy) ** 2).mean()
assert loss.placements[0].is_partial()                       # mean over a sharded dim: an averaged partial
close(torch.tensor(float(loss)), ref_loss, what="dp loss value")
loss.backward()                                              # the pending average is part of the function
for p, g, name in zip(ps, (g1, gb1, g2), ("w1", "b1", "w2")):
    close(p.grad, g, what=f"dp grad {name}")

# ---- vocab-parallel embedding + class-sharded cross entropy
torch.manual_seed(1)
V, H, T = 12, 8, 10
table, head = torch.randn(V, H), torch.randn(H, V) * 0.5
ids, lab = torch.randint(0, V, (T,)), torch.randint(0, V, (T,))
lab[3] = -100
rt, rh = table.clone().requires_grad_(True), head.clone().requires_grad_(True)
ref = F.cross_entropy(F.embedding(ids, rt) @ rh, lab, ignore_index=-100)
ref.backward()
dt = ap.shard_tensor(torch.nn.Parameter(table.clone()), mesh, [S(0)])
dh = ap.shard_tensor(torch.nn.Parameter(head.clone()), mesh, [S(1)])
emb = F.embedding(ap.shard_tensor(ids, mesh, [R_()]), dt)
assert emb.placements[0].is_partial()
logits = emb @ dh
assert logits.placements == [S(1)] and tuple(logits.shape) == (T, V // 2)
l2 = F.cross_entropy(logits, ap.shard_tensor(lab, mesh, [R_()]), ignore_index=-100)
close(torch.tensor(float(l2)), ref.detach(), what="vocab-parallel loss")
l2.backward()
close(dt.grad, rt.grad[6 * r:6 * r + 6], what="embedding table grad")
close(dh.grad, rh.grad[:, 6 * r:6 * r + 6], what="head grad")

# ---- shape ops keep what they can: [b, s, h*d] -> [b, s, h, d] with h sharded, transpose, softmax on a whole dim
q = torch.randn(2, 6, 8)
dq = ap.shard_tensor(q, mesh, [S(2)])
q4 = dq.reshape(2, 6, 2, 2)                                   # written against the local shape (4 = 2 heads x 2)
assert q4.placements == [S(2)] and q4._global_shape == (2, 6, 4, 2), (q4.placements, q4._global_shape)
close(ap.unshard_dtensor(q4.transpose(1, 2).softmax(-1)), q.reshape(2, 6, 4, 2).transpose(1, 2).softmax(-1), what="reshape/transpose/softmax")
close(ap.unshard_dtensor(dq[:, 1:4]), q[:, 1:4], what="basic indexing on whole dims")
close(ap.unshard_dtensor(torch.cat([dq, dq], dim=1)), torch.cat([q, q], 1), what="concat")
close(ap.unshard_dtensor(dq.sum(-1)), q.sum(-1), what="sum over the sharded dim")

# ---- paddle layers: shard_layer + shard_optimizer (global-norm clip over sharded grads, ShardingStage1) + shard_dataloader
paddle.seed(7)


class Net(paddle.nn.Layer):
    def __init__(self):
        super().__init__()
        self.fc1 = paddle.nn.Linear(16, 32)
        self.norm = paddle.nn.LayerNorm(32)
        self.fc2 = paddle.nn.Linear(32, 4)

    def forward(self, a):
        return self.fc2(self.norm(F.gelu(self.fc1(a))))


def build():
    paddle.seed(7)
    return Net()


def shard_fn(name, layer, m):
    if name == "fc1":
        layer.weight = ap.shard_tensor(layer.weight, m, [S(1)])
        layer.bias = ap.shard_tensor(layer.bias, m, [S(0)])
    elif name == "fc2":
        layer.weight = ap.shard_tensor(layer.weight, m, [R_()])
        layer.bias = ap.shard_tensor(layer.bias, m, [R_()])
    elif name == "norm":
        layer.weight = ap.shard_tensor(layer.weight, m, [R_()])
        layer.bias = ap.shard_tensor(layer.bias, m, [R_()])


ref_net, net = build(), ap.shard_layer(build(), mesh, shard_fn)
ref_opt = paddle.optimizer.AdamW(1e-2, parameters=ref_net.parameters(), grad_clip=paddle.nn.ClipGradByGlobalNorm(0.5))
opt = ap.shard_optimizer(paddle.optimizer.AdamW(1e-2, parameters=net.parameters(), grad_clip=paddle.nn.ClipGradByGlobalNorm(0.5)),
                         ap.ShardingStage1("x", mesh))
data = [(torch.randn(8, 16), torch.randn(8, 4)) for _ in range(2)]
loader = ap.shard_dataloader(data, mesh, shard_dims="x")
for (gx, gy), (bx, by) in zip(data, loader):
    assert bx.placements == [S(0)] and tuple(bx.shape) == (4, 16)
    rl = ((ref_net(gx) - gy) ** 2).mean()
    rl.backward()
    ref_opt.step()
    ref_opt.clear_grad()
    dl = ((net(bx) - by) ** 2).mean()
    close(torch.tensor(float(dl)), rl.detach(), tol=1e-4, what="layer loss")
    dl.backward()
    opt.step()
    opt.clear_grad()
close(ap.unshard_dtensor(net.fc1.weight), ref_net.fc1.weight.detach(), tol=2e-4, what="fc1.weight after 2 steps")
close(ap.unshard_dtensor(net.fc2.weight), ref_net.fc2.weight.detach(), tol=2e-4, what="fc2.weight after 2 steps")
close(ap.unshard_dtensor(net.norm.weight), ref_net.norm.weight.detach(), tol=2e-4, what="norm.weight after 2 steps")
# ---- to_static / DistModel with a Strategy: gradient merge over 2 calls + sharding stage 2 of the update
ref_net, net = build(), ap.shard_layer(build(), mesh, shard_fn)
ref_opt = paddle.optimizer.AdamW(1e-2, parameters=ref_net.parameters())
strategy = ap.Strategy()
strategy.gradient_merge.enable, strategy.gradient_merge.k_steps = True, 2
strategy.sharding.enable, strategy.sharding.stage = True, 2
mse = lambda out, lab: ((out - lab) ** 2).mean()
dm = ap.to_static(net, None, mse, paddle.optimizer.AdamW(1e-2, parameters=net.parameters()), strategy)
dm.train()
data = [(torch.randn(8, 16), torch.randn(8, 4)) for _ in range(4)]
for i, ((gx, gy), (bx, by)) in enumerate(zip(data, ap.shard_dataloader(data, mesh, shard_dims="x"))):
    (mse(ref_net(gx), gy) / 2).backward()
    if i % 2 == 1:
        ref_opt.step()
        ref_opt.clear_grad()
    dm(bx, by)
close(ap.unshard_dtensor(net.fc1.weight), ref_net.fc1.weight.detach(), tol=2e-4, what="DistModel fc1.weight")
close(ap.unshard_dtensor(net.fc2.bias), ref_net.fc2.bias.detach(), tol=2e-4, what="DistModel fc2.bias")
dm.eval()
close(torch.tensor(float(dm(ap.shard_tensor(data[0][0], mesh, [S(0)]), ap.shard_tensor(data[0][1], mesh, [S(0)])))),
      mse(ref_net(data[0][0]), data[0][1]).detach(), tol=1e-4, what="DistModel eval loss")
# ---- randomness: dropout draws different numbers along a sharding axis, the same numbers along a replicating one
ap.parallel_manual_seed(123)
ones = torch.ones(8, 64)
d_sh = F.dropout(ap.shard_tensor(ones, mesh, [S(0)]), 0.5)
d_re = F.dropout(ap.shard_tensor(ones, mesh, [R_()]), 0.5)
both = [None, None]
dist.all_gather_object(both, (loc(d_sh), loc(d_re)))
assert not torch.equal(both[0][0], both[1][0]), "shards of one tensor must not share a dropout mask"
assert torch.equal(both[0][1], both[1][1]), "replicas must share the dropout mask"
assert not torch.equal(loc(F.dropout(ap.shard_tensor(ones, mesh, [R_()]), 0.5)), loc(d_re)), "successive calls must differ"
sub = ap.ProcessMesh([[0, 1]], dim_names=["a", "b"])
assert sub[0].process_ids == [0, 1] and sub[0].dim_names == ["b"] and (1 in sub) and sub.get_rank_by_dim_and_process_id("b", 1) == 1

# ---- distributed checkpoint of DistTensors: saved under one sharding, loaded under another (reshard on load)
import os
import tempfile
box = [tempfile.mkdtemp() if r == 0 else None]
dist.broadcast_object_list(box, 0)
dist.save_state_dict(dict(net.state_dict()), box[0])


def other_shard_fn(name, layer, m):
    if name in ("fc1", "fc2"):
        layer.weight = ap.shard_tensor(layer.weight, m, [S(0)])      # row-sharded now (fc1 was column-sharded, fc2 whole)
        layer.bias = ap.shard_tensor(layer.bias, m, [R_()])
    elif name == "norm":
        layer.weight = ap.shard_tensor(layer.weight, m, [S(0)])
        layer.bias = ap.shard_tensor(layer.bias, m, [R_()])


paddle.seed(99)
fresh = ap.shard_layer(Net(), mesh, other_shard_fn)
sd = dict(fresh.state_dict())
dist.load_state_dict(sd, box[0])
fresh.set_state_dict(sd)
for name in ("fc1.weight", "fc1.bias", "fc2.weight", "norm.weight"):
    close(ap.unshard_dtensor(dict(fresh.named_parameters())[name]), ap.unshard_dtensor(dict(net.named_parameters())[name]),
          tol=0, what=f"checkpoint reshard {name}")
print("auto_parallel_train ok", r)
