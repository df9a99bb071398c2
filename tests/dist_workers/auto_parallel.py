"""Worker (2 ranks, gloo): reshard function set (r<->s, p->r, p->s via reduce-scatter, s->s via all-to-all, r->p, cross-mesh)
and SPMD-rule driven ops (column / row parallel matmul chains) against the single-process result."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import auto_parallel as ap

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
mesh = ap.ProcessMesh([0, 1], dim_names=["x"])
torch.manual_seed(0)
full = torch.randn(4, 6)

s0 = ap.shard_tensor(full, mesh, [ap.Shard(0)])
assert tuple(s0.shape) == (2, 6)
torch.testing.assert_close(ap.unshard_dtensor(ap.reshard(s0, mesh, [ap.Shard(1)])), full)        # s -> s (all-to-all)
s1 = ap.reshard(s0, mesh, [ap.Shard(1)])
assert tuple(s1.shape) == (4, 3)
torch.testing.assert_close(ap.dtensor_to_local(s1), full[:, 3 * r:3 * r + 3])
rep = ap.reshard(s1, mesh, [ap.Replicate()])                                                    # s -> r
torch.testing.assert_close(ap.dtensor_to_local(rep), full)
part = ap.reshard(rep, mesh, [ap.Partial()])                                                    # r -> p
torch.testing.assert_close(ap.dtensor_to_local(part), full if r == 0 else torch.zeros_like(full))
p = ap.dtensor_from_local(full * (r + 1), mesh, [ap.Partial()])
torch.testing.assert_close(ap.dtensor_to_local(ap.reshard(p, mesh, [ap.Replicate()])), full * 3)                     # p -> r
ps = ap.reshard(ap.dtensor_from_local(full * (r + 1), mesh, [ap.Partial()]), mesh, [ap.Shard(0)])   # p -> s (reduce-scatter)
torch.testing.assert_close(ap.dtensor_to_local(ps), (full * 3)[2 * r:2 * r + 2])
ps1 = ap.reshard(ap.dtensor_from_local(full * (r + 1), mesh, [ap.Partial()]), mesh, [ap.Shard(1)])
torch.testing.assert_close(ap.dtensor_to_local(ps1), (full * 3)[:, 3 * r:3 * r + 3])

# cross-mesh, same status: the shard of rank i moves to rank (1 - i)
other = ap.ProcessMesh([1, 0], dim_names=["x"])
moved = ap.reshard(s0, other, [ap.Shard(0)])
torch.testing.assert_close(ap.dtensor_to_local(moved), full[2 * (1 - r):2 * (1 - r) + 2])

# SPMD rules drive a column-parallel then row-parallel matmul: the middle stays sharded, the end is partial -> replicate
x = torch.randn(5, 8)
w1, w2 = torch.randn(8, 12), torch.randn(12, 4)
dx = ap.shard_tensor(x, mesh, [ap.Replicate()])
dw1 = ap.shard_tensor(w1, mesh, [ap.Shard(1)])
dw2 = ap.shard_tensor(w2, mesh, [ap.Shard(0)])
h = ap.dist_op("matmul", torch.matmul, dx, dw1)
assert h.placements == [ap.Shard(1)] and tuple(h.shape) == (5, 6)
h = ap.dist_op("elementwise", lambda a: torch.relu(a), h)
y = ap.dist_op("matmul", torch.matmul, h, dw2)
assert y.placements[0].is_partial() and tuple(y.shape) == (5, 4)
torch.testing.assert_close(ap.dtensor_to_local(ap.reshard(y, mesh, [ap.Replicate()])), torch.relu(x @ w1) @ w2, rtol=1e-5, atol=1e-5)
# a rule that refuses a sharding reshards the input first: softmax over a sharded dim
sm = ap.dist_op("softmax", lambda a: torch.softmax(a, -1), ap.shard_tensor(full, mesh, [ap.Shard(1)]), axis=-1)
torch.testing.assert_close(ap.unshard_dtensor(sm), torch.softmax(full, -1))
print("auto_parallel ok", r)
