"""Worker: every collective against a local expectation (gloo, 2 ranks)."""
import numpy as np
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
assert n == 2 and dist.is_initialized() and dist.get_backend() == "GLOO"

for dtype in (torch.float32, torch.bfloat16, torch.int64):
    x = torch.full((4,), float(r + 1)).to(dtype)
    dist.all_reduce(x)
    assert x.float().tolist() == [3.0] * 4, (dtype, x)
x = torch.full((4,), float(r + 1)); dist.all_reduce(x, dist.ReduceOp.MAX); assert x.tolist() == [2.0] * 4
x = torch.full((4,), float(r + 1)); dist.all_reduce(x, dist.ReduceOp.AVG); assert x.tolist() == [1.5] * 4
t = dist.all_reduce(torch.ones(2), sync_op=False); t.wait(); assert t.is_completed()

b = torch.arange(3.0) if r == 1 else torch.zeros(3)
dist.broadcast(b, src=1); assert b.tolist() == [0, 1, 2]

x = torch.full((2,), float(r + 1)); dist.reduce(x, dst=0)
if r == 0:
    assert x.tolist() == [3, 3]

lst = []
dist.all_gather(lst, torch.full((2,), float(r))); assert [t.tolist() for t in lst] == [[0, 0], [1, 1]]
out = torch.empty(4); dist.stream.all_gather(out, torch.full((2,), float(r))); assert out.tolist() == [0, 0, 1, 1]
objs = []
dist.all_gather_object(objs, {"rank": r}); assert objs == [{"rank": 0}, {"rank": 1}]

o = torch.empty(2)
dist.reduce_scatter(o, [torch.full((2,), float(r + 1)), torch.full((2,), float(10 * (r + 1)))])
assert o.tolist() == ([3, 3] if r == 0 else [30, 30])

outs = []
dist.alltoall(outs, [torch.full((2,), float(10 * r + j)) for j in range(2)])
assert [t.tolist() for t in outs] == [[float(r)] * 2, [float(10 + r)] * 2]
o = torch.empty(4); dist.alltoall_single(o, torch.arange(4.0) + 10 * r); assert o.tolist() == ([0, 1, 10, 11] if r == 0 else [2, 3, 12, 13])

s = torch.empty(2)
dist.scatter(s, [torch.zeros(2), torch.ones(2)] if r == 0 else None, src=0); assert s.tolist() == [float(r)] * 2
gl = []
dist.gather(torch.full((1,), float(r)), gl if r == 0 else None, dst=0)
if r == 0:
    assert [t.item() for t in gl] == [0.0, 1.0]

if r == 0:
    dist.send(torch.arange(3.0), dst=1)
else:
    m = torch.empty(3); dist.recv(m, src=0); assert m.tolist() == [0, 1, 2]
peer = 1 - r
snd, rcv = torch.full((2,), float(r)), torch.empty(2)
for t in dist.batch_isend_irecv([dist.P2POp(dist.isend, snd, peer), dist.P2POp(dist.irecv, rcv, peer)]):
    t.wait()
assert rcv.tolist() == [float(peer)] * 2

ol = ["x", None]
dist.broadcast_object_list(ol, src=0) if r == 0 else dist.broadcast_object_list(ol := [None, None], src=0)
assert ol[0] == "x"

g = dist.new_group([0, 1]); assert g.id >= 10 and g.rank == r
g1 = dist.new_group([1])
assert (g1.rank == 0) if r == 1 else (g1.rank == -1)
dist.barrier()
try:
    dist.stream.all_reduce(torch.ones(1), sync_op=False, use_calc_stream=True)
    raise SystemExit("expected RuntimeError")
except RuntimeError:
    pass
print("collectives ok", r)
