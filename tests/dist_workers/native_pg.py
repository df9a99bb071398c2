"""Worker (>= 2 GPUs): the native NCCL process group (own communicators, comm stream, event tasks) against
torch.distributed on the same data, through the raw class and through paddle.distributed with the routing flag."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import native_pg

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", torch.cuda.current_device())
assert native_pg.available()
g = dist.collective._get_default_group()
pg = native_pg.NativeProcessGroup.for_group(g)
torch.manual_seed(10 + r)

for dtype in (torch.float32, torch.bfloat16):
    x = torch.randn(1 << 16, device=dev).to(dtype)
    ref = x.clone()
    dist.all_reduce(ref)
    y = x.clone()
    t = pg.all_reduce(y, dist.ReduceOp.SUM, sync_op=False)        # on the group's comm stream
    assert t.wait()
    torch.testing.assert_close(y.float(), ref.float(), rtol=2e-2, atol=2e-2)
    z = x.clone()
    pg.all_reduce(z, dist.ReduceOp.AVG, sync_op=True, use_calc_stream=True)
    torch.testing.assert_close(z.float(), ref.float() / n, rtol=2e-2, atol=2e-2)

b = torch.full((1024,), float(r), device=dev)
pg.broadcast(b, src=n - 1)
assert float(b[0]) == n - 1
out = torch.empty(n * 256, device=dev)
pg.all_gather(out, torch.full((256,), float(r), device=dev))
assert out.view(n, 256)[:, 0].tolist() == [float(i) for i in range(n)]
rs = torch.empty(128, device=dev)
pg.reduce_scatter(rs, torch.arange(n * 128, device=dev, dtype=torch.float32), dist.ReduceOp.SUM)
torch.testing.assert_close(rs, torch.arange(r * 128, (r + 1) * 128, device=dev, dtype=torch.float32) * n)
a2a = torch.empty(n * 64, device=dev)
pg.all_to_all(a2a, torch.arange(n * 64, device=dev, dtype=torch.float32) + 1000 * r)
exp = torch.cat([torch.arange(r * 64, (r + 1) * 64, device=dev, dtype=torch.float32) + 1000 * p for p in range(n)])
torch.testing.assert_close(a2a, exp)
red = torch.ones(32, device=dev) * (r + 1)
pg.reduce(red, dst=0, op=dist.ReduceOp.SUM)
if r == 0:
    assert float(red[0]) == n * (n + 1) / 2
if r == 0:
    pg.send(torch.full((16,), 42.0, device=dev), dst=1)
elif r == 1:
    got = torch.empty(16, device=dev)
    pg.recv(got, src=0)
    assert float(got[3]) == 42.0
torch.cuda.synchronize()

# the async task really is asynchronous: is_completed flips without anybody calling wait()
big = torch.randn(1 << 24, device=dev)
t = pg.all_reduce(big, dist.ReduceOp.SUM, sync_op=False)
torch.cuda.synchronize()
assert t.is_completed()
t.wait()

# routing flag: paddle.distributed.* goes through the native group
paddle.set_flags({"FLAGS_use_native_process_group": True})
w = torch.ones(4096, device=dev) * (r + 1)
dist.all_reduce(w)
assert float(w[0]) == n * (n + 1) / 2
lst = torch.empty(n * 8, device=dev)
dist.all_gather(lst, torch.full((8,), float(r), device=dev))
assert lst.view(n, 8)[:, 0].tolist() == [float(i) for i in range(n)]
paddle.set_flags({"FLAGS_use_native_process_group": False})
torch.cuda.synchronize()
native_pg.shutdown()
print("native_pg ok", r, flush=True)
