"""Worker: shard-vs-whole oracle for Column/Row/Vocab parallel layers, ParallelCrossEntropy and sequence parallel
(SURVEY §4.2.2), mp=2 on gloo."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import fleet
from paddle_b200.distributed.fleet import meta_parallel as mpu
from paddle_b200.distributed.fleet.utils import sequence_parallel_utils as spu

st = fleet.DistributedStrategy()
st.hybrid_configs = {"dp_degree": 1, "mp_degree": 2, "pp_degree": 1}
fleet.init(is_collective=True, strategy=st)
hcg = fleet.get_hybrid_communicate_group()
r, n = hcg.get_model_parallel_rank(), 2
g = hcg.get_model_parallel_group()
assert hcg.get_parallel_mode() == 1 and hcg.get_model_parallel_world_size() == 2


def gather(t, dim):
    parts = []
    dist.all_gather(parts, t.detach().contiguous(), g)
    return torch.cat(parts, dim)


torch.manual_seed(3)
x = torch.randn(6, 16)
dist.broadcast(x, 0)

# ---- Column (gather_output) : whole weight = concat of shards on dim 1
col = mpu.ColumnParallelLinear(16, 8, has_bias=True, gather_output=True)
assert tuple(col.weight.shape) == (16, 4) and col.weight.is_distributed and col.weight.split_axis == 1
W, B = gather(col.weight, 1), gather(col.bias, 0)
xi = x.clone().requires_grad_(True)
y = col(xi); y.pow(2).sum().backward()
xr = x.clone().requires_grad_(True)
yr = xr @ W + B; yr.pow(2).sum().backward()
torch.testing.assert_close(y, yr, rtol=1e-5, atol=1e-5)
torch.testing.assert_close(xi.grad, xr.grad, rtol=1e-5, atol=1e-5)

# ---- Row (input not parallel): whole weight = concat on dim 0, bias replicated
row = mpu.RowParallelLinear(16, 8, has_bias=True, input_is_parallel=False)
dist.broadcast(row.bias.detach(), 0)
W = gather(row.weight, 0)
xi = x.clone().requires_grad_(True)
y = row(xi); y.pow(2).sum().backward()
xr = x.clone().requires_grad_(True)
yr = xr @ W + row.bias.detach(); yr.pow(2).sum().backward()
torch.testing.assert_close(y, yr, rtol=1e-5, atol=1e-5)
torch.testing.assert_close(xi.grad, xr.grad, rtol=1e-5, atol=1e-5)

# ---- Vocab embedding
emb = mpu.VocabParallelEmbedding(20, 8)
E = gather(emb.weight, 0)
ids = torch.randint(0, 20, (3, 5)); dist.broadcast(ids, 0)
torch.testing.assert_close(emb(ids), torch.nn.functional.embedding(ids, E))

# ---- ParallelCrossEntropy on sharded logits
logits_full = torch.randn(7, 20); dist.broadcast(logits_full, 0)
labels = torch.randint(0, 20, (7, 1)); labels[2] = -100; dist.broadcast(labels, 0)
shard = logits_full[:, r * 10:(r + 1) * 10].clone().requires_grad_(True)
loss = mpu.ParallelCrossEntropy(ignore_index=-100)(shard, labels)
assert tuple(loss.shape) == (7, 1)
w = torch.arange(1.0, 8.0)[:, None]
(loss * w).sum().backward()
lf = logits_full.clone().requires_grad_(True)
ref = torch.nn.functional.cross_entropy(lf, labels.squeeze(-1), ignore_index=-100, reduction="none")[:, None]
(ref * w).sum().backward()
torch.testing.assert_close(loss, ref, rtol=1e-5, atol=1e-5)
torch.testing.assert_close(shard.grad, lf.grad[:, r * 10:(r + 1) * 10], rtol=1e-5, atol=1e-6)

# ---- sequence parallel: [s/n, b, h] -> ColumnSP (AG->GEMM) -> RowSP (GEMM->RS)
for overlap in (False, True):
    st.hybrid_configs = {"mp_configs": {"mp_async_allreduce": overlap}}
    csp = spu.ColumnSequenceParallelLinear(16, 8, has_bias=True, gather_output=False)
    rsp = spu.RowSequenceParallelLinear(8, 16, has_bias=True, input_is_parallel=True)
    dist.broadcast(rsp.bias.detach(), 0)
    W1, B1, W2 = gather(csp.weight, 1), gather(csp.bias, 0), gather(rsp.weight, 0)
    xs = torch.randn(8, 2, 16); dist.broadcast(xs, 0)           # [s, b, h]
    local = spu.ScatterOp.apply(xs.clone().requires_grad_(True))
    local.retain_grad()
    out = rsp(torch.relu(csp(local)))
    out.pow(2).sum().backward()
    xr = xs.clone().requires_grad_(True)
    outr = torch.relu(xr @ W1 + B1) @ W2 + rsp.bias.detach()
    outr.pow(2).sum().backward()
    torch.testing.assert_close(gather(out, 0), outr, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(gather(local.grad, 0), xr.grad, rtol=1e-4, atol=1e-5)

# ---- mp_skip_c_identity: only the identity ops disappear, every all-reduce stays (results equal the whole layer)
st.hybrid_configs = {"mp_configs": {"mp_async_allreduce": False, "mp_skip_c_identity": True}}
col = mpu.ColumnParallelLinear(16, 8, has_bias=False, gather_output=False)
row = mpu.RowParallelLinear(8, 16, has_bias=False, input_is_parallel=True)
assert col.mp_skip_c_identity and row.mp_skip_c_identity
W1, W2 = gather(col.weight, 1), gather(row.weight, 0)
xi = x.clone().requires_grad_(True)
y = row(torch.tanh(col(xi))); y.pow(2).sum().backward()
xr = x.clone().requires_grad_(True)
yr = torch.tanh(xr @ W1) @ W2; yr.pow(2).sum().backward()
torch.testing.assert_close(y, yr, rtol=1e-5, atol=1e-5)           # Row forward still sums the partial products
torch.testing.assert_close(xi.grad, xr.grad, rtol=1e-5, atol=1e-5)  # Column backward still sums the partial dX
st.hybrid_configs = {"mp_configs": {"mp_skip_c_identity": False}}

# ---- RNG tracker: sharded init differs, replicated init matches
a = mpu.ColumnParallelLinear(8, 8, has_bias=False, gather_output=False).weight.detach()
parts = []; dist.all_gather(parts, a); assert not torch.allclose(parts[0], parts[1])
b = paddle.nn.Linear(4, 4).weight.detach()
parts = []; dist.all_gather(parts, b); assert torch.allclose(parts[0], parts[1])
print("tp_layers ok", r)
