"""Worker: sharded checkpoint written with mp=2 layout, metadata de-duplicates replicated tensors, loaded back into
a different layout (full tensor on every rank)."""
import os
import sys

import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import fleet
from paddle_b200.distributed.fleet import meta_parallel as mpu

st = fleet.DistributedStrategy()
st.hybrid_configs = {"dp_degree": 1, "mp_degree": 2, "pp_degree": 1}
fleet.init(is_collective=True, strategy=st)
r = dist.get_rank()
path = sys.argv[1]
col = mpu.ColumnParallelLinear(8, 12, has_bias=True, gather_output=False)
rep = paddle.nn.Linear(4, 4)
dist.broadcast(rep.weight.detach(), 0)
sd = {"col.weight": col.weight, "col.bias": col.bias, "rep.weight": rep.weight}
dist.save_state_dict(sd, path)
dist.barrier()
files = sorted(os.listdir(path))
assert "0.metadata" in files and "0_0.distcp" in files and "1_0.distcp" in files, files
# rank 1's file holds only its shards of the distributed tensors (replicated tensor de-duplicated to rank 0)
one = paddle.load(os.path.join(path, "1_0.distcp"))
assert "rep.weight" not in one and "col.weight" in one
# load into full (unsharded) targets: every rank reads both shards
parts = []
dist.all_gather(parts, col.weight.detach())
full_w = torch.cat(parts, 1)
tgt = {"col.weight": torch.zeros(8, 12), "col.bias": torch.zeros(12), "rep.weight": torch.zeros(4, 4)}
dist.load_state_dict(tgt, path)
torch.testing.assert_close(tgt["col.weight"], full_w)
torch.testing.assert_close(tgt["rep.weight"], rep.weight.detach())
print("dist_ckpt ok", r)
