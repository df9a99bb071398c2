"""Worker (4 ranks): pp=2 x dp=2 (or x sharding=2) with the gradient collectives overlapped with the backward pass
(pp_configs.dp_comm_overlap / sharding_comm_overlap) == the same run without overlap == one process on all the data."""
import sys

import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200 import nn
from paddle_b200.distributed import fleet
from paddle_b200.distributed.fleet.meta_parallel import LayerDesc, PipelineLayer

mode = sys.argv[1]          # dp | dp_overlap | sharding | sharding_overlap
ACC, MBS, H, NB = 2, 2, 8, 4
second = "sharding" if mode.startswith("sharding") else "dp"
st = fleet.DistributedStrategy()
st.hybrid_configs = {"dp_degree": 2 if second == "dp" else 1, "mp_degree": 1, "pp_degree": 2,
                     "sharding_degree": 2 if second == "sharding" else 1,
                     "pp_configs": {"dp_comm_overlap": mode == "dp_overlap", "sharding_comm_overlap": mode == "sharding_overlap"}}
st.pipeline_configs = {"accumulate_steps": ACC, "micro_batch_size": MBS}
fleet.init(is_collective=True, strategy=st)
hcg = fleet.get_hybrid_communicate_group()
r = dist.get_rank()
replica = hcg.get_data_parallel_rank() if second == "dp" else hcg.get_sharding_parallel_rank()


class Block(nn.Layer):
    def __init__(self, i):
        super().__init__()
        self.fc = nn.Linear(H, H)

    def forward(self, x):
        return x + torch.tanh(self.fc(x))


class Head(nn.Layer):
    def __init__(self):
        super().__init__()
        self.fc = nn.Linear(H, 3)

    def forward(self, x):
        return self.fc(x)


def loss_fn(out, label):
    return ((out - label) ** 2).mean()


paddle.seed(5)
with paddle.utils.unique_name.guard():
    full = [Block(i) for i in range(NB)] + [Head()]
state = [{k: v.detach().clone() for k, v in l.state_dict().items()} for l in full]
pipe = PipelineLayer([LayerDesc(Block, i) for i in range(NB)] + [LayerDesc(Head)], num_stages=2, loss_fn=loss_fn, seg_method="uniform")
for i, layer in enumerate(pipe.run_function):
    layer.set_state_dict(state[pipe._start_pos + i])
model = fleet.distributed_model(pipe)
opt = fleet.distributed_optimizer(paddle.optimizer.SGD(0.1, parameters=pipe.parameters()))
if mode == "dp_overlap":
    assert model._comm_buffers and not opt._dp_enable, (len(model._comm_buffers), opt._dp_enable)
if mode == "dp":
    assert not model._comm_buffers and opt._dp_enable

ref_opt = paddle.optimizer.SGD(0.1, parameters=[p for l in full for p in l.parameters()])
for step in range(3):
    losses = []
    for rep in range(2):
        g = torch.Generator().manual_seed(100 * step + rep)
        x = torch.randn(ACC * MBS, H, generator=g)
        y = torch.randn(ACC * MBS, 3, generator=g)
        if rep == replica:
            mine = (x, y)
        for mb in range(ACC):
            h = x[mb * MBS:(mb + 1) * MBS]
            for l in full:
                h = l(h)
            l_mb = loss_fn(h, y[mb * MBS:(mb + 1) * MBS]) / ACC / 2          # mean over the two replicas
            l_mb.backward()
    ref_opt.step()
    ref_opt.clear_grad()
    loss = model.train_batch(list(mine), opt)
    assert torch.isfinite(loss).all()
if mode == "sharding_overlap":
    assert model._comm_buffers, "sharding overlap buffers were not registered"
for i, layer in enumerate(pipe.run_function):
    for (k, a), (_, b) in zip(layer.state_dict().items(), full[pipe._start_pos + i].state_dict().items()):
        torch.testing.assert_close(a.detach(), b.detach(), rtol=1e-5, atol=1e-6, msg=lambda m: f"{mode} layer {pipe._start_pos + i} {k}: {m}")
print(f"pipeline overlap {mode} ok", r)
