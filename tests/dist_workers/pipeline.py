"""Worker: pipeline parallel (pp=2) loss/params == non-pipelined model (SURVEY §4.1 PP tests); 1F1B and interleaved."""
import sys

import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200 import nn
from paddle_b200.distributed import fleet
from paddle_b200.distributed.fleet.meta_parallel import LayerDesc, PipelineLayer

mode = sys.argv[1]          # 1f1b | vpp
ACC, MBS = 4, 2
st = fleet.DistributedStrategy()
st.hybrid_configs = {"dp_degree": 1, "mp_degree": 1, "pp_degree": 2}
st.pipeline_configs = {"accumulate_steps": ACC, "micro_batch_size": MBS}
fleet.init(is_collective=True, strategy=st)
hcg = fleet.get_hybrid_communicate_group()
r = dist.get_rank()
H = 16


class Block(nn.Layer):
    def __init__(self, i):
        super().__init__()
        self.fc = nn.Linear(H, H)

    def forward(self, x):
        return x + torch.tanh(self.fc(x))


class Head(nn.Layer):
    def __init__(self):
        super().__init__()
        self.fc = nn.Linear(H, 4)

    def forward(self, x):
        return self.fc(x)


def loss_fn(out, label):
    return ((out - label) ** 2).mean()


NB = 4
paddle.seed(21)
with paddle.utils.unique_name.guard():
    full = [Block(i) for i in range(NB)] + [Head()]
state = [{k: v.detach().clone() for k, v in l.state_dict().items()} for l in full]

descs = [LayerDesc(Block, i) for i in range(NB)] + [LayerDesc(Head)]
kw = dict(num_virtual_pipeline_stages=2, seg_method="layer:Block") if mode == "vpp" else dict(seg_method="uniform")
pipe = PipelineLayer(descs, num_stages=2, loss_fn=loss_fn, **kw)
# load the reference weights into whatever layers this stage owns
if mode == "vpp":
    for v, chunk in enumerate(pipe._model_chunks):
        for i, layer in enumerate(chunk.get_run_function()):
            layer.set_state_dict(state[pipe._start_poss[v] + i])
else:
    for i, layer in enumerate(pipe.run_function):
        layer.set_state_dict(state[pipe._start_pos + i])
model = fleet.distributed_model(pipe)
opt = fleet.distributed_optimizer(paddle.optimizer.SGD(0.1, parameters=pipe.parameters()))
assert type(model).__name__ == ("PipelineParallelWithInterleave" if mode == "vpp" else "PipelineParallel"), type(model)

ref_opt = paddle.optimizer.SGD(0.1, parameters=[p for l in full for p in l.parameters()])
for step in range(3):
    g = torch.Generator().manual_seed(step)
    x = torch.randn(ACC * MBS, H, generator=g)
    y = torch.randn(ACC * MBS, 4, generator=g)
    loss = model.train_batch([x, y], opt)
    ref_loss = 0
    for mb in range(ACC):
        h = x[mb * MBS:(mb + 1) * MBS]
        for l in full:
            h = l(h)
        l_mb = loss_fn(h, y[mb * MBS:(mb + 1) * MBS]) / ACC
        l_mb.backward()
        ref_loss += float(l_mb)
    ref_opt.step()
    ref_opt.clear_grad()
    assert abs(float(loss) - ref_loss) < 1e-5, (step, float(loss), ref_loss)
# parameters of the owned layers follow the reference
owned = []
if mode == "vpp":
    for v, chunk in enumerate(pipe._model_chunks):
        owned += [(pipe._start_poss[v] + i, l) for i, l in enumerate(chunk.get_run_function())]
else:
    owned = [(pipe._start_pos + i, l) for i, l in enumerate(pipe.run_function)]
for idx, layer in owned:
    for (k, a), (_, b) in zip(layer.state_dict().items(), full[idx].state_dict().items()):
        torch.testing.assert_close(a.detach(), b.detach(), rtol=1e-5, atol=1e-6)
sched = model.get_static_scheduler()
if mode == "1f1b":
    assert sched == ("f0;f1;b0;f2;b1;f3;b2;b3;" if r == 0 else "f0;b0;f1;b1;f2;b2;f3;b3;"), sched
ev = model.eval_batch([x, y], compute_loss=True)
assert torch.isfinite(ev).all()
print(f"pipeline {mode} ok", r)
