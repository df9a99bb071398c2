"""Worker: DataParallel (bucketed reducer) == single-process training on the concatenated batch; no_sync; unused params."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200 import nn
from paddle_b200.models import GPTConfig, GPTForCausalLM

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
cfg = GPTConfig(vocab_size=64, hidden_size=256, num_hidden_layers=2, num_attention_heads=4, max_position_embeddings=32)


def build():
    paddle.seed(5)
    with paddle.utils.unique_name.guard():
        return GPTForCausalLM(cfg)


torch.manual_seed(0)
data = torch.randint(0, 64, (4, 17))          # global batch 4, 2 per rank
ref = build()
ropt = paddle.optimizer.SGD(0.05, parameters=ref.parameters())   # SGD: linear in the grads (Adam would amplify rounding noise of ~0 grads)
m = build()
dp = paddle.DataParallel(m, comm_buffer_size=1, last_comm_buffer_size=0.25)     # small buckets: several per step
assert len(dp._reducer.buckets) > 2
opt = paddle.optimizer.SGD(0.05, parameters=m.parameters())
for step in range(3):
    loss = dp(data[2 * r:2 * r + 2, :-1], labels=data[2 * r:2 * r + 2, 1:])
    loss.backward(); opt.step(); opt.clear_grad()
    rl = ref(data[:, :-1], labels=data[:, 1:])
    rl.backward(); ropt.step(); ropt.clear_grad()
for (k, a), (_, b) in zip(m.state_dict().items(), ref.state_dict().items()):
    torch.testing.assert_close(a.detach(), b.detach(), rtol=2e-4, atol=2e-5, msg=lambda s: f"{k}: {s}")

# no_sync: grads stay local
with dp.no_sync():
    dp(data[2 * r:2 * r + 2, :-1], labels=data[2 * r:2 * r + 2, 1:]).backward()
g = m.lm_head.weight.grad.clone()
parts = []
dist.all_gather(parts, g)
assert not torch.allclose(parts[0], parts[1])
opt.clear_grad()


class Two(nn.Layer):
    def __init__(self):
        super().__init__()
        self.a, self.b = nn.Linear(4, 4), nn.Linear(4, 4)

    def forward(self, x):
        return self.a(x)          # self.b unused


paddle.seed(1)
t = paddle.DataParallel(Two(), find_unused_parameters=True)
t(torch.ones(2, 4) * (r + 1)).sum().backward()
parts = []
dist.all_gather(parts, t._layers.a.weight.grad)
torch.testing.assert_close(parts[0], parts[1])
t2 = paddle.DataParallel(Two(), find_unused_parameters=False)
try:
    t2(torch.ones(2, 4)).sum().backward()
    raise SystemExit("expected unused-parameter error")
except RuntimeError as e:
    assert "find_unused_parameters" in str(e)

# unused-parameter bitmap: `b` is unused on EVERY rank -> its gradient stays None (not zeros); a parameter used on
# one rank only still takes part in the all-reduce everywhere
assert t._layers.b.weight.grad is None and t._layers.a.weight.grad is not None


class Sometimes(nn.Layer):
    def __init__(self):
        super().__init__()
        self.a, self.b = nn.Linear(4, 4), nn.Linear(4, 4)

    def forward(self, x, use_b):
        y = self.a(x)
        return y + self.b(x) if use_b else y


paddle.seed(2)
s = paddle.DataParallel(Sometimes(), find_unused_parameters=True)
s(torch.ones(2, 4), use_b=(r == 0)).sum().backward()
parts = []
dist.all_gather(parts, s._layers.b.weight.grad)
torch.testing.assert_close(parts[0], parts[1])            # rank 1 contributed zeros, both hold the average
assert float(parts[0].abs().sum()) > 0

# sparse gradients (Embedding(sparse=True)): rows are gathered and coalesced, result == dense all-reduce
paddle.seed(4)


class Emb(nn.Layer):
    def __init__(self, sparse):
        super().__init__()
        self.e = nn.Embedding(50, 8, sparse=sparse)
        self.fc = nn.Linear(8, 1)

    def forward(self, ids):
        return self.fc(self.e(ids)).sum()


with paddle.utils.unique_name.guard():
    es = Emb(True)
with paddle.utils.unique_name.guard():
    ed = Emb(False)
ed.set_state_dict(es.state_dict())
ds, dd = paddle.DataParallel(es), paddle.DataParallel(ed)
assert any(b.sparse for b in ds._reducer.buckets)
from paddle_b200 import lib as _lib
if _lib.runtime.available() and hasattr(_lib.runtime.require(), "pdb200_reducer_create"):
    assert ds._reducer._core, "the reducer's bookkeeping must run in libpdb200_runtime.so when it is built"
ids = torch.tensor([[1, 7, 7], [3 + r, 9, 1]])
ds(ids).backward()
dd(ids).backward()
assert es.e.weight.grad.is_sparse
torch.testing.assert_close(es.e.weight.grad.to_dense(), ed.e.weight.grad, rtol=1e-5, atol=1e-6)
torch.testing.assert_close(es.fc.weight.grad, ed.fc.weight.grad, rtol=1e-5, atol=1e-6)
print("data_parallel ok", r)
