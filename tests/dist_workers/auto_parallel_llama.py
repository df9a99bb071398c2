"""Worker (2 ranks, gloo): the repo's Llama under the semi-auto API.  Parameters are sharded Megatron-style with
shard_layer (vocab-parallel embedding, column / row projections, class-sharded lm_head), nothing else is told to the
model: loss and gradients equal the single-process model."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import auto_parallel as ap
from paddle_b200.models import LlamaConfig, LlamaForCausalLM

dist.init_parallel_env()
r = dist.get_rank()
mesh = ap.ProcessMesh([0, 1], dim_names=["mp"])
R_, S = ap.Replicate, ap.Shard
cfg = LlamaConfig(vocab_size=64, hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4,
                  max_position_embeddings=32)


def build():
    paddle.seed(11)
    return LlamaForCausalLM(cfg)


PLAN = {"qkv_proj": (S(1),), "o_proj": (S(0),), "gate_up_proj": (S(1),), "down_proj": (S(0),), "embed_tokens": (S(0),),
        "lm_head": (S(1),)}


def shard_fn(name, layer, m):
    leaf = name.rsplit(".", 1)[-1]
    for pname, p in list(layer._parameters.items()):
        if p is None:
            continue
        pl = PLAN.get(leaf, (R_(),)) if pname == "weight" else (R_(),)
        layer._parameters[pname] = ap.shard_tensor(p, m, list(pl))


ref, net = build(), ap.shard_layer(build(), mesh, shard_fn)
torch.manual_seed(3)
tok = torch.randint(0, cfg.vocab_size, (2, 17))
rl = ref(tok[:, :-1], labels=tok[:, 1:])
rl.backward()
dl = net(ap.shard_tensor(tok[:, :-1], mesh, [R_()]), labels=ap.shard_tensor(tok[:, 1:], mesh, [R_()]))
torch.testing.assert_close(torch.tensor(float(dl)), rl.detach(), rtol=1e-4, atol=1e-4)
dl.backward()
checked = 0
ref_params = dict(ref.named_parameters())
for name, p in net.named_parameters():
    g = ap.unshard_dtensor(ap.dtensor_from_local(p.grad, mesh, p.placements)) if ap.is_dist(p) else p.grad
    torch.testing.assert_close(g, ref_params[name].grad, rtol=2e-3, atol=2e-4, msg=lambda m, n=name: f"{n}: {m}")
    checked += 1
assert checked == len(ref_params) and checked > 10
print("auto_parallel_llama ok", r, "params", checked)
