"""Worker: tiny Llama with context-parallel attention over the fleet sep group (sep=2) == the same model on the
whole sequence; every variant (ulysses / ring / allgather), logits and parameter gradients."""
import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.distributed import fleet
from paddle_b200.models.llama import LlamaConfig, LlamaForCausalLM

st = fleet.DistributedStrategy()
st.hybrid_configs = {"dp_degree": 1, "mp_degree": 1, "pp_degree": 1, "sharding_degree": 1, "sep_degree": 2}
fleet.init(is_collective=True, strategy=st)
r = dist.get_rank()
hcg = fleet.get_hybrid_communicate_group()
assert hcg.get_sep_parallel_world_size() == 2
B, S = 2, 32
torch.manual_seed(3)
tok = torch.randint(0, 128, (B, S))


def build(cp):
    paddle.seed(9)
    cfg = LlamaConfig(vocab_size=128, hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=64, context_parallel=cp)
    with paddle.utils.unique_name.guard():
        return LlamaForCausalLM(cfg)


ref = build("")
full = ref(tok)
full.square().mean().backward()
sl = slice(r * S // 2, (r + 1) * S // 2)
for mode in ("ulysses", "ring", "allgather"):
    m = build(mode)
    out = m(tok[:, sl])
    torch.testing.assert_close(out, full[:, sl].detach(), rtol=1e-4, atol=1e-5, msg=lambda s_: f"{mode}: {s_}")
    (out.square().sum() / full.numel()).backward()
    for (n1, p1), (n2, p2) in zip(m.named_parameters(), ref.named_parameters()):
        g = p1.grad.clone()
        dist.all_reduce(g, group=hcg.get_sep_parallel_group())       # parameters are replicated over sep: sum the halves
        torch.testing.assert_close(g, p2.grad, rtol=1e-3, atol=1e-5, msg=lambda s_: f"{mode} {n1}: {s_}")
print("llama_cp ok", r)
