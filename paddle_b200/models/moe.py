"""Mixtral-style sparse MoE decoder (BASELINE config #5: 8 experts, top-2, expert parallel over the moe group)."""
from __future__ import annotations

from dataclasses import dataclass

from paddle_b200 import nn
from paddle_b200.incubate.distributed.models.moe import MoELayer
from paddle_b200.incubate.nn import functional as IF

from .llama import LlamaAttention, LlamaConfig, LlamaPretrainingCriterion, _init


@dataclass
class MoEConfig(LlamaConfig):
    num_experts: int = 8
    top_k: int = 2
    gate_type: str = "naive"
    # stacked expert weights + device-scheduled grouped GEMMs (one launch for all local experts, no host read of the
    # counts).  None = automatic: on when a rank holds several experts; with a single local expert the "group" is one
    # GEMM and the CTA-pair kernel behind Linear is the faster way to run it.
    grouped_experts: bool | None = None

    @staticmethod
    def mixtral_8x7b(**kw):
        return MoEConfig(hidden_size=4096, intermediate_size=14336, num_hidden_layers=32, num_attention_heads=32,
                         num_key_value_heads=8, **kw)

    @staticmethod
    def tiny(**kw):
        base = dict(vocab_size=512, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
                    max_position_embeddings=128, num_experts=4, top_k=2)
        base.update(kw)
        return MoEConfig(**base)


class ExpertFFN(nn.Layer):
    def __init__(self, config):
        super().__init__()
        attr = _init(config.initializer_range)
        self.gate_up_proj = nn.Linear(config.hidden_size, 2 * config.intermediate_size, weight_attr=attr, bias_attr=False)
        self.down_proj = nn.Linear(config.intermediate_size, config.hidden_size, weight_attr=attr, bias_attr=False)

    def forward(self, x):
        return self.down_proj(IF.swiglu(self.gate_up_proj(x)))


class MoEDecoderLayer(nn.Layer):
    def __init__(self, config: MoEConfig, moe_group=None):
        super().__init__()
        world = moe_group.nranks if moe_group is not None else 1
        assert config.num_experts % world == 0
        self.input_layernorm = nn.RMSNorm(config.hidden_size, config.rms_norm_eps)
        self.self_attn = LlamaAttention(config)
        self.post_attention_layernorm = nn.RMSNorm(config.hidden_size, config.rms_norm_eps)
        n_local = config.num_experts // world
        grouped = config.grouped_experts if config.grouped_experts is not None else n_local > 1
        if grouped:
            from paddle_b200.incubate.distributed.models.moe import GroupedExpertsFFN
            experts = GroupedExpertsFFN(config.num_experts // world, config.hidden_size, config.intermediate_size,
                                        weight_attr=_init(config.initializer_range))
        else:
            experts = nn.LayerList([ExpertFFN(config) for _ in range(config.num_experts // world)])
        self.moe = MoELayer(config.hidden_size, experts, gate={"type": config.gate_type, "top_k": config.top_k},
                            moe_group=moe_group)

    def forward(self, h):
        h = h + self.self_attn(self.input_layernorm(h))
        return h + self.moe(self.post_attention_layernorm(h))


class MoEForCausalLM(nn.Layer):
    def __init__(self, config: MoEConfig, moe_group=None):
        super().__init__()
        attr = _init(config.initializer_range)
        self.embed_tokens = nn.Embedding(config.vocab_size, config.hidden_size, weight_attr=attr)
        self.layers = nn.LayerList([MoEDecoderLayer(config, moe_group) for _ in range(config.num_hidden_layers)])
        self.norm = nn.RMSNorm(config.hidden_size, config.rms_norm_eps)
        self.lm_head = nn.Linear(config.hidden_size, config.vocab_size, weight_attr=attr, bias_attr=False)
        self.criterion = LlamaPretrainingCriterion(config)

    def forward(self, input_ids, labels=None):
        h = self.embed_tokens(input_ids)
        for l in self.layers:
            h = l(h)
        logits = self.lm_head(self.norm(h))
        return logits if labels is None else self.criterion(logits, labels)
