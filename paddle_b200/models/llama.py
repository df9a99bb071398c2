"""Llama-2 family on the Paddle API (flagship model of BASELINE.json: 7B stage-3 sharding, 13B mp x pp).

Architecture follows the public Llama-2 definition (RMSNorm pre-norm, rotary embeddings on q/k, SwiGLU MLP, causal
attention, untied lm_head).  Every hot op is this framework's own kernel: fused_rms_norm with residual-add
(K18/K19), fused rotary embedding (K20), flash attention (K16/K17), SwiGLU on the packed gate|up projection (K21),
fused softmax-cross-entropy (K8), fused multi-tensor AdamW (K12).  Q/K/V and gate/up projections are single fused
GEMMs.  With ``tensor_parallel_degree > 1`` the projections become Column/RowParallelLinear (or their
sequence-parallel variants), the embedding VocabParallelEmbedding and the loss ParallelCrossEntropy.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

import paddle_b200 as paddle
from paddle_b200 import nn
from paddle_b200.incubate.nn import functional as IF
from paddle_b200.nn import functional as F


@dataclass
class LlamaConfig:
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int | None = None
    max_position_embeddings: int = 4096
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    initializer_range: float = 0.02
    tensor_parallel_degree: int = 1
    sequence_parallel: bool = False
    context_parallel: str = ""          # "", "ulysses", "ring" or "allgather": attention over the fleet sep group
    recompute: bool = False
    recompute_granularity: str = "full"
    use_flash_attention: bool = True
    dtype: str = "float32"

    def __post_init__(self):
        if self.num_key_value_heads is None:
            self.num_key_value_heads = self.num_attention_heads

    @property
    def head_dim(self):
        return self.hidden_size // self.num_attention_heads

    @staticmethod
    def llama2_7b(**kw):
        return LlamaConfig(hidden_size=4096, intermediate_size=11008, num_hidden_layers=32, num_attention_heads=32, **kw)

    @staticmethod
    def llama2_13b(**kw):
        return LlamaConfig(hidden_size=5120, intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40, **kw)

    @staticmethod
    def tiny(**kw):
        base = dict(vocab_size=512, hidden_size=128, intermediate_size=384, num_hidden_layers=2, num_attention_heads=4,
                    max_position_embeddings=256)
        base.update(kw)
        return LlamaConfig(**base)

    def num_parameters(self):
        h, i, v, L = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_hidden_layers
        kv = self.num_key_value_heads * self.head_dim
        per_layer = h * (h + 2 * kv) + h * h + 3 * h * i + 2 * h
        return L * per_layer + 2 * v * h + h

    def flops_per_token(self, seq_len):
        """Training FLOPs per token: 6 x matmul params + causal attention (fwd 2*2*s*h/2, x3 for fwd+bwd)."""
        h, i, v, L = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_hidden_layers
        kv = self.num_key_value_heads * self.head_dim
        mm = L * (h * (h + 2 * kv) + h * h + 3 * h * i) + v * h
        attn = L * 2 * 2 * seq_len * h / 2
        return 6 * mm + 3 * attn


def _init(std):
    return paddle.ParamAttr(initializer=nn.initializer.Normal(0.0, std))


def _linear_cls(config, kind):
    if config.tensor_parallel_degree <= 1:
        return None
    from paddle_b200.distributed.fleet.meta_parallel import ColumnParallelLinear, RowParallelLinear
    from paddle_b200.distributed.fleet.utils.sequence_parallel_utils import (ColumnSequenceParallelLinear,
                                                                           RowSequenceParallelLinear)
    if config.sequence_parallel:
        return ColumnSequenceParallelLinear if kind == "col" else RowSequenceParallelLinear
    return ColumnParallelLinear if kind == "col" else RowParallelLinear


class LlamaAttention(nn.Layer):
    def __init__(self, config: LlamaConfig):
        super().__init__()
        self.config = config
        tp = max(config.tensor_parallel_degree, 1)
        self.num_heads = config.num_attention_heads // tp
        self.num_kv_heads = config.num_key_value_heads // tp
        self.head_dim = config.head_dim
        h = config.hidden_size
        qkv_out = (config.num_attention_heads + 2 * config.num_key_value_heads) * self.head_dim
        attr = _init(config.initializer_range)
        Col, Row = _linear_cls(config, "col"), _linear_cls(config, "row")
        if Col is None:
            self.qkv_proj = nn.Linear(h, qkv_out, weight_attr=attr, bias_attr=False)
            self.o_proj = nn.Linear(h, h, weight_attr=attr, bias_attr=False)
        else:
            self.qkv_proj = Col(h, qkv_out, weight_attr=attr, has_bias=False, gather_output=False)
            self.o_proj = Row(h, h, weight_attr=attr, has_bias=False, input_is_parallel=True)

    def forward(self, x, position_ids=None):
        """x: [b, s, h]  (sequence parallel: [s/n, b, h] -> projections gather to [s, b, ...])."""
        sp = self.config.sequence_parallel and self.config.tensor_parallel_degree > 1
        qkv = self.qkv_proj(x)
        if sp:
            qkv = qkv.transpose(0, 1)        # [s, b, .] -> [b, s, .]
        b, s = qkv.shape[0], qkv.shape[1]
        nq, nkv, d = self.num_heads, self.num_kv_heads, self.head_dim
        cp_group = _sep_group() if self.config.context_parallel else None
        if cp_group is not None and cp_group.nranks > 1:
            # this rank holds tokens [rank*s, (rank+1)*s) of the sequence: rotate with the global positions
            sin, cos = _rope_tables(s * cp_group.nranks, d, self.config.rope_theta, qkv.device)
            if position_ids is None:
                position_ids = (torch.arange(s, device=qkv.device) + cp_group.rank * s)[None].expand(b, s)
            q, k, v = paddle.ops.qkv_split_rope(qkv, nq, nkv, sin, cos, position_ids, use_neox_rotary_style=False)
            from paddle_b200.distributed.fleet.utils import context_parallel as cp
            fn = {"ulysses": cp.ulysses_attention, "ring": cp.ring_flash_attention,
                  "allgather": cp.allgather_kv_attention}[self.config.context_parallel]
            out = fn(q, k, v, cp_group, causal=True)
        else:
            sin, cos = _rope_tables(s, d, self.config.rope_theta, qkv.device)
            q, k, v = paddle.ops.qkv_split_rope(qkv, nq, nkv, sin, cos, position_ids, use_neox_rotary_style=False)
            out, _ = F.flash_attention(q, k, v, causal=True, training=self.training)
        out = out.reshape(b, s, nq * d)
        if sp:
            out = out.transpose(0, 1).contiguous()
        return self.o_proj(out)


def _sep_group():
    """The context-parallel (sep) group of the current fleet topology, if any."""
    try:
        from paddle_b200.distributed import fleet
        hcg = fleet.get_hybrid_communicate_group()
        return hcg.get_sep_parallel_group() if hcg is not None and hcg.get_sep_parallel_world_size() > 1 else None
    except Exception:  # noqa: BLE001  (fleet not initialised)
        return None


_rope_cache = {}


def _rope_tables(seq, d, theta, device):
    key = (seq, d, float(theta), str(device))
    if key not in _rope_cache:
        inv = theta ** (-torch.arange(0, d, 2, dtype=torch.float32, device=device) / d)
        ang = torch.arange(seq, dtype=torch.float32, device=device)[:, None] * inv[None]
        ang = torch.cat([ang, ang], -1)           # rotate-half layout (HF Llama)
        _rope_cache[key] = (ang.sin().reshape(1, seq, 1, d), ang.cos().reshape(1, seq, 1, d))
    return _rope_cache[key]


def _rope(q, k, position_ids, theta):
    s, d = q.shape[1], q.shape[-1]
    sin, cos = _rope_tables(s, d, theta, q.device)
    return IF.fused_rotary_position_embedding(q, k, None, sin=sin, cos=cos, position_ids=position_ids,
                                              use_neox_rotary_style=False)


class LlamaMLP(nn.Layer):
    def __init__(self, config: LlamaConfig):
        super().__init__()
        h, i = config.hidden_size, config.intermediate_size
        attr = _init(config.initializer_range)
        Col, Row = _linear_cls(config, "col"), _linear_cls(config, "row")
        self.tp = max(config.tensor_parallel_degree, 1)
        if Col is None:
            self.gate_up_proj = nn.Linear(h, 2 * i, weight_attr=attr, bias_attr=False)
            self.down_proj = nn.Linear(i, h, weight_attr=attr, bias_attr=False)
        else:
            # per-rank layout [gate_shard | up_shard] so swiglu splits the local output in halves
            self.gate_up_proj = Col(h, 2 * i, weight_attr=attr, has_bias=False, gather_output=False)
            self.down_proj = Row(i, h, weight_attr=attr, has_bias=False, input_is_parallel=True)

    def forward(self, x):
        return self.down_proj(IF.swiglu(self.gate_up_proj(x)))


class LlamaDecoderLayer(nn.Layer):
    def __init__(self, config: LlamaConfig, layer_idx=0):
        super().__init__()
        self.config = config
        self.input_layernorm = nn.RMSNorm(config.hidden_size, config.rms_norm_eps)
        self.self_attn = LlamaAttention(config)
        self.post_attention_layernorm = nn.RMSNorm(config.hidden_size, config.rms_norm_eps)
        self.mlp = LlamaMLP(config)
        if config.sequence_parallel and config.tensor_parallel_degree > 1:
            from paddle_b200.distributed.fleet.utils.sequence_parallel_utils import mark_as_sequence_parallel_parameter
            mark_as_sequence_parallel_parameter(self.input_layernorm.weight)
            mark_as_sequence_parallel_parameter(self.post_attention_layernorm.weight)

    def _body(self, hidden):
        normed = self.input_layernorm(hidden)
        attn = self.self_attn(normed)
        # residual add fused into the norm kernel: normed2 = norm(attn + hidden), resid = attn + hidden
        normed2, resid = IF.fused_rms_norm(attn, self.post_attention_layernorm.weight, None,
                                           self.post_attention_layernorm._epsilon, attn.dim() - 1, residual=hidden)
        return resid + self.mlp(normed2)

    def forward(self, hidden):
        if self.config.recompute and self.training and torch.is_grad_enabled():
            from paddle_b200.distributed.fleet.recompute import recompute
            return recompute(self._body, hidden)
        return self._body(hidden)


class LlamaModel(nn.Layer):
    def __init__(self, config: LlamaConfig):
        super().__init__()
        self.config = config
        attr = _init(config.initializer_range)
        if config.tensor_parallel_degree > 1:
            from paddle_b200.distributed.fleet.meta_parallel import VocabParallelEmbedding
            self.embed_tokens = VocabParallelEmbedding(config.vocab_size, config.hidden_size, weight_attr=attr)
        else:
            self.embed_tokens = nn.Embedding(config.vocab_size, config.hidden_size, weight_attr=attr)
        self.layers = nn.LayerList([LlamaDecoderLayer(config, i) for i in range(config.num_hidden_layers)])
        self.norm = nn.RMSNorm(config.hidden_size, config.rms_norm_eps)
        if config.sequence_parallel and config.tensor_parallel_degree > 1:
            from paddle_b200.distributed.fleet.utils.sequence_parallel_utils import mark_as_sequence_parallel_parameter
            mark_as_sequence_parallel_parameter(self.norm.weight)

    def forward(self, input_ids):
        h = self.embed_tokens(input_ids)
        sp = self.config.sequence_parallel and self.config.tensor_parallel_degree > 1
        if sp:
            from paddle_b200.distributed.fleet.utils.sequence_parallel_utils import ScatterOp
            h = ScatterOp.apply(h.transpose(0, 1).contiguous())      # [b, s, h] -> [s/n, b, h]
        for layer in self.layers:
            h = layer(h)
        h = self.norm(h)
        if sp:
            from paddle_b200.distributed.fleet.utils.sequence_parallel_utils import GatherOp
            h = GatherOp.apply(h).transpose(0, 1).contiguous()
        return h


class LlamaPretrainingCriterion(nn.Layer):
    """Mean token cross-entropy; vocab-parallel when the logits are sharded."""

    def __init__(self, config: LlamaConfig, ignore_index=-100):
        super().__init__()
        self.ignore_index = ignore_index
        self.tp = config.tensor_parallel_degree > 1
        if self.tp:
            from paddle_b200.distributed.fleet.meta_parallel import ParallelCrossEntropy
            self.loss_func = ParallelCrossEntropy(ignore_index=ignore_index)

    def forward(self, logits, labels):
        if self.tp:
            loss = self.loss_func(logits, labels.unsqueeze(-1)).squeeze(-1)
        else:
            loss = paddle.ops.softmax_cross_entropy(logits, labels, self.ignore_index)
        valid = (labels != self.ignore_index)
        return loss.sum() / valid.sum().clamp(min=1).to(loss.dtype)


class LlamaForCausalLM(nn.Layer):
    def __init__(self, config: LlamaConfig):
        super().__init__()
        self.config = config
        self.llama = LlamaModel(config)
        attr = _init(config.initializer_range)
        if config.tensor_parallel_degree > 1:
            from paddle_b200.distributed.fleet.meta_parallel import ColumnParallelLinear
            self.lm_head = ColumnParallelLinear(config.hidden_size, config.vocab_size, weight_attr=attr, has_bias=False,
                                                gather_output=False)
        else:
            self.lm_head = nn.Linear(config.hidden_size, config.vocab_size, weight_attr=attr, bias_attr=False)
        self.criterion = LlamaPretrainingCriterion(config)

    def forward(self, input_ids, labels=None):
        h = self.llama(input_ids)
        logits = self.lm_head(h)
        if labels is None:
            return logits
        return self.criterion(logits, labels)


# ----------------------------------------------------------------------------------------- pipeline variant
class _EmbeddingPipe(nn.Layer):
    def __init__(self, config):
        super().__init__()
        attr = _init(config.initializer_range)
        if config.tensor_parallel_degree > 1:
            from paddle_b200.distributed.fleet.meta_parallel import VocabParallelEmbedding
            self.embed_tokens = VocabParallelEmbedding(config.vocab_size, config.hidden_size, weight_attr=attr)
        else:
            self.embed_tokens = nn.Embedding(config.vocab_size, config.hidden_size, weight_attr=attr)
        self.sp = config.sequence_parallel and config.tensor_parallel_degree > 1

    def forward(self, input_ids):
        h = self.embed_tokens(input_ids)
        if self.sp:
            from paddle_b200.distributed.fleet.utils.sequence_parallel_utils import ScatterOp
            h = ScatterOp.apply(h.transpose(0, 1).contiguous())
        return h


class _HeadPipe(nn.Layer):
    def __init__(self, config):
        super().__init__()
        self.norm = nn.RMSNorm(config.hidden_size, config.rms_norm_eps)
        attr = _init(config.initializer_range)
        self.sp = config.sequence_parallel and config.tensor_parallel_degree > 1
        if config.tensor_parallel_degree > 1:
            from paddle_b200.distributed.fleet.meta_parallel import ColumnParallelLinear
            self.lm_head = ColumnParallelLinear(config.hidden_size, config.vocab_size, weight_attr=attr, has_bias=False,
                                                gather_output=False)
        else:
            self.lm_head = nn.Linear(config.hidden_size, config.vocab_size, weight_attr=attr, bias_attr=False)

    def forward(self, h):
        h = self.norm(h)
        if self.sp:
            from paddle_b200.distributed.fleet.utils.sequence_parallel_utils import GatherOp
            h = GatherOp.apply(h).transpose(0, 1).contiguous()
        return self.lm_head(h)


def llama_pipeline_descs(config: LlamaConfig):
    from paddle_b200.distributed.fleet.meta_parallel import LayerDesc
    descs = [LayerDesc(_EmbeddingPipe, config)]
    descs += [LayerDesc(LlamaDecoderLayer, config, i) for i in range(config.num_hidden_layers)]
    descs.append(LayerDesc(_HeadPipe, config))
    return descs


def LlamaForCausalLMPipe(config: LlamaConfig, **pipe_kwargs):
    from paddle_b200.distributed.fleet.meta_parallel import PipelineLayer
    crit = LlamaPretrainingCriterion(config)
    return PipelineLayer(llama_pipeline_descs(config), loss_fn=crit, seg_method=pipe_kwargs.pop("seg_method", "layer:LlamaDecoderLayer"),
                         **pipe_kwargs)
