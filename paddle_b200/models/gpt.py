"""GPT (decoder-only, LayerNorm + GELU) - BASELINE config #1 uses the tiny variant (2 layers, hidden 256) under
DataParallel on CPU/gloo.  Reference model shape: test/collective/fleet hybrid-parallel GPT test models."""
from __future__ import annotations

from dataclasses import dataclass

import torch

from paddle_b200 import nn
from paddle_b200.nn import functional as F


@dataclass
class GPTConfig:
    vocab_size: int = 1024
    hidden_size: int = 256
    num_hidden_layers: int = 2
    num_attention_heads: int = 4
    intermediate_size: int | None = None
    max_position_embeddings: int = 512
    dropout: float = 0.0
    layer_norm_eps: float = 1e-5

    def __post_init__(self):
        if self.intermediate_size is None:
            self.intermediate_size = 4 * self.hidden_size

    @staticmethod
    def tiny(**kw):
        return GPTConfig(**kw)


class GPTBlock(nn.Layer):
    def __init__(self, c: GPTConfig):
        super().__init__()
        self.c = c
        self.ln1 = nn.LayerNorm(c.hidden_size, c.layer_norm_eps)
        self.qkv = nn.Linear(c.hidden_size, 3 * c.hidden_size)
        self.proj = nn.Linear(c.hidden_size, c.hidden_size)
        self.ln2 = nn.LayerNorm(c.hidden_size, c.layer_norm_eps)
        self.fc1 = nn.Linear(c.hidden_size, c.intermediate_size)
        self.fc2 = nn.Linear(c.intermediate_size, c.hidden_size)
        self.drop = nn.Dropout(c.dropout)

    def forward(self, x):
        b, s, h = x.shape
        nh = self.c.num_attention_heads
        qkv = self.qkv(self.ln1(x)).reshape(b, s, 3, nh, h // nh)
        out, _ = F.flash_attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=True, training=self.training)
        x = x + self.drop(self.proj(out.reshape(b, s, h)))
        return x + self.drop(self.fc2(F.gelu(self.fc1(self.ln2(x)))))


class GPTForCausalLM(nn.Layer):
    def __init__(self, c: GPTConfig):
        super().__init__()
        self.c = c
        self.wte = nn.Embedding(c.vocab_size, c.hidden_size)
        self.wpe = nn.Embedding(c.max_position_embeddings, c.hidden_size)
        self.blocks = nn.LayerList([GPTBlock(c) for _ in range(c.num_hidden_layers)])
        self.ln_f = nn.LayerNorm(c.hidden_size, c.layer_norm_eps)
        self.lm_head = nn.Linear(c.hidden_size, c.vocab_size, bias_attr=False)

    def forward(self, input_ids, labels=None):
        s = input_ids.shape[1]
        pos = torch.arange(s, device=input_ids.device)
        x = self.wte(input_ids) + self.wpe(pos)[None]
        for blk in self.blocks:
            x = blk(x)
        logits = self.lm_head(self.ln_f(x))
        if labels is None:
            return logits
        return F.cross_entropy(logits.reshape(-1, logits.shape[-1]), labels.reshape(-1))
