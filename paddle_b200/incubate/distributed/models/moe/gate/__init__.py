"""MoE gates (reference: python/paddle/incubate/distributed/models/moe/gate/{base,naive,gshard,switch}_gate.py)."""
from __future__ import annotations

import math

import torch

from paddle_b200 import ops
from paddle_b200.nn import functional as F
from paddle_b200.nn.layer.common import Linear
from paddle_b200.nn.layer.layers import Layer

from ..utils import limit_by_capacity


class BaseGate(Layer):
    def __init__(self, num_expert, world_size):
        super().__init__()
        self.world_size = world_size
        self.num_expert = num_expert
        self.tot_expert = world_size * num_expert
        self.loss = None

    def forward(self, x):
        raise NotImplementedError("Please implement the forward function.")

    def set_loss(self, loss):
        self.loss = loss

    def get_loss(self, clear=True):
        loss = self.loss
        if clear:
            self.loss = None
        return loss


class NaiveGate(BaseGate):
    def __init__(self, d_model, num_expert, world_size, topk=2):
        super().__init__(num_expert, world_size)
        self.gate = Linear(d_model, self.tot_expert)
        self.gate.weight.name = "gate_" + self.gate.weight.name
        self.gate.bias.name = "gate_" + self.gate.bias.name
        self.top_k = topk

    def forward(self, inp, return_all_scores=False):
        gate = self.gate(inp)
        val, idx = torch.topk(gate, k=self.top_k, dim=-1, largest=True, sorted=False)
        if return_all_scores:
            return val, idx, gate
        return val, idx


class GShardGate(NaiveGate):
    def __init__(self, d_model, num_expert, world_size, topk=2, capacity=(1.2, 2.4), random_routing=True, group=None):
        assert topk == 2, "topk should be 2 in gshard"
        super().__init__(d_model, num_expert, world_size)
        self.capacity = capacity
        self.random_routing = random_routing
        self.group = group

    def forward(self, x):
        topk_val, topk_idx, gate_score = super().forward(x, return_all_scores=True)
        s = gate_score.shape[0]
        top1_idx = topk_idx.flatten()
        c_e = torch.zeros(self.tot_expert, device=x.device).index_add_(
            0, top1_idx, torch.ones_like(top1_idx, dtype=torch.float32)) / s
        m_e = torch.mean(F.softmax(gate_score.float(), axis=1), dim=0)
        self.set_loss(torch.mean(c_e * m_e) * (self.num_expert ** 2))
        cap_rate = self.capacity[0 if self.training else 1]
        capacity = math.ceil(cap_rate * x.shape[0])
        _, _, topk_idx = limit_by_capacity(topk_idx, self.num_expert, self.world_size, capacity, group=self.group)
        if self.random_routing:
            prob = torch.rand(gate_score.shape[0], dtype=torch.float32, device=x.device)
            topk_idx = ops.random_routing(topk_idx, topk_val, prob)
        return topk_val, topk_idx


class SwitchGate(NaiveGate):
    def __init__(self, d_model, num_expert, world_size, topk=1, switch_eps=0.1, capacity=(1.2, 2.4), group=None):
        assert topk == 1, "topk should be 1 in switch"
        super().__init__(d_model, num_expert, world_size, topk=1)
        self.switch_eps = switch_eps
        self.capacity = capacity
        self.group = group

    def forward(self, inp):
        score = self.gate(inp)
        if self.training:
            noise = torch.rand_like(score.float()) * 2 * self.switch_eps + 1.0 - self.switch_eps
            score = score * noise.to(score.dtype)
        score = F.softmax(score.float(), axis=-1)
        top1_score, top1_idx = torch.topk(score, k=1, dim=-1, largest=True)
        cap_rate = self.capacity[0 if self.training else 1]
        capacity = math.ceil(cap_rate * inp.shape[0])
        _, _, top1_idx = limit_by_capacity(top1_idx, self.num_expert, self.world_size, capacity, group=self.group)
        valid_idx = top1_idx[top1_idx > -1]
        fraction_expert = torch.zeros(self.tot_expert, device=inp.device).index_add_(
            0, valid_idx, torch.ones_like(valid_idx, dtype=torch.float32)) / max(valid_idx.numel(), 1)
        prob_expert = score.sum(0) / max(valid_idx.numel(), 1)
        self.set_loss((fraction_expert * prob_expert).sum() * self.tot_expert)
        return top1_score.to(inp.dtype), top1_idx
