"""All local experts of an MoE layer as ONE module with stacked weights, computed by device-scheduled grouped GEMMs
(K9/K10, csrc/kernels/gemm_grouped.cu).

The reference runs ``for idx, expert in enumerate(self.experts)`` over slices whose sizes it first reads back from the
device (python/paddle/incubate/distributed/models/moe/moe_layer.py:441-470).  Here the experts are a SwiGLU FFN with
weights ``gate_up_weight [E, d, 2*ff]`` and ``down_weight [E, ff, d]``; ``forward(x, counts)`` takes the expert-major
slab and the per-expert row counts as a DEVICE tensor, and launches three kernels (grouped gate_up GEMM, SwiGLU,
grouped down GEMM) whatever the number of experts - no host synchronisation in forward or backward.  Rows of the
slab beyond ``sum(counts)`` are ignored (their outputs are unspecified and never read by the combine step)."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.nn.layer.layers import Layer


class GroupedExpertsFFN(Layer):
    def __init__(self, num_experts, d_model, d_ffn, weight_attr=None, dtype=None):
        super().__init__()
        self.num_experts, self.d_model, self.d_ffn = num_experts, d_model, d_ffn
        self.gate_up_weight = self.create_parameter([num_experts, d_model, 2 * d_ffn], attr=weight_attr, dtype=dtype or self._dtype)
        self.down_weight = self.create_parameter([num_experts, d_ffn, d_model], attr=weight_attr, dtype=dtype or self._dtype)

    def __len__(self):
        return self.num_experts

    def forward(self, x, counts):
        h = ops.grouped_linear(x, self.gate_up_weight, counts)
        a = ops.swiglu(h)
        return ops.grouped_linear(a, self.down_weight, counts)

    @classmethod
    def from_experts(cls, experts):
        """Stack a list of ``ExpertFFN``-shaped layers (``gate_up_proj`` / ``down_proj`` Linears without bias)."""
        e0 = experts[0]
        d, ff2 = e0.gate_up_proj.weight.shape
        out = cls(len(experts), d, ff2 // 2, dtype=e0.gate_up_proj.weight.dtype)
        with torch.no_grad():
            for i, e in enumerate(experts):
                out.gate_up_weight[i].copy_(e.gate_up_proj.weight)
                out.down_weight[i].copy_(e.down_proj.weight)
        return out
