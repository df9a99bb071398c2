"""Op layer: each function runs the hand-written sm_100a kernel from libpdb200_kernels.so when its inputs live on a
GPU, and the equivalent plain-PyTorch composition on CPU (which is also the numerical oracle used by the tests).
Kernel ids (K..) refer to SURVEY.md §8."""
from .activation import fused_bias_act, swiglu  # noqa: F401
from .attention import flash_attention, flash_attn_varlen  # noqa: F401
from .linear import linear  # noqa: F401
from .loss import (ce_combine, ce_local_backward, ce_local_stats, softmax_cross_entropy)  # noqa: F401
from .moe import (assign_pos, grouped_linear, limit_by_capacity, number_count, prune_gate_by_capacity, random_routing,  # noqa: F401
                  topk_softmax)
from .multi_tensor import (TensorTable, adamw_step, grads_table, l2norm_sq, momentum_step, squared_l2_norm,  # noqa: F401
                           unscale_and_check)
from .norm import layer_norm, rms_norm  # noqa: F401
from .rope import fused_rope, qkv_split_rope  # noqa: F401
from . import gemm  # noqa: F401
