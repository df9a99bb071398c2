from .gate import BaseGate, GShardGate, NaiveGate, SwitchGate  # noqa: F401
from .grad_clip import ClipGradForMOEByGlobalNorm  # noqa: F401
from .grouped_experts import GroupedExpertsFFN  # noqa: F401
from .moe_layer import MoELayer  # noqa: F401

ClipGradByGlobalNorm = ClipGradForMOEByGlobalNorm
