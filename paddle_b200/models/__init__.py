"""Model families shipped with the framework: the five BASELINE.json configs are built from these."""
from .gpt import GPTConfig, GPTForCausalLM  # noqa: F401
from .llama import (LlamaConfig, LlamaDecoderLayer, LlamaForCausalLM, LlamaForCausalLMPipe, LlamaModel,  # noqa: F401
                    LlamaPretrainingCriterion)
from .moe import ExpertFFN, MoEConfig, MoEForCausalLM  # noqa: F401
from ..vision.models.resnet import resnet50  # noqa: F401
