"""paddle.incubate.nn (reference: python/paddle/incubate/nn/)."""
from . import functional  # noqa: F401
from .layers import (FusedBiasDropoutResidualLayerNorm, FusedDropoutAdd, FusedEcMoe, FusedFeedForward, FusedLinear,  # noqa: F401
                     FusedMultiHeadAttention, FusedMultiTransformer, FusedTransformerEncoderLayer)
