"""paddle.incubate.nn (reference: python/paddle/incubate/nn/)."""
from . import functional  # noqa: F401
from .layers import FusedLinear, FusedDropoutAdd  # noqa: F401
