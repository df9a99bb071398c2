"""paddle.incubate (reference: python/paddle/incubate/)."""
from . import distributed, nn  # noqa: F401
