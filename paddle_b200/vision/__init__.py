"""paddle.vision (reference: python/paddle/vision/)."""
from . import datasets, models, transforms  # noqa: F401
