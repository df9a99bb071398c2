"""paddle.quantization.observers (reference: python/paddle/quantization/observers/__init__.py)."""
from .. import AbsmaxObserver, AbsmaxObserverLayer, GroupWiseWeightObserver, GroupWiseWeightObserverLayer  # noqa: F401

__all__ = ["AbsmaxObserver", "GroupWiseWeightObserver"]
