"""paddle.quantization.quanters (reference: python/paddle/quantization/quanters/__init__.py)."""
from .. import FakeQuanterWithAbsMaxObserver, FakeQuanterWithAbsMaxObserverLayer  # noqa: F401

__all__ = ["FakeQuanterWithAbsMaxObserver"]
