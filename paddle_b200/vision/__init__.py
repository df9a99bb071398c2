"""paddle.vision (reference: python/paddle/vision/__init__.py)."""
from . import datasets, models, ops, transforms  # noqa: F401
from .image import get_image_backend, image_load, set_image_backend  # noqa: F401

__all__ = ["set_image_backend", "get_image_backend", "image_load"]
