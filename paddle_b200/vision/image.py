"""paddle.vision image back end selection (reference: python/paddle/vision/image.py)."""
from .datasets import get_image_backend, image_load, set_image_backend  # noqa: F401

__all__ = ["set_image_backend", "get_image_backend", "image_load"]
