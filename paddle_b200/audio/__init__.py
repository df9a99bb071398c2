"""paddle.audio (reference: python/paddle/audio/__init__.py)."""
from . import backends, datasets, features, functional  # noqa: F401
from .backends import info, load, save  # noqa: F401

__all__ = ["functional", "features", "datasets", "backends", "load", "info", "save"]
