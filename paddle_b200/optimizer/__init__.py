"""paddle.optimizer (reference: python/paddle/optimizer/__init__.py)."""
from . import lr  # noqa: F401
from .optimizer import SGD, Adagrad, Adam, AdamW, Lamb, Momentum, Optimizer, RMSProp  # noqa: F401
