"""paddle.optimizer (reference: python/paddle/optimizer/__init__.py)."""
from . import lr  # noqa: F401
from .optimizer import (ASGD, LBFGS, SGD, Adadelta, Adagrad, Adam, Adamax, AdamW, Lamb, Momentum, NAdam, Optimizer, RAdam,  # noqa: F401
                        RMSProp, Rprop)
