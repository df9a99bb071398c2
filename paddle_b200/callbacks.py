"""paddle.callbacks (reference: python/paddle/callbacks.py): re-export of the hapi callbacks."""
from .hapi.callbacks import (Callback, EarlyStopping, LRScheduler, ModelCheckpoint, ProgBarLogger, ReduceLROnPlateau,  # noqa: F401
                             VisualDL, WandbCallback)

__all__ = ["Callback", "ProgBarLogger", "ModelCheckpoint", "VisualDL", "LRScheduler", "EarlyStopping", "ReduceLROnPlateau",
           "WandbCallback"]
