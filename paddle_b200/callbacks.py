"""paddle.callbacks (reference: python/paddle/callbacks.py): re-export of the hapi callbacks."""
from .hapi.callbacks import *  # noqa: F401,F403
from .hapi.callbacks import Callback, EarlyStopping, LRScheduler, ModelCheckpoint, ProgBarLogger  # noqa: F401
