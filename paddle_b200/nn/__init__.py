"""paddle.nn (reference: python/paddle/nn/__init__.py)."""
from . import functional, initializer  # noqa: F401
from .clip import ClipGradByGlobalNorm, ClipGradByNorm, ClipGradByValue  # noqa: F401
from .layer import *  # noqa: F401,F403
from .layer.layers import Layer, LayerDict, LayerList, ParameterList, Sequential  # noqa: F401
from . import utils  # noqa: F401
from . import quant  # noqa: F401
