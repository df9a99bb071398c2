from .layers import Layer, LayerList, Sequential, ParameterList, LayerDict, HookRemoveHelper  # noqa: F401
from .common import *  # noqa: F401,F403
from .extra import *  # noqa: F401,F403
from .conv_pool import *  # noqa: F401,F403,E402  (the full conv / pool / pad / resample layer set)
from .transformer import *  # noqa: F401,F403,E402
