"""paddle.tensor namespace (reference: python/paddle/tensor/__init__.py)."""
from .creation import *  # noqa: F401,F403
from .math import *  # noqa: F401,F403
from .manipulation import *  # noqa: F401,F403
from .extra import *  # noqa: F401,F403
from .compat import *  # noqa: F401,F403
from . import compat, creation, extra, math, manipulation  # noqa: F401
