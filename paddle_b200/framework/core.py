from ..base.core import *  # noqa: F401,F403
