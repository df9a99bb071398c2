"""paddle.base compatibility namespace (reference: python/paddle/base/)."""
from . import core, framework  # noqa: F401
from ..framework.param_attr import ParamAttr  # noqa: F401
from ..utils import unique_name  # noqa: F401
from . import dygraph  # noqa: F401
