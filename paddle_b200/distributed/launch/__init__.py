"""python -m paddle.distributed.launch (reference: python/paddle/distributed/launch/)."""
from .main import launch  # noqa: F401
