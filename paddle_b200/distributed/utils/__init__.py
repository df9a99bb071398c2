from . import moe_utils, nccl_utils, log_utils  # noqa: F401
