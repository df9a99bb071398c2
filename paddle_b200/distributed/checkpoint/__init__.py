from . import metadata, utils  # noqa: F401
from .load_state_dict import load_state_dict  # noqa: F401
from .save_state_dict import save_state_dict  # noqa: F401
