"""paddle.base.framework (reference: python/paddle/base/framework.py)."""
import torch

from ..framework.flags import get_flags, set_flags  # noqa: F401
from ..framework.place import current_place as _current_expected_place  # noqa: F401
from ..framework.tensor_patch import EagerParamBase, ParamBase, Parameter  # noqa: F401

Variable = torch.Tensor


def in_dygraph_mode():
    return True


def in_dynamic_mode():
    return True


def in_dynamic_or_pir_mode():
    return True


def in_pir_mode():
    return False


def _non_static_mode():
    return True


def _set_expected_place(place):
    from ..framework.place import set_device
    set_device(place)


def _current_expected_place_():
    return _current_expected_place()
