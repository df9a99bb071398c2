"""paddle.framework (reference: python/paddle/framework/__init__.py)."""
from . import dtype, flags, place, random, tensor_patch  # noqa: F401
from .dtype import get_default_dtype, set_default_dtype  # noqa: F401
from .flags import get_flags, set_flags  # noqa: F401
from .io import async_save, load, save  # noqa: F401
from .param_attr import ParamAttr  # noqa: F401
from .place import CPUPlace, CUDAPinnedPlace, CUDAPlace  # noqa: F401
from .random import get_cuda_rng_state, get_rng_state, seed, set_cuda_rng_state, set_rng_state  # noqa: F401
from .tensor_patch import EagerParamBase, Parameter  # noqa: F401
from . import core  # noqa: F401


def in_dynamic_mode():
    return True


def in_dygraph_mode():
    return True


def in_pir_mode():
    return False


def in_dynamic_or_pir_mode():
    return True


def _current_expected_place():
    return place.current_place()
