"""paddle.base.dygraph shims."""
import contextlib

import torch

no_grad = torch.no_grad


@contextlib.contextmanager
def guard(place=None):
    yield


def to_variable(value, name=None, zero_copy=None, dtype=None):
    from ..tensor.creation import to_tensor
    return to_tensor(value, dtype=dtype)
