"""paddle.nn.utils (reference: python/paddle/nn/utils/*.py)."""
import torch

from ..clip import clip_grad_norm_, clip_grad_value_  # noqa: F401


def parameters_to_vector(parameters, name=None):
    return torch.cat([p.reshape(-1) for p in parameters])


def vector_to_parameters(vec, parameters, name=None):
    off = 0
    with torch.no_grad():
        for p in parameters:
            n = p.numel()
            p.copy_(vec[off:off + n].reshape(p.shape))
            off += n


def weight_norm(layer, name="weight", dim=0):
    """w = g * v / ||v|| re-parameterisation (reference: nn/utils/weight_norm_hook.py); parameters ``<name>_g/_v``."""
    import torch
    return torch.nn.utils.weight_norm(layer, name=name, dim=dim if dim is not None else None)


def remove_weight_norm(layer, name="weight"):
    import torch
    return torch.nn.utils.remove_weight_norm(layer, name=name)


def spectral_norm(layer, name="weight", n_power_iterations=1, eps=1e-12, dim=None):
    import torch
    return torch.nn.utils.spectral_norm(layer, name=name, n_power_iterations=n_power_iterations, eps=eps, dim=dim)
