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


def _norm_except_dim(v, dim):
    """||v|| over every dim but ``dim``, as a 1-D tensor of ``v.shape[dim]`` entries (a 1-entry tensor for dim=None)."""
    import torch
    if dim is None or dim == -1:
        return v.norm().reshape(1)
    dims = [d for d in range(v.dim()) if d != dim]
    return torch.linalg.vector_norm(v, 2, dim=dims) if dims else v.abs()


def _weight_from(g, v, dim):
    if dim is None or dim == -1:
        return v * (g / v.norm())
    shape = [1] * v.dim()
    shape[dim] = -1
    return v * (g / _norm_except_dim(v, dim)).reshape(shape)


class _WeightNormHook:
    def __init__(self, name, dim):
        self.name, self.dim = name, dim

    def __call__(self, layer, inputs):
        g, v = getattr(layer, self.name + "_g"), getattr(layer, self.name + "_v")
        object.__setattr__(layer, self.name, _weight_from(g, v, self.dim))


def weight_norm(layer, name="weight", dim=0):
    """w = g * v / ||v|| re-parameterisation (reference: python/paddle/nn/utils/weight_norm_hook.py).  ``<name>`` is
    replaced by the parameters ``<name>_g`` - 1-D, one magnitude per slice along ``dim`` (one entry in all for
    ``dim=None``) - and ``<name>_v`` (the direction, shaped like the weight); the weight is recomputed before every
    forward."""
    import torch
    from ...framework.tensor_patch import Parameter
    if hasattr(layer, "_pd_weight_norm_" + name):
        raise RuntimeError(f"Cannot register two weight_norm hooks on the same parameter {name}")
    w = getattr(layer, name)
    nd = w.dim()
    if dim is not None:
        if not -nd <= dim < nd:
            raise ValueError(f"dim must set between [-R, R), R means the dimension of weight, got {dim}")
        dim = dim % nd if dim != -1 else None
    with torch.no_grad():
        g = Parameter(_norm_except_dim(w.detach(), dim).clone(), name=getattr(w, "name", name) + "_g")
        v = Parameter(w.detach().clone(), name=getattr(w, "name", name) + "_v")
    del layer._parameters[name]
    layer.register_parameter(name + "_g", g)
    layer.register_parameter(name + "_v", v)
    object.__setattr__(layer, name, _weight_from(g, v, dim).detach())
    hook = _WeightNormHook(name, dim)
    handle = layer.register_forward_pre_hook(hook)

    def _drop_graph(lyr, inputs, output):      # between forwards the attribute is a plain tensor (readable, deep-copyable)
        object.__setattr__(lyr, name, getattr(lyr, name).detach())
    post = layer.register_forward_post_hook(_drop_graph)
    object.__setattr__(layer, "_pd_weight_norm_" + name, (hook, handle, post))
    return layer


def remove_weight_norm(layer, name="weight"):
    import torch
    from ...framework.tensor_patch import Parameter
    entry = getattr(layer, "_pd_weight_norm_" + name, None)
    if entry is None:
        raise ValueError(f"weight_norm of '{name}' not found in {layer}")
    hook, handle, post = entry
    post.remove()
    g, v = getattr(layer, name + "_g"), getattr(layer, name + "_v")
    with torch.no_grad():
        w = Parameter(_weight_from(g, v, hook.dim).detach().clone(), name=v.name[:-2] if v.name.endswith("_v") else name)
    handle.remove()
    del layer._parameters[name + "_g"], layer._parameters[name + "_v"]
    if name in layer.__dict__:
        object.__delattr__(layer, name)
    layer.register_parameter(name, w)
    object.__delattr__(layer, "_pd_weight_norm_" + name)
    return layer


def spectral_norm(layer, name="weight", n_power_iterations=1, eps=1e-12, dim=None):
    import torch
    return torch.nn.utils.spectral_norm(layer, name=name, n_power_iterations=n_power_iterations, eps=eps, dim=dim)
