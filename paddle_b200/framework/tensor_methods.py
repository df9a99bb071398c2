"""Paddle-style calls of tensor METHODS.

``paddle.Tensor`` is ``torch.Tensor``, so ``x.max(dim=1)`` keeps meaning what torch says.  Paddle code, however, writes
``x.max(axis=1)`` (values only), ``x.split(3, axis=1)`` (three sections, not sections of three), ``x.gather(index, axis=0)``
(row selection), ``x.sort(axis=0, descending=True)`` (values only), ``x.flatten(start_axis=1)`` ... - and torch either
rejects those keywords or, worse, accepts ``axis`` as an alias of ``dim`` and answers with torch's semantics.

Rule installed here: a method call that uses a Paddle-only keyword (``axis``, ``num_or_sections``, ``start_axis``,
``perm``, ``repeat_times``, ``unbiased`` ...) is routed to the ``paddle.*`` function of the same name, which implements the
reference semantics (python/paddle/tensor/{math,manipulation,search,stat}.py, attached to Tensor by
python/paddle/tensor/__init__.py ``tensor_method_func``); every other call reaches torch's own method untouched, so
torch's internals and torch-style user code see no difference."""
from __future__ import annotations

import torch

_PADDLE_KW = {"axis", "num_or_sections", "start_axis", "stop_axis", "perm", "repeat_times", "unbiased", "shifts_axis",
              "largest_axis", "name", "transpose_x", "transpose_y", "overwrite", "descending_axis"}

# method -> keywords whose presence marks a Paddle-style call
_ROUTES = {
    "max": {"axis"}, "min": {"axis"}, "median": {"axis"}, "sum": {"axis"}, "mean": {"axis"}, "prod": {"axis"},
    "std": {"axis", "unbiased"}, "var": {"axis", "unbiased"}, "all": {"axis"}, "any": {"axis"},
    "argmax": {"axis"}, "argmin": {"axis"}, "cumsum": {"axis"}, "logsumexp": {"axis"}, "amax": {"axis"}, "amin": {"axis"},
    "norm": {"axis"}, "topk": {"axis"}, "sort": {"axis"}, "argsort": {"axis"}, "kthvalue": {"axis"}, "mode": {"axis"},
    "split": {"axis", "num_or_sections"}, "chunk": {"axis"}, "unbind": {"axis"}, "gather": {"axis"},
    "index_select": {"axis"}, "squeeze": {"axis"}, "unsqueeze": {"axis"}, "flatten": {"start_axis", "stop_axis"},
    "roll": {"axis"}, "flip": {"axis"}, "tile": {"repeat_times"}, "expand": {"shape"}, "reshape": {"shape"},
    "matmul": {"transpose_x", "transpose_y"}, "scatter": {"overwrite"}, "unique": {"axis"}, "repeat_interleave": {"axis"},
    "take_along_axis": {"axis"}, "put_along_axis": {"axis"}, "diff": {"axis"}, "count_nonzero": {"axis"}, "nanmean": {"axis"},
    "nansum": {"axis"}, "quantile": {"axis"}, "moveaxis": set(), "concat": {"axis"}, "stack": {"axis"},
}


def _wrap(name, orig, paddle_fn, marks):
    def method(self, *args, **kwargs):
        if kwargs and not marks.isdisjoint(kwargs):
            return paddle_fn(self, *args, **kwargs)
        if orig is None:
            return paddle_fn(self, *args, **kwargs)
        return orig(self, *args, **kwargs)
    method.__name__ = name
    method.__doc__ = getattr(orig, "__doc__", None)
    method._pdb200_routed = True
    return method


def install(namespace):
    """``namespace``: the ``paddle`` module dict (name -> function with the reference semantics)."""
    for name, marks in _ROUTES.items():
        fn = namespace.get(name)
        if fn is None or not callable(fn):
            continue
        orig = getattr(torch.Tensor, name, None)
        if getattr(orig, "_pdb200_routed", False):
            continue
        setattr(torch.Tensor, name, _wrap(name, orig, fn, marks))
    # methods torch does not have at all: attach the paddle function as is
    for name in ("take_along_axis", "put_along_axis", "gather_nd", "scatter_nd_add", "index_sample", "strided_slice", "slice",
                 "crop", "unstack", "rank", "add_n", "stanh", "increment", "is_empty", "equal_all", "floor_mod", "mod",
                 "multiplex", "broadcast_to", "expand_as", "reverse", "astype", "tolist", "top_p_sampling", "lerp", "masked_fill",
                 # the rest of the reference's tensor_method_func list
                 "create_parameter", "create_tensor", "histogramdd", "reduce_as", "multigammaln", "block_diag", "isin",
                 "broadcast_shape", "gammaincc", "gammainc", "greater_than", "less_than", "is_tensor", "reshape_", "scatter_nd",
                 "shard_index", "tensordot", "is_integer", "gammaln", "broadcast_tensors", "cdist", "as_complex", "as_real",
                 "put_along_axis_", "bucketize", "trapezoid", "cumulative_trapezoid", "polar", "vander", "i0e", "i1", "i1e",
                 "atleast_1d", "atleast_2d", "atleast_3d", "combinations"):
        if not hasattr(torch.Tensor, name) and callable(namespace.get(name)):
            setattr(torch.Tensor, name, namespace[name])
    linalg = namespace.get("linalg")
    for name in ("cond", "householder_product", "pca_lowrank", "svd_lowrank", "eigvals", "eigvalsh", "multi_dot", "lu_unpack", "pinv"):
        fn = getattr(linalg, name, None) if linalg is not None else None
        if fn is not None and not hasattr(torch.Tensor, name):
            setattr(torch.Tensor, name, fn)
