"""Distributed tensors that propagate their sharding through ordinary ops (P35 / N11 / N12).

Reference behaviour: ``paddle.distributed.shard_tensor`` returns a DistTensor, every op called on DistTensors runs the
op's SPMD rule (paddle/phi/infermeta/spmd_rules), reshards the inputs the rule wants resharded
(paddle/phi/core/distributed/auto_parallel/reshard/*), computes on the local shards and returns DistTensors; the
backward does the same with the rules' backward halves (python/paddle/distributed/auto_parallel/api.py:180-520,
paddle/phi/api/lib/api_gen_utils.cc, generated dist branches of the eager API).

Here a DistTensor IS its local shard (a ``torch.Tensor`` subclass carrying ``process_mesh`` / ``placements`` /
``_global_shape``), and ``__torch_function__`` is the dist branch: known functions map to an SPMD rule
(``spmd_rules``) plus a local computation, unknown tensor-producing functions fall back to "replicate everything, run,
result replicated" (the reference's ``replicated`` rule), metadata accessors act on the local shard.

Gradients need no second rule set.  Autograd runs on the local graph; three facts make its result the right shard of
the global gradient:
  * every reshard step is an autograd function whose backward is the dual transfer (all-gather <-> slice,
    reduce-scatter <-> all-gather, all-to-all <-> all-to-all, all-reduce <-> identity, / n for averaged partials);
  * the gradient of a tensor that is Partial over a mesh axis is the global gradient on every rank, the gradient of a
    Shard is the shard of the global gradient, the gradient of a Replicate is the global gradient on every rank;
  * the one case the local graph gets wrong: an input that is Replicate over a mesh axis along which the op's output is
    Shard or Partial (ranks along that axis computed different things from the same value) receives only this rank's
    contribution.  Such inputs are passed through an identity whose backward all-reduces over that axis - which is
    exactly data-parallel gradient synchronisation for replicated parameters and Megatron's ``c_identity`` for
    replicated activations in front of a column-parallel matmul.
"""
from __future__ import annotations

import math
import warnings

import torch
import torch.nn.functional as TF

from .. import collective as C
from .. import communication as comm
from ...framework.tensor_patch import Parameter
from . import spmd_rules as R


def _P():
    from . import Partial, Replicate, Shard
    return Partial, Replicate, Shard


# ----------------------------------------------------------------------------------------------------- tensor classes
class DistTensor(torch.Tensor):
    """The local shard of a distributed tensor.  ``.shape`` is the LOCAL shape; ``._global_shape`` the global one."""

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        return _torch_function(func, args, kwargs or {})


class DistParameter(Parameter):
    """A parameter whose local value is a shard (or a replica) of a distributed tensor."""

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        return _torch_function(func, args, kwargs or {})


def is_dist(t):
    return isinstance(t, torch.Tensor) and hasattr(t, "placements") and hasattr(t, "process_mesh")


def _attach(t, mesh, placements, global_shape):
    _, _, Shard = _P()
    t.process_mesh = mesh
    t.placements = list(placements)
    coord = mesh.coord_of(C.get_rank()) or tuple([0] * mesh.ndim)
    off = [0] * len(global_shape)
    for ax, pl in enumerate(placements):
        if isinstance(pl, Shard) and pl.dim < len(global_shape) and pl.dim < t.dim():
            off[pl.dim] += coord[ax] * t.shape[pl.dim]
    t._dist_meta = (tuple(global_shape), tuple(off))
    t._global_shape = tuple(global_shape)
    return t


def wrap(local, mesh, placements, global_shape):
    """Label a local result.  Non-leaf results stay attached to the autograd graph (``as_subclass`` aliases them)."""
    if isinstance(local, (DistTensor, DistParameter)):
        out = local
    else:
        with torch._C.DisableTorchFunctionSubclass():
            out = local.as_subclass(DistTensor)
    return _attach(out, mesh, placements, global_shape)


def make_leaf(local, mesh, placements, global_shape, requires_grad=False):
    out = torch.Tensor._make_subclass(DistTensor, local.detach(), requires_grad)
    return _attach(out, mesh, placements, global_shape)


def make_parameter(local, mesh, placements, global_shape, like=None):
    p = DistParameter(local.detach(), like.requires_grad if like is not None else True,
                      name=getattr(like, "name", None))
    if like is not None:
        for k, v in like.__dict__.items():
            if k not in ("_pd_name", "process_mesh", "placements", "_dist_meta", "_global_shape"):
                p.__dict__.setdefault(k, v)
    return _attach(p, mesh, placements, global_shape)


def local_of(t):
    """The same storage and autograd node as a plain tensor (no dispatch through this module)."""
    if isinstance(t, (DistTensor, DistParameter)):
        with torch._C.DisableTorchFunctionSubclass():
            return t.as_subclass(torch.Tensor)
    return t


# --------------------------------------------------------------------------------------------- differentiable reshard
def _axis_group(mesh, ax):
    return mesh.axis_group(ax) if mesh.shape[ax] > 1 else None


def _gather(t, group, dim, n):
    parts = [torch.empty_like(t) for _ in range(n)]
    comm.all_gather(parts, t.contiguous(), group)
    return torch.cat(parts, dim=dim)


def _transfer(t, mesh, ax, a, b):
    """One mesh axis, placement a -> b, on a plain local tensor, no autograd."""
    Partial, Replicate, Shard = _P()
    n = mesh.shape[ax]
    if a == b or n == 1:
        if isinstance(a, Replicate) and isinstance(b, Shard):
            return t                                   # axis of size 1: the whole dim is the shard
        return t
    g = _axis_group(mesh, ax)
    coord = (mesh.coord_of(C.get_rank()) or tuple([0] * mesh.ndim))[ax]
    if isinstance(a, Shard) and isinstance(b, Replicate):
        return _gather(t, g, a.dim, n)
    if isinstance(a, Replicate) and isinstance(b, Shard):
        return t.chunk(n, dim=b.dim)[coord].contiguous()
    if isinstance(a, Partial) and isinstance(b, Replicate):
        out = t.clone()
        comm.all_reduce(out, a.reduce_type, g)
        return out
    if isinstance(a, Partial) and isinstance(b, Shard):
        moved = t.movedim(b.dim, 0).contiguous()
        out = torch.empty((moved.shape[0] // n,) + tuple(moved.shape[1:]), dtype=t.dtype, device=t.device)
        comm.reduce_scatter(out, moved, a.reduce_type, g)
        return out.movedim(0, b.dim).contiguous()
    if isinstance(a, Shard) and isinstance(b, Shard):
        pieces = [c.contiguous() for c in t.chunk(n, dim=b.dim)]
        recv = [torch.empty_like(pieces[0]) for _ in range(n)]
        comm.alltoall(recv, pieces, g)
        return torch.cat(recv, dim=a.dim)
    if isinstance(a, Replicate) and isinstance(b, Partial):
        return t if coord == 0 else torch.zeros_like(t)
    if isinstance(a, Shard) and isinstance(b, Partial):
        full = _gather(t, g, a.dim, n)
        return full if coord == 0 else torch.zeros_like(full)
    raise NotImplementedError(f"reshard {a} -> {b}")


def _grad_placement(p):
    """Placement of the gradient of a tensor placed as p (module docstring)."""
    Partial, Replicate, Shard = _P()
    return Replicate() if isinstance(p, Partial) else p


class _Transfer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t, mesh, ax, a, b):
        ctx.cfg = (mesh, ax, a, b)
        return _transfer(t, mesh, ax, a, b)

    @staticmethod
    def backward(ctx, g):
        Partial, Replicate, Shard = _P()
        mesh, ax, a, b = ctx.cfg
        n = mesh.shape[ax]
        ga, gb = _grad_placement(a), _grad_placement(b)
        out = _transfer(g.contiguous(), mesh, ax, gb, ga)
        if isinstance(a, Partial) and n > 1:
            if a.reduce_type == C.ReduceOp.AVG:
                out = out / n                              # y = mean_i x_i
            elif a.reduce_type != C.ReduceOp.SUM:
                raise NotImplementedError("gradient through a max / min / prod partial")
        if isinstance(a, Replicate) and isinstance(b, Partial) and n > 1:
            # forward kept rank 0's copy only; every rank's replica must still receive the global gradient
            out = g
        return out, None, None, None, None


def _step(t, mesh, ax, a, b):
    if a == b or mesh.shape[ax] == 1:
        return t
    if t.requires_grad and torch.is_grad_enabled():
        return _Transfer.apply(t, mesh, ax, a, b)
    return _transfer(t, mesh, ax, a, b)


def reshard_local(t, mesh, src, dst):
    """Local tensor placed as ``src`` -> local tensor placed as ``dst`` (same mesh), differentiable.

    Mesh axes whose change touches no tensor dim that another axis shards move directly (one collective each: slice,
    all-gather, all-reduce, reduce-scatter or all-to-all).  When two axes meet on a tensor dim (before or after), the
    changing axes go through Replicate: nested shards of one dim are gathered innermost axis first and cut outermost
    first, which is the order ``shard_tensor`` cuts in."""
    Partial, Replicate, Shard = _P()
    changing = [ax for ax, (a, b) in enumerate(zip(src, dst)) if a != b and mesh.shape[ax] > 1]
    if not changing:
        return t

    def dims(pl, skip):
        return {p.dim for ax, p in enumerate(pl) if isinstance(p, Shard) and ax != skip and mesh.shape[ax] > 1}
    conflict = False
    for ax in changing:
        touched = {p.dim for p in (src[ax], dst[ax]) if isinstance(p, Shard)}
        if touched & (dims(src, ax) | dims(dst, ax)):
            conflict = True
    if not conflict:
        for ax in changing:
            t = _step(t, mesh, ax, src[ax], dst[ax])
        return t
    cur = list(src)
    for ax in sorted(changing, reverse=True):              # innermost shard first
        if not isinstance(cur[ax], Replicate):
            t = _step(t, mesh, ax, cur[ax], Replicate())
            cur[ax] = Replicate()
    for ax in sorted(changing):                            # outermost cut first
        t = _step(t, mesh, ax, cur[ax], dst[ax])
        cur[ax] = dst[ax]
    return t


class _GradAllReduce(torch.autograd.Function):
    """Identity whose backward sums the gradient over mesh axes (see the module docstring, third fact)."""

    @staticmethod
    def forward(ctx, t, mesh, axes):
        ctx.cfg = (mesh, tuple(axes))
        return t.view_as(t)

    @staticmethod
    def backward(ctx, g):
        mesh, axes = ctx.cfg
        g = g.contiguous().clone()
        for ax in axes:
            comm.all_reduce(g, C.ReduceOp.SUM, _axis_group(mesh, ax))
        return g, None, None


# ------------------------------------------------------------------------------------------------------- rule runner
def _mesh_of(tensors):
    for t in tensors:
        if is_dist(t):
            return t.process_mesh
    return None


def _spec_of(t, mesh):
    if is_dist(t):
        return R.DistTensorSpec.from_placements(t._global_shape, t.placements)
    return R.DistTensorSpec(tuple(t.shape), [-1] * t.dim())          # a plain tensor is a replica


def run_rule(rule, local_fn, tensors, attrs=None, out_reduce=None):
    """Run an op on distributed tensors.

    rule       name in ``spmd_rules`` (or a callable with the same signature)
    local_fn   called with the resharded LOCAL tensors (plain ``torch.Tensor``), returns a tensor or a tuple of tensors,
               one per output spec of the rule (extra outputs of the rule are ignored)
    tensors    the op's tensor inputs, DistTensors or plain tensors (treated as replicated); ``None`` entries allowed
    attrs      keyword arguments of the rule
    out_reduce reduce type of partial outputs (ReduceOp.SUM unless given, e.g. ReduceOp.AVG for mean, ReduceOp.MAX)
    """
    Partial, Replicate, Shard = _P()
    present = [t for t in tensors if t is not None]
    mesh = _mesh_of(present)
    specs = [_spec_of(t, mesh) for t in present]
    fn = rule if callable(rule) else R.get_spmd_rule(rule)
    new_specs, out_specs = fn(*specs, **(attrs or {}))
    nd = mesh.ndim
    if not in_mesh(mesh):
        # this rank holds placeholders only: nothing to compute, the outputs are placeholders tied into the graph
        ref = next(t for t in present if isinstance(t, torch.Tensor))
        locs = [local_of(t) for t in present]
        need_graph = torch.is_grad_enabled() and any(l.requires_grad for l in locs)
        raw = _Skip.apply(len(out_specs), ref.dtype, ref.device, *locs) if need_graph else tuple(
            torch.empty(0, dtype=ref.dtype, device=ref.device) for _ in out_specs)
        outs = []
        for o, sp in zip(raw, out_specs):
            pl = sp.placements(nd)
            if out_reduce is not None:
                pl = [Partial(out_reduce) if isinstance(p, Partial) else p for p in pl]
            w = wrap(o, mesh, pl, tuple(sp.shape))
            w._placeholder = True
            outs.append(w)
        return outs[0] if len(outs) == 1 else tuple(outs)
    out_axes_busy = set()
    for sp in out_specs:
        out_axes_busy |= {ax for ax in sp.dims_mapping if ax >= 0} | set(sp.partial_dims)
    locals_ = []
    for t, want in zip(present, new_specs):
        dst = want.placements(nd)
        if is_dist(t):
            src, loc = list(t.placements), local_of(t)
        else:
            src, loc = [Replicate() for _ in range(nd)], t
        loc = reshard_local(loc, mesh, src, dst)
        sync = [ax for ax in range(nd) if mesh.shape[ax] > 1 and isinstance(dst[ax], Replicate) and ax in out_axes_busy]
        if sync and loc.requires_grad and torch.is_grad_enabled() and loc.is_floating_point():
            loc = _GradAllReduce.apply(loc, mesh, sync)
        locals_.append(loc)
    it = iter(locals_)
    full = [next(it) if t is not None else None for t in tensors]
    with torch._C.DisableTorchFunctionSubclass():
        res = local_fn(*full)
    single = not isinstance(res, (tuple, list))
    outs = [res] if single else list(res)
    labelled = []
    for o, sp in zip(outs, out_specs):
        if not isinstance(o, torch.Tensor):
            labelled.append(o)
            continue
        pl = sp.placements(nd)
        if out_reduce is not None:
            pl = [Partial(out_reduce) if isinstance(p, Partial) else p for p in pl]
        gshape = list(o.shape)
        for ax, p in enumerate(pl):
            if isinstance(p, Shard):
                gshape[p.dim] *= mesh.shape[ax]
        labelled.append(wrap(o, mesh, pl, tuple(gshape)))
    labelled += outs[len(out_specs):]
    return labelled[0] if single else tuple(labelled)


def replicate_all(args, kwargs):
    """Every DistTensor in (args, kwargs) -> its full value as a plain tensor (differentiable)."""
    _, Replicate, _ = _P()

    def conv(x):
        if is_dist(x):
            mesh = x.process_mesh
            return reshard_local(local_of(x), mesh, list(x.placements), [Replicate() for _ in range(mesh.ndim)])
        if isinstance(x, (list, tuple)):
            return type(x)(conv(v) for v in x)
        if isinstance(x, dict):
            return {k: conv(v) for k, v in x.items()}
        return x
    return conv(tuple(args)), conv(dict(kwargs))


# ---------------------------------------------------------------------------------------- the __torch_function__ table
_META = {
    "shape", "size", "dim", "ndim", "ndimension", "numel", "nelement", "dtype", "device", "requires_grad", "is_leaf", "grad",
    "grad_fn", "data_ptr", "stride", "is_contiguous", "element_size", "nbytes", "itemsize", "is_cuda", "is_cpu",
    "is_floating_point", "is_complex", "is_signed", "__repr__", "__str__", "__format__", "__len__", "__hash__", "backward",
    "register_hook", "register_post_accumulate_grad_hook", "retain_grad", "retains_grad", "requires_grad_", "item", "tolist",
    "numpy", "storage_offset", "untyped_storage", "_version", "layout", "names", "is_sparse", "is_quantized", "is_meta",
    "output_nr", "__reduce_ex__", "__deepcopy__", "data", "_base", "name", "is_shared", "get_device", "is_pinned", "_is_view",
    "is_inference", "is_neg", "is_conj", "dim_order", "is_nested", "is_mkldnn", "__bool__", "__int__", "__float__",
    "__index__", "__contains__", "__iter__", "is_same_size", "_backward_hooks", "grad_fn", "__setstate__", "__dlpack__",
    "storage", "_typed_storage", "is_set_to", "__array__", "cpu", "cuda", "pin_memory", "record_stream", "share_memory_",
    "zero_", "fill_", "normal_", "uniform_", "copy_", "set_", "add_", "sub_", "mul_", "div_", "addcmul_", "addcdiv_", "lerp_",
    "sqrt_", "clamp_", "detach_", "resize_", "random_", "bernoulli_", "exponential_", "trunc_normal_",
}
_TABLE = {}


def _name_of(func):
    n = getattr(func, "__name__", None)
    if n in ("__get__", "__set__", "__delete__"):
        n = getattr(getattr(func, "__self__", None), "__name__", n)
    return n or str(func)


def _register(*funcs):
    def deco(h):
        for f in funcs:
            if f is not None:
                _TABLE[f] = h
        return h
    return deco


def _same_placement_result(func, args, kwargs):
    """Metadata accessors and in-place maths on tensors that share a placement: act on the local shards."""
    with torch._C.DisableTorchFunctionSubclass():
        return func(*args, **kwargs)


_warned = set()


def _torch_function(func, args, kwargs):
    h = _TABLE.get(func)
    if h is not None:
        return h(*args, **kwargs)
    name = _name_of(func)
    if name in _META or name.startswith("_foreach_") or (name.endswith("_") and not name.startswith("__")):
        return _same_placement_result(func, args, kwargs)
    dist = [a for a in _flatten((args, kwargs)) if is_dist(a)]
    if not dist:
        return _same_placement_result(func, args, kwargs)
    _, Replicate, _ = _P()
    mesh = dist[0].process_mesh
    sharded = any(not isinstance(p, Replicate) for t in dist for p in t.placements)
    if sharded and name not in _warned:
        _warned.add(name)
        warnings.warn(f"auto_parallel: no SPMD rule for '{name}', inputs are gathered and the op runs replicated")
    a2, k2 = replicate_all(args, kwargs)
    with torch._C.DisableTorchFunctionSubclass():
        res = func(*a2, **k2)
    rep = [Replicate() for _ in range(mesh.ndim)]

    def lab(o):
        if isinstance(o, torch.Tensor):
            return wrap(o, mesh, rep, tuple(o.shape))
        if isinstance(o, (list, tuple)):
            return type(o)(lab(v) for v in o)
        return o
    return lab(res)


def _flatten(x):
    if isinstance(x, (list, tuple)):
        for v in x:
            yield from _flatten(v)
    elif isinstance(x, dict):
        for v in x.values():
            yield from _flatten(v)
    else:
        yield x


# ---- matmul family ----------------------------------------------------------------------------------------------
@_register(torch.matmul, torch.Tensor.matmul, torch.Tensor.__matmul__, torch.mm, torch.Tensor.mm, torch.bmm, torch.Tensor.bmm)
def _h_matmul(x, y, **kw):
    return run_rule("matmul", lambda a, b: torch.matmul(a, b), [x, y])


@_register(torch.Tensor.__rmatmul__)
def _h_rmatmul(y, x):
    return _h_matmul(x, y)


def dist_linear(x, w, b=None, weight_is_out_in=False):
    """x @ w (+ b); ``weight_is_out_in``: torch layout [out, in] (``F.linear``), else Paddle's [in, out]."""
    y = run_rule("matmul", (lambda a, ww: torch.matmul(a, ww.t())) if weight_is_out_in else (lambda a, ww: torch.matmul(a, ww)),
                 [x, w], {"trans_y": bool(weight_is_out_in)})
    return y if b is None else _h_binary(torch.add)(y, b)


@_register(TF.linear)
def _h_linear(x, w, bias=None):
    return dist_linear(x, w, bias, weight_is_out_in=True)


# ---- element-wise -----------------------------------------------------------------------------------------------
def _h_binary(op):
    def h(a, b, *rest, **kw):
        ta, tb = isinstance(a, torch.Tensor), isinstance(b, torch.Tensor)
        if ta and tb:
            return run_rule("elementwise", lambda x, y: op(x, y, *rest, **kw), [a, b])
        if ta:
            return run_rule("elementwise", lambda x: op(x, b, *rest, **kw), [a])
        return run_rule("elementwise", lambda y: op(a, y, *rest, **kw), [b])
    return h


for _op, _names in ((torch.add, ("add", "__add__", "__radd__")), (torch.sub, ("sub", "__sub__")), (torch.mul, ("mul", "__mul__", "__rmul__")),
                    (torch.div, ("div", "__truediv__", "true_divide")), (torch.maximum, ("maximum",)), (torch.minimum, ("minimum",)),
                    (torch.pow, ("pow", "__pow__")), (torch.eq, ("eq", "__eq__")), (torch.ne, ("ne", "__ne__")),
                    (torch.gt, ("gt", "__gt__")), (torch.lt, ("lt", "__lt__")), (torch.ge, ("ge", "__ge__")), (torch.le, ("le", "__le__"))):
    _TABLE[_op] = _h_binary(_op)
    for _n in _names:
        if hasattr(torch.Tensor, _n):
            _TABLE[getattr(torch.Tensor, _n)] = _h_binary(_op)
_TABLE[torch.Tensor.__rsub__] = lambda a, b: _h_binary(torch.sub)(b, a)
_TABLE[torch.Tensor.__rtruediv__] = lambda a, b: _h_binary(torch.div)(b, a)
_TABLE[torch.true_divide] = _h_binary(torch.div)


def _h_unary(op):
    def h(x, *rest, **kw):
        return run_rule("elementwise", lambda a: op(a, *rest, **kw), [x])
    return h


for _f in (torch.relu, TF.relu, TF.gelu, TF.silu, torch.tanh, torch.sigmoid, torch.exp, torch.log, torch.sqrt, torch.rsqrt, torch.neg,
           torch.abs, torch.square, torch.clone, torch.nan_to_num, torch.clamp, torch.erf, torch.sin, torch.cos,
           torch.Tensor.relu, torch.Tensor.tanh, torch.Tensor.sigmoid, torch.Tensor.exp, torch.Tensor.log, torch.Tensor.sqrt,
           torch.Tensor.rsqrt, torch.Tensor.neg, torch.Tensor.__neg__, torch.Tensor.abs, torch.Tensor.square, torch.Tensor.clone,
           torch.Tensor.contiguous, torch.Tensor.to, torch.Tensor.float, torch.Tensor.half, torch.Tensor.bfloat16,
           torch.Tensor.double, torch.Tensor.long, torch.Tensor.int, torch.Tensor.bool, torch.Tensor.type, torch.Tensor.clamp,
           torch.Tensor.detach, torch.detach, torch.Tensor.type_as, torch.zeros_like, torch.ones_like, torch.empty_like,
           torch.Tensor.masked_fill):
    _TABLE[_f] = _h_unary(_f)


@_register(TF.dropout, torch.dropout)
def _h_dropout(x, p=0.5, training=True, inplace=False, **kw):
    """Element-wise, with the generator chosen by the placement (auto_parallel/random.py): different numbers along
    sharding axes, the same numbers along replicating axes, once ``parallel_manual_seed`` switched the control on."""
    from . import random as PR
    train = kw.pop("train", training)

    def fn(a):
        with PR.rng_for(x.process_mesh, x.placements, a.device):
            return TF.dropout(a, p, train)
    return run_rule("elementwise", fn, [x])


@_register(torch.where)
def _h_where(c, a, b):
    ts = [t for t in (c, a, b)]
    idx = [i for i, t in enumerate(ts) if isinstance(t, torch.Tensor)]

    def fn(*loc):
        vals = list(ts)
        for i, v in zip(idx, loc):
            vals[i] = v
        return torch.where(*vals)
    return run_rule("elementwise", fn, [ts[i] for i in idx])


# ---- reductions -------------------------------------------------------------------------------------------------
def _h_reduce(op, kind):
    red = {"sum": C.ReduceOp.SUM, "mean": C.ReduceOp.AVG, "max": C.ReduceOp.MAX, "min": C.ReduceOp.MIN}[kind]

    def h(x, dim=None, keepdim=False, **kw):
        if isinstance(dim, torch.Tensor) or (kind in ("max", "min") and dim is not None and op in (torch.max, torch.min)):
            a2, k2 = replicate_all((x, dim), dict(keepdim=keepdim, **kw))      # (values, indices) forms: replicated
            with torch._C.DisableTorchFunctionSubclass():
                return op(*a2, **k2)
        axis = None if dim is None or dim == [] or dim == () else dim

        def fn(a):
            if axis is None:
                return op(a, **kw) if not keepdim else op(a, dim=list(range(a.dim())), keepdim=True, **kw)
            return op(a, dim=axis, keepdim=keepdim, **kw)
        return run_rule(kind, fn, [x], {"axis": axis, "keepdim": bool(keepdim)}, out_reduce=red)
    return h


for _f, _m, _k in ((torch.sum, torch.Tensor.sum, "sum"), (torch.mean, torch.Tensor.mean, "mean"),
                   (torch.amax, torch.Tensor.amax, "max"), (torch.amin, torch.Tensor.amin, "min")):
    _TABLE[_f] = _TABLE[_m] = _h_reduce(_f, _k)


# ---- normalisation / softmax --------------------------------------------------------------------------------------
@_register(TF.softmax, torch.softmax, torch.Tensor.softmax)
def _h_softmax(x, dim=-1, **kw):
    return run_rule("softmax", lambda a: TF.softmax(a, dim=dim, **kw), [x], {"axis": dim})


@_register(TF.log_softmax, torch.log_softmax, torch.Tensor.log_softmax)
def _h_log_softmax(x, dim=-1, **kw):
    return run_rule("softmax", lambda a: TF.log_softmax(a, dim=dim, **kw), [x], {"axis": dim})


@_register(TF.layer_norm)
def _h_layer_norm(x, normalized_shape, weight=None, bias=None, eps=1e-5):
    begin = x.dim() - len(normalized_shape)
    return run_rule("layer_norm", lambda a, w, b: (TF.layer_norm(a, normalized_shape, w, b, eps),), [x, weight, bias],
                    {"begin_norm_axis": begin})[0]


@_register(getattr(TF, "rms_norm", None))
def _h_rms_norm(x, normalized_shape, weight=None, eps=None):
    begin = x.dim() - len(normalized_shape)
    return run_rule("rms_norm", lambda a, w: (TF.rms_norm(a, normalized_shape, w, eps),), [x, weight],
                    {"begin_norm_axis": begin})[0]


# ---- embedding / cross entropy ----------------------------------------------------------------------------------------
@_register(TF.embedding)
def _h_embedding(ids, weight, padding_idx=None, *a, **kw):
    _, _, Shard = _P()
    mesh = _mesh_of([ids, weight])
    row_axes = [ax for ax, p in enumerate(weight.placements) if isinstance(p, Shard) and p.dim == 0] if is_dist(weight) else []

    def fn(i, w):
        if not row_axes:
            return TF.embedding(i, w, padding_idx)
        # vocab-parallel: this rank holds rows [lo, lo + n); other ids contribute zeros to the partial sum
        coord = mesh.coord_of(C.get_rank())[row_axes[0]]
        lo = coord * w.shape[0]
        inside = (i >= lo) & (i < lo + w.shape[0])
        out = TF.embedding(torch.where(inside, i - lo, torch.zeros_like(i)), w)
        return out * inside.unsqueeze(-1).to(out.dtype)
    return run_rule("embedding", fn, [ids, weight])


@_register(TF.cross_entropy)
def _h_cross_entropy(logits, label, weight=None, size_average=None, ignore_index=-100, reduce=None, reduction="mean",
                     label_smoothing=0.0):
    """Per-token losses under the cross-entropy rule (a class-sharded logits tensor is reduced with partial statistics:
    max, sum-exp and the label's logit are combined over the class mesh axis), then the reduction's own rule."""
    _, _, Shard = _P()
    if weight is not None or label_smoothing or logits.dim() != 2:
        a2, k2 = replicate_all((logits, label), {})
        with torch._C.DisableTorchFunctionSubclass():
            out = TF.cross_entropy(*a2, weight=local_of(weight) if weight is not None else None, ignore_index=ignore_index,
                                   reduction=reduction, label_smoothing=label_smoothing)
        mesh = _mesh_of([logits, label])
        return wrap(out, mesh, [_P()[1]() for _ in range(mesh.ndim)], tuple(out.shape))
    mesh = _mesh_of([logits, label])
    cls_axes = [ax for ax, p in enumerate(logits.placements) if isinstance(p, Shard) and p.dim == 1] if is_dist(logits) else []

    def fn(lg, lab):
        if not cls_axes:
            loss = TF.cross_entropy(lg, lab, ignore_index=ignore_index, reduction="none")
        else:
            ax = cls_axes[0]
            g = _axis_group(mesh, ax)
            coord = mesh.coord_of(C.get_rank())[ax]
            lo = coord * lg.shape[1]
            lf = lg.float()
            m = lf.detach().amax(dim=1)
            comm.all_reduce(m, C.ReduceOp.MAX, g)
            e = torch.exp(lf - m[:, None])
            se = _AllReduceSum.apply(e.sum(dim=1), g)
            inside = (lab >= lo) & (lab < lo + lg.shape[1])
            picked = lf.gather(1, torch.where(inside, lab - lo, torch.zeros_like(lab))[:, None])[:, 0] * inside
            picked = _AllReduceSum.apply(picked, g)
            loss = (torch.log(se) + m - picked) * (lab != ignore_index)
        return lg, loss.unsqueeze(1)
    _, loss = run_rule("cross_entropy_with_softmax", fn, [logits, label], {"axis": 1})
    loss = _h_reduce_view(loss)
    if reduction == "none":
        return loss
    if reduction == "sum":
        return _TABLE[torch.sum](loss)
    valid = _TABLE[torch.sum](_h_binary(torch.ne)(label, ignore_index).to(torch.float32))
    return _h_binary(torch.div)(_TABLE[torch.sum](loss), valid)


def _h_reduce_view(loss):
    """[n, 1] -> [n] keeping the placements."""
    return _h_reshape_to(loss, (loss._global_shape[0],))


class _AllReduceSum(torch.autograd.Function):
    """y = sum over the group of x; every rank uses y identically afterwards, so dL/dx = dL/dy."""

    @staticmethod
    def forward(ctx, x, group):
        y = x.clone()
        comm.all_reduce(y, C.ReduceOp.SUM, group)
        return y

    @staticmethod
    def backward(ctx, g):
        return g, None


# ---- shape ops -------------------------------------------------------------------------------------------------------
def _local_shape(spec, mesh):
    shp = list(spec.shape)
    for d, ax in enumerate(spec.dims_mapping):
        if ax >= 0:
            shp[d] //= mesh.shape[ax]
    return tuple(shp)


def _h_reshape_to(x, gshape):
    mesh = x.process_mesh
    gshape = tuple(int(s) for s in gshape)
    _, outs = R.reshape(_spec_of(x, mesh), list(gshape))
    target = _local_shape(outs[0], mesh)
    return run_rule("reshape", lambda a: a.reshape(target), [x], {"shape": list(gshape)})


def _global_target(x, shape):
    """A reshape target written against the LOCAL shape (what user code sees through ``.shape``) -> the global target:
    entries equal to a sharded local dim size are scaled back up when the element counts require it."""
    shape = [int(s) for s in (shape[0] if len(shape) == 1 and isinstance(shape[0], (list, tuple, torch.Size)) else shape)]
    g_numel = math.prod(x._global_shape)
    l_numel = x.numel()
    if -1 in shape or math.prod(shape) == g_numel:
        if -1 in shape:
            known = -math.prod(shape)
            if known and g_numel % known == 0 and l_numel != g_numel:
                # ambiguous only when local sizes were used for the known dims; resolve against the local count first
                if l_numel % known == 0:
                    loc = [s if s != -1 else l_numel // known for s in shape]
                    return _scale_up(x, loc)
            return shape
        return shape
    if math.prod(shape) == l_numel:
        return _scale_up(x, shape)
    raise ValueError(f"cannot reshape a distributed tensor of global shape {x._global_shape} to {shape}")


def _scale_up(x, local_target):
    """Local target shape -> global target shape: each sharded input dim's factor goes to the output dim it leads."""
    mesh = x.process_mesh
    spec = _spec_of(x, mesh)
    out = list(local_target)
    lshape = list(x.shape)
    i = j = 0
    while i < len(lshape) and j < len(out):
        a, b, i0, j0 = lshape[i], out[j], i, j
        while a != b:
            if a < b:
                i += 1
                a *= lshape[i]
            else:
                j += 1
                b *= out[j]
        for k in range(i0, i + 1):
            ax = spec.dims_mapping[k]
            if ax >= 0:
                out[j0] *= mesh.shape[ax]
        i += 1
        j += 1
    return out


@_register(torch.reshape, torch.Tensor.reshape, torch.Tensor.view)
def _h_reshape(x, *shape):
    return _h_reshape_to(x, _global_target(x, shape))


@_register(torch.flatten, torch.Tensor.flatten)
def _h_flatten(x, start_dim=0, end_dim=-1):
    g = list(x._global_shape)
    nd = len(g)
    s, e = start_dim % nd, end_dim % nd
    return _h_reshape_to(x, g[:s] + [math.prod(g[s:e + 1])] + g[e + 1:])


@_register(torch.unsqueeze, torch.Tensor.unsqueeze)
def _h_unsqueeze(x, dim):
    g = list(x._global_shape)
    d = dim % (len(g) + 1)
    return _h_reshape_to(x, g[:d] + [1] + g[d:])


@_register(torch.squeeze, torch.Tensor.squeeze)
def _h_squeeze(x, dim=None):
    g = list(x._global_shape)
    if dim is None:
        tgt = [s for s in g if s != 1]
    else:
        dims = [d % len(g) for d in (dim if isinstance(dim, (list, tuple)) else [dim])]
        tgt = [s for i, s in enumerate(g) if not (i in dims and s == 1)]
    return _h_reshape_to(x, tgt)


@_register(torch.transpose, torch.Tensor.transpose)
def _h_transpose(x, d0, d1):
    perm = list(range(x.dim()))
    perm[d0], perm[d1] = perm[d1], perm[d0]
    return run_rule("transpose", lambda a: a.permute(perm), [x], {"perm": perm})


@_register(torch.permute, torch.Tensor.permute)
def _h_permute(x, *dims):
    perm = [int(d) % x.dim() for d in (dims[0] if len(dims) == 1 and isinstance(dims[0], (list, tuple)) else dims)]
    return run_rule("transpose", lambda a: a.permute(perm), [x], {"perm": perm})


@_register(torch.t, torch.Tensor.t)
def _h_t(x):
    return _h_transpose(x, 0, 1) if x.dim() == 2 else x


_TABLE[torch.Tensor.T.__get__] = lambda x: _h_permute(x, *reversed(range(x.dim())))
_TABLE[torch.Tensor.mT.__get__] = lambda x: _h_transpose(x, -2, -1)


@_register(torch.cat, torch.concat, torch.concatenate)
def _h_cat(tensors, dim=0, **kw):
    tensors = list(tensors)
    return run_rule(lambda *specs, axis=0: R.concat(list(specs), axis), lambda *loc: torch.cat(loc, dim=dim), tensors, {"axis": dim})


def _split_like(is_chunk):
    """torch.split takes a piece size (or sizes), torch.chunk a number of pieces."""

    def h(x, arg, dim=0):
        d = dim % x.dim()
        g = x._global_shape[d]
        if is_chunk:
            secs = [g // arg] * arg if g % arg == 0 else None
        elif isinstance(arg, int):
            secs = [arg] * (g // arg) if g % arg == 0 else None
        else:
            secs = list(arg)
        if secs is None:
            a2, _ = replicate_all((x,), {})
            mesh = x.process_mesh
            with torch._C.DisableTorchFunctionSubclass():
                res = torch.chunk(a2[0], arg, dim) if is_chunk else torch.split(a2[0], arg, dim)
            return tuple(wrap(o, mesh, [_P()[1]() for _ in range(mesh.ndim)], tuple(o.shape)) for o in res)
        return run_rule("split", lambda a: torch.split(a, secs, dim=d), [x], {"num_or_sections": secs, "axis": d})
    return h


for _f in (torch.split, torch.Tensor.split):
    _TABLE[_f] = _split_like(False)
for _f in (torch.chunk, torch.Tensor.chunk):
    _TABLE[_f] = _split_like(True)


@_register(torch.Tensor.__getitem__)
def _h_getitem(x, idx):
    """Basic indexing that leaves every sharded dim whole (full slices there) runs on the local shard; anything else
    gathers first."""
    Partial, Replicate, Shard = _P()
    items = list(idx) if isinstance(idx, tuple) else [idx]
    simple = all(isinstance(i, (slice, int)) or i is None or i is Ellipsis for i in items) and not any(
        isinstance(p, Partial) for p in x.placements)
    if simple:
        nd = x.dim()
        n_real = sum(1 for i in items if i is not None and i is not Ellipsis)
        expanded = []
        for i in items:
            if i is Ellipsis:
                expanded += [slice(None)] * (nd - n_real)
            else:
                expanded.append(i)
        expanded += [slice(None)] * (nd - sum(1 for i in expanded if i is not None))
        sharded_dims = {p.dim: ax for ax, p in enumerate(x.placements) if isinstance(p, Shard)}
        new_pl = {}
        in_d = out_d = 0
        ok = True
        for i in expanded:
            if i is None:
                out_d += 1
                continue
            if in_d in sharded_dims:
                if not (isinstance(i, slice) and i == slice(None)):
                    ok = False
                    break
                new_pl[sharded_dims[in_d]] = out_d
            if isinstance(i, int):
                in_d += 1
            else:
                in_d += 1
                out_d += 1
        if ok:
            with torch._C.DisableTorchFunctionSubclass():
                loc = local_of(x)[idx]
            mesh = x.process_mesh
            pl = [Shard(new_pl[ax]) if ax in new_pl else Replicate() for ax in range(mesh.ndim)]
            gshape = list(loc.shape)
            for ax, d in new_pl.items():
                gshape[d] *= mesh.shape[ax]
            return wrap(loc, mesh, pl, tuple(gshape))
    a2, _ = replicate_all((x,), {})
    mesh = x.process_mesh
    with torch._C.DisableTorchFunctionSubclass():
        out = a2[0][local_of(idx) if isinstance(idx, torch.Tensor) else idx]
    return wrap(out, mesh, [Replicate() for _ in range(mesh.ndim)], tuple(out.shape))


# ---- values that only make sense whole --------------------------------------------------------------------------------
def _resolved(t):
    """A Partial tensor as a plain local tensor holding the reduced value (differentiable); others unchanged."""
    Partial, Replicate, Shard = _P()
    if is_dist(t) and any(isinstance(p, Partial) for p in t.placements):
        mesh = t.process_mesh
        dst = [Replicate() if isinstance(p, Partial) else p for p in t.placements]
        return reshard_local(local_of(t), mesh, list(t.placements), dst)
    return local_of(t)


@_register(torch.Tensor.backward)
def _h_backward(self, gradient=None, retain_graph=None, create_graph=False, inputs=None):
    """``loss.backward()`` on a loss that is still a partial sum (e.g. the mean over a batch-sharded dim): the pending
    reduction is part of the function being differentiated, so it is applied first."""
    with torch._C.DisableTorchFunctionSubclass():
        if is_placeholder(self):
            # not a member of the loss's mesh: nothing to seed, but the walk back through this rank's graph must happen
            # (it reaches the cross-mesh transfers that bring this rank's share of the gradient)
            loc = local_of(self)
            if loc.requires_grad:
                torch.Tensor.backward(loc, torch.zeros_like(loc), retain_graph, create_graph, inputs=inputs)
            return
        torch.Tensor.backward(_resolved(self), local_of(gradient) if gradient is not None else None, retain_graph, create_graph,
                              inputs=inputs)


def _value_reader(method):
    def h(self, *a, **kw):
        if is_placeholder(self):
            return float("nan")                     # the value lives on the ranks of the tensor's mesh
        with torch._C.DisableTorchFunctionSubclass():
            return method(_resolved(self).detach(), *a, **kw)
    return h


for _m in ("item", "tolist", "numpy", "__float__", "__int__", "__bool__"):
    _TABLE[getattr(torch.Tensor, _m)] = _value_reader(getattr(torch.Tensor, _m))


# ---------------------------------------------------------------------------------------- tensors on other ranks' meshes
# A rank that is not a member of a tensor's mesh holds a PLACEHOLDER for it: a zero-element local tensor with the right
# dtype, the global shape and the placements.  Ops on placeholders compute nothing, but keep the autograd graph
# connected, so that ``loss.backward()`` on every rank walks back to the cross-mesh transfers, whose backward moves the
# gradient to the ranks that own the producer - pipeline stages as meshes, with no schedule object (the reference runs
# dygraph semi-auto pipelines the same way: reshard between the stage meshes, api.py reshard + same_status reshard).
def in_mesh(mesh):
    return mesh.coord_of(C.get_rank()) is not None


def is_placeholder(t):
    return getattr(t, "_placeholder", False)


def make_placeholder(mesh, placements, global_shape, dtype, device, like=None):
    loc = torch.empty(0, dtype=dtype, device=device)
    out = wrap(loc if like is None else like, mesh, placements, global_shape)
    out._placeholder = True
    return out


class _Skip(torch.autograd.Function):
    """n zero-element outputs that depend (for autograd's bookkeeping only) on the inputs."""

    @staticmethod
    def forward(ctx, n_out, dtype, device, *inputs):
        ctx.shapes = [(tuple(i.shape), i.dtype, i.device) if isinstance(i, torch.Tensor) else None for i in inputs]
        outs = tuple(torch.empty(0, dtype=dtype, device=device) for _ in range(n_out))
        return outs

    @staticmethod
    def backward(ctx, *grads):
        return (None, None, None) + tuple(None if s is None else torch.zeros(s[0], dtype=s[1], device=s[2]) for s in ctx.shapes)


def _local_shape_of(global_shape, mesh, placements):
    _, _, Shard = _P()
    shp = list(global_shape)
    for ax, p in enumerate(placements):
        if isinstance(p, Shard):
            shp[p.dim] //= mesh.shape[ax]
    return tuple(shp)


class _CrossMesh(torch.autograd.Function):
    """Same placements, other ranks: rank i of the source mesh hands its shard to rank i of the destination mesh."""

    @staticmethod
    def forward(ctx, t, src_mesh, dst_mesh, local_shape, dtype):
        me = C.get_rank()
        ctx.cfg = (src_mesh, dst_mesh, tuple(local_shape), dtype, t.device)
        out = None
        for s_rank, d_rank in zip(src_mesh.process_ids, dst_mesh.process_ids):
            if s_rank == d_rank:
                if me == s_rank:
                    out = t.view_as(t)
                continue
            if me == s_rank:
                comm.send(t.detach().contiguous(), d_rank)
            elif me == d_rank:
                out = torch.empty(local_shape, dtype=dtype, device=t.device)
                comm.recv(out, s_rank)
        return out if out is not None else torch.empty(0, dtype=dtype, device=t.device)

    @staticmethod
    def backward(ctx, g):
        src_mesh, dst_mesh, local_shape, dtype, device = ctx.cfg
        me = C.get_rank()
        out = None
        for s_rank, d_rank in zip(src_mesh.process_ids, dst_mesh.process_ids):
            if s_rank == d_rank:
                if me == s_rank:
                    out = g
                continue
            if me == d_rank:
                comm.send(g.detach().contiguous(), s_rank)
            elif me == s_rank:
                out = torch.empty(local_shape, dtype=dtype, device=device)
                comm.recv(out, d_rank)
        if out is None:
            out = torch.empty(0, dtype=dtype, device=device)
        return out, None, None, None, None


def cross_mesh(t, dst_mesh):
    """DistTensor on its mesh -> the same placements on ``dst_mesh`` (a mesh of the same shape)."""
    src_mesh = t.process_mesh
    assert list(src_mesh.shape) == list(dst_mesh.shape), "cross-mesh reshard needs meshes of the same shape"
    lshape = _local_shape_of(t._global_shape, src_mesh, t.placements)
    loc = local_of(t)
    if torch.is_grad_enabled() and loc.requires_grad:
        out = _CrossMesh.apply(loc, src_mesh, dst_mesh, lshape, loc.dtype)
    else:
        out = _CrossMesh.forward(type("ctx", (), {})(), loc.detach(), src_mesh, dst_mesh, lshape, loc.dtype)
    res = wrap(out, dst_mesh, t.placements, t._global_shape)
    if not in_mesh(dst_mesh):
        res._placeholder = True
    return res
