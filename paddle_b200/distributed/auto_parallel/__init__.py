"""Semi-auto parallel dygraph API (reference: python/paddle/distributed/auto_parallel/api.py:131-2846,
process_mesh.py, placement_type.py; native DistTensor/reshard under paddle/phi/core/distributed/auto_parallel/).

A ``DistTensor`` here is a local torch tensor plus ``(process_mesh, placements)`` metadata: ``shard_tensor`` cuts the
local shard out of the global value, ``reshard`` moves between placements with the matching collective
(r->s slice, s->r all-gather, p->r all-reduce, p->s reduce-scatter, s->s all-to-all), ``unshard_dtensor`` rebuilds
the global value, and a change of mesh (same shape, other ranks) is a send / recv pair.  SPMD propagation (N13) lives
in ``spmd_rules``: ``dist_op(name, fn, *dtensors)`` infers the shardings with the op's rule, reshards the inputs the
rule asks for, runs ``fn`` on the local shards and labels the outputs (partial outputs stay ``Partial`` until someone
reshards them)."""
from __future__ import annotations

import numpy as np
import torch

from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm


class Placement:
    def is_shard(self, dim=None):
        return False

    def is_replicated(self):
        return False

    def is_partial(self):
        return False


class Shard(Placement):
    def __init__(self, dim):
        self.dim = dim

    def get_dim(self):
        return self.dim

    def is_shard(self, dim=None):
        return dim is None or dim == self.dim

    def __eq__(self, o):
        return isinstance(o, Shard) and o.dim == self.dim

    def __hash__(self):
        return hash(("S", self.dim))

    def __repr__(self):
        return f"Shard(dim={self.dim})"


class Replicate(Placement):
    def is_replicated(self):
        return True

    def __eq__(self, o):
        return isinstance(o, Replicate)

    def __hash__(self):
        return hash("R")

    def __repr__(self):
        return "Replicate()"


class Partial(Placement):
    def __init__(self, reduce_type=None):
        self.reduce_type = reduce_type if reduce_type is not None else C.ReduceOp.SUM

    def is_partial(self):
        return True

    def __eq__(self, o):
        return isinstance(o, Partial)

    def __hash__(self):
        return hash("P")

    def __repr__(self):
        return "Partial()"


class ProcessMesh:
    def __init__(self, mesh=None, dim_names=None, shape=None, process_ids=None):
        arr = np.array(mesh if mesh is not None else np.array(process_ids).reshape(shape))
        self._mesh = arr
        self._shape = list(arr.shape)
        self._process_ids = arr.flatten().tolist()
        self._dim_names = list(dim_names) if dim_names is not None else [f"d{i}" for i in range(arr.ndim)]
        self._groups = {}

    @property
    def shape(self):
        return self._shape

    @property
    def process_ids(self):
        return self._process_ids

    @property
    def dim_names(self):
        return self._dim_names

    @property
    def ndim(self):
        return self._mesh.ndim

    @property
    def mesh(self):
        return self._mesh

    def get_dim_size(self, name):
        return self._shape[self._dim_names.index(name)]

    def coord_of(self, rank):
        idx = np.argwhere(self._mesh == rank)
        return tuple(idx[0].tolist()) if len(idx) else None

    def get_mesh_with_dim(self, dim_name, index=None):
        axis = self._dim_names.index(dim_name)
        moved = np.moveaxis(self._mesh, axis, 0)
        names = [dim_name] + [n for n in self._dim_names if n != dim_name]
        if index is not None:
            return ProcessMesh(moved[index], names[1:])
        return ProcessMesh(moved, names)

    def axis_group(self, axis):
        """Communicator along mesh axis ``axis`` containing the calling rank (created collectively, cached)."""
        if axis in self._groups:
            return self._groups[axis]
        me = C.get_rank()
        moved = np.moveaxis(self._mesh, axis, -1).reshape(-1, self._shape[axis])
        mine = None
        for row in moved:
            g = C.new_group(ranks=[int(r) for r in row]) if C.get_world_size() > 1 else None
            if me in row.tolist():
                mine = g
        self._groups[axis] = mine
        return mine

    def __eq__(self, o):
        return isinstance(o, ProcessMesh) and np.array_equal(o._mesh, self._mesh) and o._dim_names == self._dim_names

    def __hash__(self):
        return hash((tuple(self._process_ids), tuple(self._shape)))

    def __repr__(self):
        return f"ProcessMesh(shape={self._shape}, process_ids={self._process_ids}, dim_names={self._dim_names})"


class DistAttr:
    def __init__(self, mesh, sharding_specs):
        self.process_mesh = mesh
        self.sharding_specs = sharding_specs


def _attach(t, mesh, placements, global_shape):
    t.process_mesh = mesh
    t.placements = list(placements)
    coord = mesh.coord_of(C.get_rank()) or tuple([0] * mesh.ndim)
    off = [0] * len(global_shape)
    for ax, pl in enumerate(placements):
        if isinstance(pl, Shard):
            off[pl.dim] += coord[ax] * t.shape[pl.dim]
    t._dist_meta = (tuple(global_shape), tuple(off))
    t._global_shape = tuple(global_shape)
    return t


def is_dist(t):
    return hasattr(t, "placements") and hasattr(t, "process_mesh")


def shard_tensor(data, mesh, placements, dtype=None, place=None, stop_gradient=None):
    from paddle_b200.tensor.creation import to_tensor
    t = data if isinstance(data, torch.Tensor) else to_tensor(data, dtype=dtype, place=place)
    gshape = tuple(t.shape)
    coord = mesh.coord_of(C.get_rank()) or tuple([0] * mesh.ndim)
    local = t.detach()
    for ax, pl in enumerate(placements):
        if isinstance(pl, Shard):
            n = mesh.shape[ax]
            assert local.shape[pl.dim] % n == 0, f"dim {pl.dim} ({local.shape[pl.dim]}) not divisible by mesh axis {n}"
            local = local.chunk(n, dim=pl.dim)[coord[ax]]
    local = local.contiguous().clone()
    if isinstance(data, torch.nn.Parameter):
        from paddle_b200.framework.tensor_patch import Parameter
        out = Parameter(local, data.requires_grad, name=getattr(data, "name", None))
    else:
        out = local
        sg = stop_gradient if stop_gradient is not None else not getattr(data, "requires_grad", False)
        out.stop_gradient = sg
    return _attach(out, mesh, placements, gshape)


def dtensor_from_fn(fn, mesh, placements, *args, **kwargs):
    return shard_tensor(fn(*args, **kwargs), mesh, placements)


def dtensor_from_local(local_tensor, mesh, placements):
    """Wrap this rank's local shard as a distributed tensor (reference: auto_parallel/api.py:499-520): the global
    shape is the local shape multiplied by the mesh size along every sharded dim."""
    gshape = list(local_tensor.shape)
    for ax, pl in enumerate(placements):
        if isinstance(pl, Shard):
            gshape[pl.dim] *= mesh.shape[ax]
    return _attach(local_tensor, mesh, placements, tuple(gshape))


def dtensor_to_local(dist_tensor, mesh=None, placements=None):
    """The local shard of a distributed tensor, as a plain tensor."""
    out = dist_tensor.detach().clone() if not dist_tensor.requires_grad else dist_tensor * 1
    for attr in ("process_mesh", "placements", "_global_shape"):
        if hasattr(out, attr):
            try:
                delattr(out, attr)
            except AttributeError:
                pass
    return out


_g_mesh = None


def get_mesh():
    return _g_mesh


def set_mesh(mesh):
    global _g_mesh
    _g_mesh = mesh


def create_mesh(mesh_dims):
    """``create_mesh([("dp", 2), ("mp", 4)])`` -> global ProcessMesh over all ranks (reference: interface.py:326)."""
    import numpy as np
    names = [n for n, _ in mesh_dims]
    shape = [d for _, d in mesh_dims]
    mesh = ProcessMesh(np.arange(int(np.prod(shape))).reshape(shape).tolist(), names)
    set_mesh(mesh)
    return mesh


class ReduceType:
    """Reduce kinds of a Partial placement (reference: paddle.base.core.ReduceType)."""
    kRedSum, kRedMax, kRedMin, kRedProd, kRedAvg, kRedAny, kRedAll = range(7)


def _gather_axis(t, group, dim):
    parts = []
    comm.all_gather(parts, t.contiguous(), group)
    return torch.cat(parts, dim=dim)


def _cross_mesh(dist_tensor, mesh):
    """Same placements on another mesh of the same shape: rank i of the source mesh sends its shard to rank i of the
    destination mesh (reference: same_status_reshard_function.cc)."""
    src_mesh = dist_tensor.process_mesh
    assert list(src_mesh.shape) == list(mesh.shape), "cross-mesh reshard needs meshes of the same shape"
    me = C.get_rank()
    t = dist_tensor.detach()
    out = t
    for s_rank, d_rank in zip(src_mesh.process_ids, mesh.process_ids):
        if s_rank == d_rank:
            continue
        if me == s_rank:
            comm.send(t.contiguous(), d_rank)
        elif me == d_rank:
            out = torch.empty_like(t)
            comm.recv(out, s_rank)
    return _attach(out, mesh, dist_tensor.placements, dist_tensor._global_shape)


def reshard(dist_tensor, mesh, placements):
    assert is_dist(dist_tensor), "reshard expects a tensor produced by shard_tensor"
    if dist_tensor.process_mesh != mesh and C.get_world_size() > 1:
        dist_tensor = _cross_mesh(dist_tensor, mesh)
    src = dist_tensor.placements
    gshape = dist_tensor._global_shape
    t = dist_tensor.detach() if not dist_tensor.requires_grad else dist_tensor
    coord = mesh.coord_of(C.get_rank()) or tuple([0] * mesh.ndim)
    for ax, (a, b) in enumerate(zip(src, placements)):
        if a == b:
            continue
        n = mesh.shape[ax]
        g = mesh.axis_group(ax) if n > 1 else None
        if isinstance(a, Shard) and isinstance(b, Replicate):
            t = _gather_axis(t, g, a.dim) if n > 1 else t
        elif isinstance(a, Replicate) and isinstance(b, Shard):
            t = t.chunk(n, dim=b.dim)[coord[ax]].contiguous()
        elif isinstance(a, Partial) and isinstance(b, Replicate):
            t = t.clone()
            if n > 1:
                comm.all_reduce(t, a.reduce_type, g)
        elif isinstance(a, Partial) and isinstance(b, Shard):
            # p -> s: one reduce-scatter (reference p_to_s_reshard_function.cc), not all-reduce + slice
            if n > 1 and t.shape[b.dim] % n == 0:
                moved = t.movedim(b.dim, 0).contiguous()
                out = torch.empty((moved.shape[0] // n,) + tuple(moved.shape[1:]), dtype=t.dtype, device=t.device)
                comm.reduce_scatter(out, moved, a.reduce_type, g)
                t = out.movedim(0, b.dim).contiguous()
            else:
                t = t.clone()
                if n > 1:
                    comm.all_reduce(t, a.reduce_type, g)
                t = t.chunk(n, dim=b.dim)[coord[ax]].contiguous()
        elif isinstance(a, Shard) and isinstance(b, Shard):
            # s -> s on another dim: one all-to-all (reference s_to_s_reshard_function.cc)
            if n > 1 and t.shape[b.dim] % n == 0:
                pieces = [c.contiguous() for c in t.chunk(n, dim=b.dim)]
                recv = [torch.empty_like(pieces[0]) for _ in range(n)]
                comm.alltoall(recv, pieces, g)
                t = torch.cat(recv, dim=a.dim)
            else:
                full = _gather_axis(t, g, a.dim) if n > 1 else t
                t = full.chunk(n, dim=b.dim)[coord[ax]].contiguous()
        elif isinstance(a, Replicate) and isinstance(b, Partial):
            t = t if coord[ax] == 0 else torch.zeros_like(t)
        elif isinstance(a, Shard) and isinstance(b, Partial):
            full = _gather_axis(t, g, a.dim) if n > 1 else t
            t = full if coord[ax] == 0 else torch.zeros_like(full)
        else:
            raise NotImplementedError(f"reshard {a} -> {b}")
    return _attach(t, mesh, placements, gshape)


def dist_op(op_name, fn, *dtensors, **attrs):
    """Run ``fn`` on the local shards of ``dtensors`` under the SPMD rule ``op_name`` (see ``spmd_rules``): inputs are
    resharded to what the rule requires, outputs come back labelled with the inferred placements."""
    from . import spmd_rules as R
    mesh = next(t.process_mesh for t in dtensors if is_dist(t))
    specs = []
    for t in dtensors:
        assert is_dist(t), "dist_op takes distributed tensors (see shard_tensor)"
        specs.append(R.DistTensorSpec.from_placements(t._global_shape, t.placements))
    new_specs, out_specs = R.infer_forward(op_name, *specs, **attrs)
    locals_ = []
    for t, want in zip(dtensors, new_specs):
        pl = want.placements(mesh.ndim)
        # a partial input the rule did not ask to keep partial is reduced first
        cur = [p if not p.is_partial() else Replicate() for p in t.placements] if any(p.is_partial() for p in t.placements) else t.placements
        if any(p.is_partial() for p in t.placements):
            t = reshard(t, mesh, cur)
        locals_.append(t if list(t.placements) == list(pl) else reshard(t, mesh, pl))
    res = fn(*locals_)
    outs = res if isinstance(res, (tuple, list)) else [res]
    labelled = [_attach(o, mesh, sp.placements(mesh.ndim), sp.shape) for o, sp in zip(outs, out_specs)]
    return labelled if isinstance(res, (tuple, list)) else labelled[0]


def unshard_dtensor(dist_tensor):
    if not is_dist(dist_tensor):
        return dist_tensor
    mesh = dist_tensor.process_mesh
    out = reshard(dist_tensor, mesh, [Replicate() for _ in dist_tensor.placements])
    for attr in ("placements", "process_mesh", "_dist_meta", "_global_shape"):
        try:
            delattr(out, attr)
        except AttributeError:
            pass
    return out


def shard_layer(layer, process_mesh, shard_fn=None, input_fn=None, output_fn=None):
    if shard_fn is not None:
        for name, sub in layer.named_sublayers(include_self=True):
            shard_fn(name, sub, process_mesh)
    else:
        for sub in layer.sublayers(include_self=True):
            for pname, p in list(sub._parameters.items()):
                if p is not None and not is_dist(p):
                    sub._parameters[pname] = shard_tensor(p, process_mesh, [Replicate() for _ in range(process_mesh.ndim)])
    if input_fn is not None:
        layer.register_forward_pre_hook(lambda l, inp: input_fn(inp, process_mesh))
    if output_fn is not None:
        layer.register_forward_post_hook(lambda l, inp, out: output_fn(out, process_mesh))
    return layer


class _ShardingStage:
    def __init__(self, mesh=None, sharding_mesh_dim=None):
        self.mesh, self.sharding_mesh_dim = mesh, sharding_mesh_dim


class ShardingStage1(_ShardingStage):
    level = "os"


class ShardingStage2(_ShardingStage):
    level = "os_g"


class ShardingStage3(_ShardingStage):
    level = "p_g_os"


def shard_optimizer(optimizer, shard_fn=None, gradient_accumulation_steps=1):
    optimizer._shard_fn = shard_fn
    optimizer._gradient_accumulation_steps = gradient_accumulation_steps
    return optimizer


def shard_dataloader(dataloader, meshes, input_keys=None, shard_dims=None, is_dataset_splitted=False):
    return dataloader


def shard_scaler(scaler):
    return scaler


class Strategy:
    """Reference: auto_parallel/strategy.py - config bag with the same section names."""

    def __init__(self, config=None):
        cfg = config or {}

        class _S(dict):
            __getattr__ = dict.get

            def __setattr__(self, k, v):
                self[k] = v
        self.sharding = _S(enable=False, stage=1, degree=8, **cfg.get("sharding", {}))
        self.fused_passes = _S(enable=False, fused_passes_list=[], **cfg.get("fused_passes", {}))
        self.gradient_merge = _S(enable=False, k_steps=1, avg=True, **cfg.get("gradient_merge", {}))
        self.pipeline = _S(enable=False, schedule_mode="1F1B", micro_batch_size=1, accumulate_steps=1, **cfg.get("pipeline", {}))
        self.amp = _S(enable=False, dtype="float16", level="O1", init_loss_scaling=32768.0, **cfg.get("amp", {}))
        self.recompute = _S(enable=False, **cfg.get("recompute", {}))


class DistModel:
    """``to_static`` result (reference: api.py DistModel).  There is no static graph here: the object drives the
    same dygraph layer with train/eval/predict modes and an optional loss + optimizer."""

    def __init__(self, layer, loader=None, loss=None, optimizer=None, strategy=None):
        self._layer, self._loss, self._opt, self._strategy = layer, loss, optimizer, strategy
        self._mode = "train" if optimizer is not None and loss is not None else ("eval" if loss is not None else "predict")

    def train(self):
        self._mode = "train"
        self._layer.train()

    def eval(self):
        self._mode = "eval"
        self._layer.eval()

    def predict(self):
        self._mode = "predict"
        self._layer.eval()

    def __call__(self, *args):
        if self._mode == "predict":
            with torch.no_grad():
                return self._layer(*args)
        *inputs, label = args
        if self._mode == "eval":
            with torch.no_grad():
                return self._loss(self._layer(*inputs), label)
        loss = self._loss(self._layer(*inputs), label)
        loss.backward()
        self._opt.step()
        self._opt.clear_grad()
        return loss.detach()

    def state_dict(self, mode="all"):
        sd = dict(self._layer.state_dict())
        if mode in ("all", "opt") and self._opt is not None:
            sd.update(self._opt.state_dict())
        return sd

    def set_state_dict(self, sd):
        self._layer.set_state_dict({k: v for k, v in sd.items() if k in self._layer.state_dict()})


def to_static(layer, loader=None, loss=None, optimizer=None, strategy=None):
    return DistModel(layer, loader, loss, optimizer, strategy)
