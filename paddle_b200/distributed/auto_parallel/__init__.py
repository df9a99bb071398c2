"""Semi-auto parallel dygraph API (reference: python/paddle/distributed/auto_parallel/api.py:131-2846,
process_mesh.py, placement_type.py; native DistTensor/reshard under paddle/phi/core/distributed/auto_parallel/).

A ``DistTensor`` (``dtensor.py``) is the local shard as a ``torch.Tensor`` subclass carrying ``(process_mesh,
placements)``: ``shard_tensor`` cuts the local shard out of the global value, ``reshard`` moves between placements with
the matching collective (r->s slice, s->r all-gather, p->r all-reduce, p->s reduce-scatter, s->s all-to-all; every step
differentiable, its backward is the dual transfer), ``unshard_dtensor`` rebuilds the global value, and a change of mesh
(same shape, other ranks) is a send / recv pair.  Ordinary ops on DistTensors propagate the sharding by themselves:
``__torch_function__`` looks up the op's SPMD rule (N13, ``spmd_rules``), reshards the inputs the rule asks for, runs the
op on the local shards and labels the outputs (partial outputs stay ``Partial`` until someone needs them whole);
``dist_op(name, fn, *dtensors)`` is the same machinery for ops that are not torch functions.  ``shard_layer`` /
``shard_optimizer`` / ``shard_dataloader`` / ``shard_scaler`` build a semi-auto training step out of it."""
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
    def unique_id(self):
        """A number identifying the mesh (shape + process ids + dim names), equal for equal meshes."""
        import zlib
        return zlib.crc32(repr((self._shape, self._process_ids, self._dim_names)).encode())

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

    def __getitem__(self, index):
        """Sub-mesh by numpy-style indexing of the rank array (reference: process_mesh.py __getitem__)."""
        sub = self._mesh[index]
        if np.ndim(sub) == 0:
            return ProcessMesh([int(sub)], [self._dim_names[-1]])
        idx = index if isinstance(index, tuple) else (index,)
        kept = [n for i, n in enumerate(self._dim_names) if i >= len(idx) or not isinstance(idx[i], (int, np.integer))]
        return ProcessMesh(sub, kept)

    def __contains__(self, rank):
        return int(rank) in self._process_ids

    def get_rank_by_dim_and_process_id(self, dim, process_id):
        """Coordinate of ``process_id`` along mesh dim ``dim`` (name or index); -1 if the rank is not in the mesh."""
        c = self.coord_of(process_id)
        if c is None:
            return -1
        return c[self._dim_names.index(dim) if isinstance(dim, str) else int(dim)]

    def get_submesh_with_dim(self, dim_name):
        """The 1-D mesh along ``dim_name`` that contains the calling rank."""
        axis = self._dim_names.index(dim_name)
        c = self.coord_of(C.get_rank())
        if c is None:
            return None
        idx = tuple(slice(None) if i == axis else c[i] for i in range(self.ndim))
        return ProcessMesh(self._mesh[idx], [dim_name])

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
    from . import dtensor as D
    return D._attach(t, mesh, placements, global_shape)


def is_dist(t):
    return isinstance(t, torch.Tensor) and hasattr(t, "placements") and hasattr(t, "process_mesh")


def shard_tensor(data, mesh, placements, dtype=None, place=None, stop_gradient=None):
    """Cut this rank's shard out of the global ``data`` (reference: api.py:131-300).  Parameters come back as
    parameters (``DistParameter``), other tensors as leaf ``DistTensor``s."""
    from paddle_b200.tensor.creation import to_tensor
    from . import dtensor as D
    t = data if isinstance(data, torch.Tensor) else to_tensor(data, dtype=dtype, place=place)
    gshape = tuple(t.shape)
    if mesh.coord_of(C.get_rank()) is None and C.get_world_size() > 1:
        # not a member of the mesh: keep a zero-element placeholder (dtype, global shape, placements)
        empty = torch.empty(0, dtype=t.dtype, device=t.device)
        if isinstance(data, torch.nn.Parameter):
            out = D.make_parameter(empty, mesh, placements, gshape, like=data)
        else:
            out = D.make_leaf(empty, mesh, placements, gshape, requires_grad=False)
        out._placeholder = True
        return out
    coord = mesh.coord_of(C.get_rank()) or tuple([0] * mesh.ndim)
    local = D.local_of(t).detach()
    for ax, pl in enumerate(placements):
        if isinstance(pl, Shard):
            n = mesh.shape[ax]
            assert local.shape[pl.dim] % n == 0, f"dim {pl.dim} ({local.shape[pl.dim]}) not divisible by mesh axis {n}"
            local = local.chunk(n, dim=pl.dim)[coord[ax]]
        elif isinstance(pl, Partial) and coord[ax] != 0:
            local = torch.zeros_like(local)
    local = local.contiguous().clone()
    if isinstance(data, torch.nn.Parameter):
        return D.make_parameter(local, mesh, placements, gshape, like=data)
    sg = stop_gradient if stop_gradient is not None else not getattr(data, "requires_grad", False)
    out = D.make_leaf(local, mesh, placements, gshape, requires_grad=not sg and local.is_floating_point())
    return out


def dtensor_from_fn(fn, mesh, placements, *args, **kwargs):
    return shard_tensor(fn(*args, **kwargs), mesh, placements)


def dtensor_from_local(local_tensor, mesh, placements):
    """Wrap this rank's local shard as a distributed tensor (reference: auto_parallel/api.py:499-520): the global
    shape is the local shape multiplied by the mesh size along every sharded dim.  Differentiable: the result stays
    attached to ``local_tensor``'s autograd graph."""
    from . import dtensor as D
    gshape = list(local_tensor.shape)
    for ax, pl in enumerate(placements):
        if isinstance(pl, Shard):
            gshape[pl.dim] *= mesh.shape[ax]
    return D.wrap(D.local_of(local_tensor).view_as(local_tensor) if local_tensor.requires_grad else local_tensor,
                  mesh, placements, tuple(gshape))


def dtensor_to_local(dist_tensor, mesh=None, placements=None):
    """The local shard of a distributed tensor, as a plain tensor (same autograd graph)."""
    from . import dtensor as D
    loc = D.local_of(dist_tensor)
    return loc.view_as(loc) if loc.requires_grad else loc.detach()


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
    destination mesh (reference: same_status_reshard_function.cc).  Differentiable; ranks outside a mesh hold
    placeholders (dtensor.py)."""
    from . import dtensor as D
    return D.cross_mesh(dist_tensor, mesh)


def reshard(dist_tensor, mesh, placements):
    """Move a distributed tensor to other placements (and, for a mesh of the same shape, other ranks).  One collective
    per mesh axis that changes (reference: reshard/{r_to_s,s_to_r,p_to_r,p_to_s,s_to_s,r_to_p,same_status}_reshard_function.cc);
    differentiable - the backward of each step is the dual transfer."""
    from . import dtensor as D
    assert is_dist(dist_tensor), "reshard expects a tensor produced by shard_tensor"
    if dist_tensor.process_mesh != mesh and C.get_world_size() > 1:
        dist_tensor = _cross_mesh(dist_tensor, mesh)
    gshape = dist_tensor._global_shape
    if not D.in_mesh(mesh):
        out = D.wrap(D.local_of(dist_tensor).view_as(D.local_of(dist_tensor)), mesh, placements, gshape)
        out._placeholder = True
        return out
    loc = D.reshard_local(D.local_of(dist_tensor), mesh, list(dist_tensor.placements), list(placements))
    if loc is D.local_of(dist_tensor) or loc.data_ptr() == dist_tensor.data_ptr():
        loc = loc.view_as(loc) if loc.requires_grad else loc.detach()        # a fresh label, not a relabelled input
    return D.wrap(loc, mesh, placements, gshape)


def dist_op(op_name, fn, *dtensors, **attrs):
    """Run ``fn`` on the local shards of ``dtensors`` under the SPMD rule ``op_name`` (see ``spmd_rules``): inputs are
    resharded to what the rule requires, outputs come back labelled with the inferred placements.  This is what every
    torch function does by itself on DistTensors; use it for ops that are not torch functions."""
    from . import dtensor as D
    return D.run_rule(op_name, fn, list(dtensors), attrs)


def unshard_dtensor(dist_tensor):
    """The global value as a plain tensor (reference: api.py unshard_dtensor)."""
    from . import dtensor as D
    if not is_dist(dist_tensor):
        return dist_tensor
    mesh = dist_tensor.process_mesh
    loc = D.reshard_local(D.local_of(dist_tensor), mesh, list(dist_tensor.placements), [Replicate() for _ in range(mesh.ndim)])
    return loc.view_as(loc) if loc.requires_grad else loc.detach().clone()


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
    """``shard_fn`` objects of ``shard_optimizer`` (reference: api.py ShardingStage1/2/3): partition the optimizer work
    for parameters that are REPLICATED along mesh dim ``sharding_mesh_dim``.  Here all three levels partition the
    optimizer states and the update (each rank along the dim owns every n-th such parameter, updates it and broadcasts
    the result); gradients and parameters themselves stay whole on every rank, which is stage-1 memory behaviour."""
    level = "os"

    def __init__(self, sharding_mesh_dim=None, mesh=None):
        if isinstance(sharding_mesh_dim, ProcessMesh):                   # the older (mesh, dim) argument order
            sharding_mesh_dim, mesh = mesh, sharding_mesh_dim
        self.mesh, self.sharding_mesh_dim = mesh, sharding_mesh_dim

    def axis(self, mesh):
        d = self.sharding_mesh_dim
        if d is None:
            return 0
        return mesh.dim_names.index(d) if isinstance(d, str) else int(d)


class ShardingStage1(_ShardingStage):
    level = "os"


class ShardingStage2(_ShardingStage):
    level = "os_g"


class ShardingStage3(_ShardingStage):
    level = "p_g_os"


def _sharded_axes(p):
    return tuple(ax for ax, pl in enumerate(p.placements) if isinstance(pl, Shard)) if is_dist(p) else ()


class _DistGlobalNormClip:
    """ClipGradByGlobalNorm over distributed parameters: the squared norm of a sharded gradient is summed over the mesh
    axes it is sharded on; replicated gradients count once (reference: auto_parallel clip handling in
    python/paddle/nn/clip.py ``_dygraph_clip`` for dist tensors)."""

    def __init__(self, inner):
        self.inner = inner
        self.clip_norm = inner.clip_norm

    @torch.no_grad()
    def __call__(self, params_grads):
        from . import dtensor as D
        buckets = {}
        mesh = None
        for p, g in params_grads:
            if g is None or not getattr(p, "need_clip", True):
                continue
            mesh = mesh or (p.process_mesh if is_dist(p) else None)
            key = _sharded_axes(p)
            sq = D.local_of(g).float().pow(2).sum()
            buckets[key] = buckets.get(key, 0) + sq
        if not buckets:
            return params_grads
        total = None
        for axes, sq in buckets.items():
            sq = sq.reshape(1).clone()
            for ax in axes:
                if mesh.shape[ax] > 1:
                    comm.all_reduce(sq, C.ReduceOp.SUM, mesh.axis_group(ax))
            total = sq if total is None else total + sq
        coef = self.inner.clip_coef(total)
        for p, g in params_grads:
            if g is not None and getattr(p, "need_clip", True):
                D.local_of(g).mul_(coef.to(g.dtype))
        return params_grads

    _dygraph_clip = __call__


class _ShardOptimizer:
    """``shard_optimizer`` result: the wrapped optimizer sees local shards (its accumulators are created with the local
    shapes, so they are sharded exactly like their parameters); this wrapper adds what needs the mesh - the global-norm
    clip over sharded gradients and, with a ShardingStage ``shard_fn``, the partition of the update."""

    def __init__(self, optimizer, shard_fn=None, gradient_accumulation_steps=1):
        self.__dict__["_inner"] = optimizer
        self.__dict__["_shard_fn"] = shard_fn
        self.__dict__["_gradient_accumulation_steps"] = gradient_accumulation_steps
        self.__dict__["_acc"] = 0
        from paddle_b200.nn.clip import ClipGradByGlobalNorm
        clip = getattr(optimizer, "_grad_clip", None)
        if isinstance(clip, ClipGradByGlobalNorm):
            optimizer._grad_clip = _DistGlobalNormClip(clip)

    def __getattr__(self, k):
        return getattr(self._inner, k)

    def __setattr__(self, k, v):
        setattr(self._inner, k, v)

    def _params(self):
        return [p for p in self._inner._parameter_list if isinstance(p, torch.Tensor)]

    @torch.no_grad()
    def step(self):
        from . import dtensor as D
        self.__dict__["_acc"] += 1
        if self._acc % max(int(self._gradient_accumulation_steps), 1) != 0:
            return                                                      # keep accumulating into .grad
        stage = self._shard_fn if isinstance(self._shard_fn, _ShardingStage) else None
        if stage is None:
            return self._inner.step()
        params = [p for p in self._params() if is_dist(p)]
        mesh = stage.mesh or (params[0].process_mesh if params else None)
        if mesh is None:
            return self._inner.step()
        ax = stage.axis(mesh)
        n = mesh.shape[ax]
        if n == 1:
            return self._inner.step()
        me = mesh.coord_of(C.get_rank())[ax]
        group = mesh.axis_group(ax)
        clip = getattr(self._inner, "_grad_clip", None)
        if clip is not None:                                            # the norm needs every gradient: clip first
            clip([(p, p.grad) for p in self._params() if p.grad is not None])
        owned = {}
        k = 0
        for p in params:
            if isinstance(p.placements[ax], Replicate):
                owned[id(p)] = k % n
                k += 1
        hidden = []
        for p in params:
            o = owned.get(id(p))
            if o is not None and o != me and p.grad is not None:
                hidden.append((p, p.grad))
                p.grad = None
        self._inner._grad_clip = None
        try:
            self._inner.step()
        finally:
            self._inner._grad_clip = clip
            for p, g in hidden:
                p.grad = g
        for p in params:
            o = owned.get(id(p))
            if o is not None:
                comm.broadcast(D.local_of(p).data, group.ranks[o], group)

    def clear_grad(self, *a, **kw):
        if self._acc % max(int(self._gradient_accumulation_steps), 1) == 0:
            return self._inner.clear_grad(*a, **kw)

    def state_dict(self):
        return self._inner.state_dict()

    def set_state_dict(self, sd):
        return self._inner.set_state_dict(sd)


def shard_optimizer(optimizer, shard_fn=None, gradient_accumulation_steps=1):
    """Reference: api.py shard_optimizer (:1034-1500)."""
    return _ShardOptimizer(optimizer, shard_fn, gradient_accumulation_steps)


class _ShardDataLoader:
    """``shard_dataloader`` result: every tensor of a batch becomes a DistTensor sharded along its batch dim over mesh
    dim ``shard_dims`` (reference: api.py ShardDataloader :2560-2846).  ``is_dataset_splitted``: the wrapped loader
    already yields this rank's part; otherwise it yields the global batch and the local part is cut out."""

    def __init__(self, loader, meshes, input_keys=None, shard_dims=None, is_dataset_splitted=False):
        self._loader = loader
        self._meshes = meshes if isinstance(meshes, (list, tuple)) else [meshes]
        self._keys = input_keys
        self._dims = shard_dims if isinstance(shard_dims, (list, tuple)) else [shard_dims] * len(self._meshes)
        self._split = is_dataset_splitted

    def __len__(self):
        return len(self._loader)

    def _one(self, t, mesh, dim):
        if not isinstance(t, torch.Tensor):
            return t
        pl = [Replicate() for _ in range(mesh.ndim)]
        if dim is not None:
            ax = mesh.dim_names.index(dim) if isinstance(dim, str) else int(dim)
            pl[ax] = Shard(0)
        return dtensor_from_local(t, mesh, pl) if self._split else shard_tensor(t, mesh, pl)

    def __iter__(self):
        for batch in self._loader:
            if isinstance(batch, dict):
                keys = self._keys or list(batch)
                yield {k: (self._one(v, self._meshes[min(i, len(self._meshes) - 1)], self._dims[min(i, len(self._dims) - 1)])
                           if k in keys else v) for i, (k, v) in enumerate(batch.items())}
            elif isinstance(batch, (list, tuple)):
                yield type(batch)(self._one(v, self._meshes[min(i, len(self._meshes) - 1)], self._dims[min(i, len(self._dims) - 1)])
                                  for i, v in enumerate(batch))
            else:
                yield self._one(batch, self._meshes[0], self._dims[0])


def shard_dataloader(dataloader, meshes, input_keys=None, shard_dims=None, is_dataset_splitted=False):
    return _ShardDataLoader(dataloader, meshes, input_keys, shard_dims, is_dataset_splitted)


def shard_scaler(scaler):
    """A GradScaler whose found-inf flag is agreed on by every rank (reference: api.py shard_scaler :1562-1640): with
    sharded gradients an overflow is visible only on the rank that holds the offending shard."""
    def sync(optimizer, _scaler=scaler):
        if C.get_world_size() > 1 and _scaler._found_inf is not None:
            comm.all_reduce(_scaler._found_inf, C.ReduceOp.MAX)
    scaler._sync_found_inf = sync
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


from .random import parallel_manual_seed  # noqa: E402,F401


class DistModel:
    """``to_static`` result (reference: api.py DistModel :1900-2400).  There is no static graph here: the object drives
    the same dygraph layer with train / eval / predict modes and applies the ``Strategy`` itself:
      sharding        -> ``shard_optimizer`` with ShardingStage{1,2,3} over the mesh dim named "dp" (or the first one)
      gradient_merge  -> the update runs every ``k_steps`` calls (``avg``: the accumulated gradient is the mean)
      amp             -> forward and loss under ``paddle.amp.auto_cast`` (+ a dynamic loss scaler for float16)
    """

    def __init__(self, layer, loader=None, loss=None, optimizer=None, strategy=None):
        self._layer, self._loss, self._strategy = layer, loss, strategy or Strategy()
        self._loader = loader
        self._mode = "train" if optimizer is not None and loss is not None else ("eval" if loss is not None else "predict")
        st = self._strategy
        self._k = max(int(st.gradient_merge.k_steps), 1) if st.gradient_merge.enable else 1
        self._avg = bool(st.gradient_merge.avg)
        self._calls = 0
        self._scaler = None
        if optimizer is not None and st.sharding.enable and not isinstance(optimizer, _ShardOptimizer):
            mesh = next((p.process_mesh for p in layer.parameters() if is_dist(p)), get_mesh())
            stage = {1: ShardingStage1, 2: ShardingStage2, 3: ShardingStage3}[int(st.sharding.stage)]
            dim = "dp" if mesh is not None and "dp" in mesh.dim_names else (mesh.dim_names[0] if mesh is not None else None)
            optimizer = shard_optimizer(optimizer, stage(dim, mesh))
        self._opt = optimizer
        if st.amp.enable and str(st.amp.dtype) == "float16" and optimizer is not None:
            from paddle_b200.amp import GradScaler
            self._scaler = shard_scaler(GradScaler(init_loss_scaling=float(st.amp.init_loss_scaling)))

    def train(self):
        self._mode = "train"
        self._layer.train()

    def eval(self):
        self._mode = "eval"
        self._layer.eval()

    def predict(self):
        self._mode = "predict"
        self._layer.eval()

    def _forward(self, inputs, label=None):
        st = self._strategy
        if st.amp.enable:
            from paddle_b200.amp import auto_cast
            with auto_cast(True, level=str(st.amp.level).upper(), dtype=str(st.amp.dtype)):
                out = self._layer(*inputs)
                return out if label is None else self._loss(out, label)
        out = self._layer(*inputs)
        return out if label is None else self._loss(out, label)

    def __call__(self, *args):
        if self._mode == "predict":
            with torch.no_grad():
                return self._forward(args)
        *inputs, label = args
        if self._mode == "eval":
            with torch.no_grad():
                return self._forward(inputs, label)
        loss = self._forward(inputs, label)
        scaled = loss / self._k if (self._k > 1 and self._avg) else loss
        if self._scaler is not None:
            self._scaler.scale(scaled).backward()
        else:
            scaled.backward()
        self._calls += 1
        if self._calls % self._k == 0:
            if self._scaler is not None:
                self._scaler.step(self._opt)
                self._scaler.update()
            else:
                self._opt.step()
            self._opt.clear_grad()
        return loss.detach()

    def state_dict(self, mode="all"):
        sd = dict(self._layer.state_dict()) if mode in ("all", "param") else {}
        if mode in ("all", "opt") and self._opt is not None:
            sd.update(self._opt.state_dict())
        return sd

    def set_state_dict(self, sd):
        own = self._layer.state_dict()
        self._layer.set_state_dict({k: v for k, v in sd.items() if k in own})
        if self._opt is not None:
            rest = {k: v for k, v in sd.items() if k not in own}
            if rest:
                self._opt.set_state_dict(rest)

    def dist_main_program(self, mode=None):
        """There is no program object: the dygraph layer is the program."""
        return self._layer

    def dist_startup_program(self, mode=None):
        """Parameters are initialised when the layer is built; there is nothing to run at start-up."""
        return None

    def serial_main_program(self, mode=None):
        return self._layer

    def serial_startup_program(self, mode=None):
        return None


def to_static(layer, loader=None, loss=None, optimizer=None, strategy=None):
    return DistModel(layer, loader, loss, optimizer, strategy)
