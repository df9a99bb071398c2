"""SPMD inference rules (N13; reference: paddle/phi/infermeta/spmd_rules/*.cc - matmul.cc, elementwise.cc,
reduction.cc, embedding.cc, layer_norm.cc, softmax.cc, transpose.cc, reshape.cc, concat.cc, split.cc,
cross_entropy_with_softmax.cc, flash_attention.cc, replicated.cc, and the notation helpers in utils.cc).

Given the sharding of an op's inputs, a rule answers two questions: how must the inputs be (re)sharded so that the op
can run shard-locally, and how are the outputs sharded (and which of them are partial sums).  Shardings are written as
``dims_mapping``: one entry per tensor dim, the mesh axis that dim is split over or -1.

One engine serves almost every rule: the op is written in einsum-like notation (``"mk,kn->mn"``), the mesh axis of
every letter is merged over all inputs (the same letter must be split the same way everywhere; a conflict falls back
to replicated, and a mesh axis may shard only one letter), inputs are re-expressed with the merged mapping, outputs
read their letters' axes, and a letter that is sharded but absent from an output (a contracted dim) makes that output
partial over the letter's mesh axis.
"""
from __future__ import annotations

import string
from dataclasses import dataclass, field


@dataclass
class DistTensorSpec:
    shape: tuple
    dims_mapping: list
    partial_dims: set = field(default_factory=set)      # mesh axes over which the tensor is a partial sum

    def __post_init__(self):
        self.shape = tuple(self.shape)
        self.dims_mapping = list(self.dims_mapping)
        self.partial_dims = set(self.partial_dims)
        assert len(self.shape) == len(self.dims_mapping), (self.shape, self.dims_mapping)

    def replicated(self):
        return DistTensorSpec(self.shape, [-1] * len(self.shape))

    def placements(self, mesh_ndim):
        """dims_mapping + partial_dims -> [Placement] per mesh axis (the dygraph API's form)."""
        from . import Partial, Replicate, Shard
        out = [Replicate() for _ in range(mesh_ndim)]
        for d, ax in enumerate(self.dims_mapping):
            if ax >= 0:
                out[ax] = Shard(d)
        for ax in self.partial_dims:
            out[ax] = Partial()
        return out

    @staticmethod
    def from_placements(shape, placements):
        dm, part = [-1] * len(shape), set()
        for ax, p in enumerate(placements):
            if p.is_shard():
                dm[p.get_dim()] = ax
            elif p.is_partial():
                part.add(ax)
        return DistTensorSpec(shape, dm, part)


# --------------------------------------------------------------------------------------------- einsum engine
def _merge(notations, specs):
    """letter -> mesh axis, merged over the inputs."""
    axis_of = {}
    for note, spec in zip(notations, specs):
        for letter, ax, size in zip(note, spec.dims_mapping, spec.shape):
            if letter == "1" or size == 1:          # broadcast dims never carry a sharding
                continue
            if ax < 0:
                axis_of.setdefault(letter, -1)
            elif axis_of.get(letter, -1) < 0:
                axis_of[letter] = ax
            elif axis_of[letter] != ax:
                axis_of[letter] = -1               # two inputs disagree: replicate that dim
    # a mesh axis can shard one letter only: the first (in input order) keeps it
    seen = {}
    for note in notations:
        for letter in note:
            ax = axis_of.get(letter, -1)
            if ax >= 0:
                if ax in seen and seen[ax] != letter:
                    axis_of[letter] = -1
                else:
                    seen[ax] = letter
    return axis_of


def einsum_infer(notation, specs, out_shapes=None):
    lhs, rhs = notation.replace(" ", "").split("->")
    in_notes, out_notes = lhs.split(","), rhs.split(",")          # "ab->" has one output, a scalar
    assert len(in_notes) == len(specs), (notation, len(specs))
    axis_of = _merge(in_notes, specs)
    new_inputs = []
    for note, spec in zip(in_notes, specs):
        dm = [(-1 if (l == "1" or s == 1) else axis_of.get(l, -1)) for l, s in zip(note, spec.shape)]
        new_inputs.append(DistTensorSpec(spec.shape, dm))
    size_of = {l: s for note, spec in zip(in_notes, specs) for l, s in zip(note, spec.shape) if s != 1}
    outs = []
    for i, note in enumerate(out_notes):
        shape = out_shapes[i] if out_shapes else tuple(size_of.get(l, 1) for l in note)
        dm = [axis_of.get(l, -1) for l in note]
        partial = {ax for l, ax in axis_of.items() if ax >= 0 and l not in note and any(l in n for n in in_notes)}
        outs.append(DistTensorSpec(shape, dm, partial))
    return new_inputs, outs


def _letters(n, skip=""):
    pool = [c for c in string.ascii_lowercase if c not in skip]
    return "".join(pool[:n])


# --------------------------------------------------------------------------------------------- rules
def matmul(x, y, trans_x=False, trans_y=False):
    """Batched matmul with broadcasting of the batch dims (reference: matmul.cc)."""
    nx, ny = len(x.shape), len(y.shape)
    batch = max(nx, ny) - 2
    bl = _letters(max(batch, 0), skip="mnk")
    xn = bl[batch - (nx - 2):] + ("km" if trans_x else "mk") if nx >= 2 else "k"
    yn = bl[batch - (ny - 2):] + ("nk" if trans_y else "kn") if ny >= 2 else "k"
    out = bl + ("m" if nx >= 2 else "") + ("n" if ny >= 2 else "")
    return einsum_infer(f"{xn},{yn}->{out}", [x, y])


def elementwise(*specs):
    """Broadcasting elementwise op of any arity (reference: elementwise.cc)."""
    nd = max(len(s.shape) for s in specs)
    letters = _letters(nd)
    notes = []
    for s in specs:
        off = nd - len(s.shape)
        notes.append("".join("1" if size == 1 else letters[off + i] for i, size in enumerate(s.shape)))
    shape = tuple(max(s.shape[i - (nd - len(s.shape))] if i >= nd - len(s.shape) else 1 for s in specs) for i in range(nd))
    return einsum_infer(",".join(notes) + "->" + letters, list(specs), [shape])


def reduction(x, axis=None, keepdim=False):
    """sum / mean / max ...: a reduced dim that is sharded leaves the output partial (reference: reduction.cc)."""
    nd = len(x.shape)
    axes = list(range(nd)) if axis is None else [a % nd for a in (axis if isinstance(axis, (list, tuple)) else [axis])]
    letters = _letters(nd)
    out = "".join(("1" if keepdim else "") if i in axes else letters[i] for i in range(nd))
    shape = tuple(1 if i in axes else x.shape[i] for i in range(nd) if keepdim or i not in axes)
    ins, outs = einsum_infer(f"{letters}->{out.replace('1', '')}", [x], None)
    o = outs[0]
    if keepdim:
        dm, it = [], iter(o.dims_mapping)
        for i in range(nd):
            dm.append(-1 if i in axes else next(it))
        o = DistTensorSpec(shape, dm, o.partial_dims)
    return ins, [o]


def embedding(ids, weight):
    """Row-sharded (vocab-parallel) tables give partial outputs; column-sharded ones shard the hidden dim (embedding.cc)."""
    idl = _letters(len(ids.shape), skip="vh")
    return einsum_infer(f"{idl},vh->{idl}h", [ids, weight])


def layer_norm(x, scale=None, bias=None, begin_norm_axis=-1):
    """Normalised dims must be whole on every rank; leading dims keep their sharding (layer_norm.cc).  Outputs:
    y, mean, variance."""
    nd = len(x.shape)
    b = begin_norm_axis % nd
    dm = [ax if i < b else -1 for i, ax in enumerate(x.dims_mapping)]
    xin = DistTensorSpec(x.shape, dm)
    ins = [xin] + [s.replicated() for s in (scale, bias) if s is not None]
    stat = DistTensorSpec(x.shape[:b], dm[:b])
    return ins, [DistTensorSpec(x.shape, dm), stat, stat]


rms_norm = layer_norm


def softmax(x, axis=-1):
    nd = len(x.shape)
    a = axis % nd
    dm = [(-1 if i == a else ax) for i, ax in enumerate(x.dims_mapping)]
    return [DistTensorSpec(x.shape, dm)], [DistTensorSpec(x.shape, dm)]


def cross_entropy_with_softmax(logits, label, axis=-1):
    """Class-dim sharding is kept (vocab-parallel CE computes with partial statistics); the loss is replicated over that
    mesh axis after the op's own reduction (cross_entropy_with_softmax.cc).  Outputs: softmax, loss."""
    nd = len(logits.shape)
    a = axis % nd
    dm = list(logits.dims_mapping)
    lab_dm = [(-1 if i == a else dm[i]) for i in range(nd)][: len(label.shape)]
    loss_shape = tuple(1 if i == a else s for i, s in enumerate(logits.shape))
    loss_dm = [(-1 if i == a else dm[i]) for i in range(nd)]
    return ([DistTensorSpec(logits.shape, dm), DistTensorSpec(label.shape, lab_dm)],
            [DistTensorSpec(logits.shape, dm), DistTensorSpec(loss_shape, loss_dm)])


def transpose(x, perm):
    dm = [x.dims_mapping[p] for p in perm]
    return [DistTensorSpec(x.shape, x.dims_mapping)], [DistTensorSpec(tuple(x.shape[p] for p in perm), dm)]


def reshape(x, shape):
    """A sharded input dim survives when it is the LEADING factor of the output dim(s) it maps to (reshape.cc's
    dim-transformation: flatten keeps the first dim's sharding, split keeps it on the first piece); otherwise that dim
    is replicated first."""
    shape = list(shape)
    numel = 1
    for s in x.shape:
        numel *= s
    if -1 in shape:
        known = 1
        for s in shape:
            if s != -1:
                known *= s
        shape[shape.index(-1)] = numel // max(known, 1)
    in_dm = list(x.dims_mapping)
    out_dm = [-1] * len(shape)
    i = j = 0
    while i < len(x.shape) and j < len(shape):
        a, b, i0, j0 = x.shape[i], shape[j], i, j
        while a != b:
            if a < b:
                i += 1
                a *= x.shape[i]
            else:
                j += 1
                b *= shape[j]
        # input dims i0..i map onto output dims j0..j: only the leading input dim may stay sharded, and its sharding
        # moves to the leading output dim of the group (flatten keeps the first dim's split, a split keeps it on the
        # first piece); every other input dim of the group has to be whole
        out_dm[j0] = in_dm[i0]
        for k in range(i0 + 1, i + 1):
            in_dm[k] = -1
        i += 1
        j += 1
    return [DistTensorSpec(x.shape, in_dm)], [DistTensorSpec(tuple(shape), out_dm)]


def concat(specs, axis=0):
    nd = len(specs[0].shape)
    a = axis % nd
    letters = _letters(nd)
    notes = ["".join("1" if i == a else letters[i] for i in range(nd)) for _ in specs]
    fake = [DistTensorSpec(tuple(1 if i == a else s.shape[i] for i in range(nd)),
                           [(-1 if i == a else s.dims_mapping[i]) for i in range(nd)]) for s in specs]
    ins, outs = einsum_infer(",".join(notes) + "->" + "".join("1" if i == a else letters[i] for i in range(nd)), fake)
    dm = ins[0].dims_mapping if ins else [-1] * nd
    merged = [(-1 if i == a else _pick(ins, i)) for i in range(nd)]
    shape = tuple(sum(s.shape[a] for s in specs) if i == a else specs[0].shape[i] for i in range(nd))
    return [DistTensorSpec(s.shape, merged) for s in specs], [DistTensorSpec(shape, merged)]


def _pick(ins, i):
    for s in ins:
        if s.dims_mapping[i] >= 0:
            return s.dims_mapping[i]
    return -1


def split(x, num_or_sections, axis=0):
    nd = len(x.shape)
    a = axis % nd
    dm = [(-1 if i == a else ax) for i, ax in enumerate(x.dims_mapping)]
    secs = [x.shape[a] // num_or_sections] * num_or_sections if isinstance(num_or_sections, int) else list(num_or_sections)
    outs = [DistTensorSpec(tuple(s if i == a else x.shape[i] for i in range(nd)), dm) for s in secs]
    return [DistTensorSpec(x.shape, dm)], outs


def flash_attention(q, k, v):
    """[b, s, h, d]: batch and heads may be sharded (data / tensor parallel); the sequence dim of k, v and the head dim
    d must be whole (flash_attention.cc).  Outputs: out, softmax_lse [b, h, s]."""
    ins, outs = einsum_infer("bqhd,bkhd,bkhd->bqhd,bhq", [q, k, v])
    fix = []
    for s, note in zip(ins, ("bqhd", "bkhd", "bkhd")):
        dm = [(-1 if (l in "kd") else ax) for l, ax in zip(note, s.dims_mapping)]
        fix.append(DistTensorSpec(s.shape, dm))
    out = DistTensorSpec(q.shape, [fix[0].dims_mapping[0], fix[0].dims_mapping[1], fix[0].dims_mapping[2], -1])
    lse = DistTensorSpec((q.shape[0], q.shape[2], q.shape[1]), [out.dims_mapping[0], out.dims_mapping[2], out.dims_mapping[1]])
    return fix, [out, lse]


def fused_rope(q, k=None, v=None):
    """Position-wise: every input keeps its sharding except the rotary (last) dim (fused_rope.cc)."""
    ins = [DistTensorSpec(s.shape, s.dims_mapping[:-1] + [-1]) for s in (q, k, v) if s is not None]
    return ins, [DistTensorSpec(s.shape, s.dims_mapping) for s in ins]


def qkv_split_rope(qkv, num_q_heads=1, num_kv_heads=1):
    """Packed projection [b, s, (hq + 2 hk) d] -> q [b, s, hq, d], k, v [b, s, hk, d]: the packed dim mixes q, k and v, so
    it must be whole; batch and sequence shardings pass through (the rotary tables are indexed by position ids)."""
    d = qkv.shape[-1] // (num_q_heads + 2 * num_kv_heads)
    dm = list(qkv.dims_mapping[:-1]) + [-1]
    lead = tuple(qkv.shape[:-1])
    outs = [DistTensorSpec(lead + (h, d), dm[:-1] + [-1, -1]) for h in (num_q_heads, num_kv_heads, num_kv_heads)]
    return [DistTensorSpec(qkv.shape, dm)], outs


def swiglu(x, y=None):
    if y is None:
        dm = x.dims_mapping[:-1] + [-1]          # the packed gate|up dim must be whole to be split in halves
        return [DistTensorSpec(x.shape, dm)], [DistTensorSpec(x.shape[:-1] + (x.shape[-1] // 2,), dm)]
    return elementwise(x, y)


def replicated(*specs, n_outputs=1, out_shapes=None):
    """Fallback for ops without a rule: everything replicated (replicated.cc)."""
    ins = [s.replicated() for s in specs]
    shapes = out_shapes or [specs[0].shape] * n_outputs
    return ins, [DistTensorSpec(s, [-1] * len(s)) for s in shapes]


_RULES = {"matmul": matmul, "elementwise": elementwise, "add": elementwise, "subtract": elementwise, "multiply": elementwise,
          "divide": elementwise, "sum": reduction, "mean": reduction, "max": reduction, "min": reduction,
          "reduction": reduction, "embedding": embedding, "layer_norm": layer_norm, "rms_norm": rms_norm,
          "softmax": softmax, "cross_entropy_with_softmax": cross_entropy_with_softmax, "transpose": transpose,
          "reshape": reshape, "concat": concat, "split": split, "flash_attention": flash_attention,
          "fused_rotary_position_embedding": fused_rope, "swiglu": swiglu, "qkv_split_rope": qkv_split_rope,
          "default": replicated}


def get_spmd_rule(op_name):
    """Rule lookup by op name; unknown ops get the replicated fallback (reference: SpmdRuleFactory)."""
    return _RULES.get(op_name, replicated)


def infer_forward(op_name, *specs, **attrs):
    return get_spmd_rule(op_name)(*specs, **attrs)
