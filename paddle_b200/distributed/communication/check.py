"""Static and dynamic argument checks of the collectives (reference: paddle/phi/core/distributed/check/
static_check.cc CommStaticCheck::{CheckShape, SameShape, ScatterLikeShape, GatherLikeShape} and
nccl_dynamic_check.cc NCCLDynamicCheck::CheckShape / CheckDataType).

Static checks run on every call (they are a few integer compares): tensors must be on the communicator's device kind,
dtypes must agree, and element counts must satisfy the collective's size relation.  The dynamic check
(``FLAGS_enable_nccl_dynamic_check``) broadcasts the root's dtype id and shape and compares them on every rank, which
turns a silent hang or corruption from mismatched arguments into an error that names the rank."""
from __future__ import annotations

import torch

from paddle_b200.framework.flags import flag


class CommCheckError(ValueError):
    pass


def _place_ok(t, backend):
    if backend == "nccl":
        return t.is_cuda
    if backend == "gloo":
        return not t.is_cuda
    return True          # no communicator yet (single process): nothing to violate


def check_place(t, backend, what="tensor"):
    if not _place_ok(t, backend):
        raise CommCheckError(f"{what} lives on {t.device} but the {backend} communicator needs "
                             f"{'a CUDA' if backend == 'nccl' else 'a CPU'} tensor")


def same_shape(out, inp, backend, op):
    """all_reduce / broadcast / reduce style: out and in hold the same number of elements of one dtype."""
    check_place(out, backend, f"{op} output")
    check_place(inp, backend, f"{op} input")
    if out.dtype != inp.dtype:
        raise CommCheckError(f"{op}: input dtype {inp.dtype} != output dtype {out.dtype}")
    if out.numel() != inp.numel():
        raise CommCheckError(f"{op}: input has {inp.numel()} elements, output {out.numel()}")


def scatter_like(out, inp, nranks, backend, op):
    """reduce_scatter / scatter: in = nranks x out."""
    check_place(out, backend, f"{op} output")
    check_place(inp, backend, f"{op} input")
    if out.dtype != inp.dtype:
        raise CommCheckError(f"{op}: input dtype {inp.dtype} != output dtype {out.dtype}")
    if inp.numel() != out.numel() * nranks:
        raise CommCheckError(f"{op}: input must hold nranks x output elements ({nranks} x {out.numel()} = "
                             f"{out.numel() * nranks}), got {inp.numel()}")


def gather_like(out, inp, nranks, backend, op):
    """all_gather / gather: out = nranks x in."""
    check_place(out, backend, f"{op} output")
    check_place(inp, backend, f"{op} input")
    if out.dtype != inp.dtype:
        raise CommCheckError(f"{op}: input dtype {inp.dtype} != output dtype {out.dtype}")
    if out.numel() != inp.numel() * nranks:
        raise CommCheckError(f"{op}: output must hold nranks x input elements ({nranks} x {inp.numel()} = "
                             f"{inp.numel() * nranks}), got {out.numel()}")


_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64, torch.int32, torch.int16, torch.int8,
           torch.uint8, torch.bool, torch.complex64, torch.complex128]


def dynamic_check(tensor, root, group, op):
    """Every rank compares its dtype / shape with the root's (one small broadcast).  No-op unless the flag is on."""
    if not flag("enable_nccl_dynamic_check"):
        return
    import torch.distributed as tdist
    dev = tensor.device if tensor.is_cuda else torch.device("cpu")
    meta = [_DTYPES.index(tensor.dtype) if tensor.dtype in _DTYPES else -1, tensor.dim()] + list(tensor.shape)
    buf = torch.tensor(meta + [0] * (16 - len(meta)), dtype=torch.int64, device=dev)
    mine = buf.clone()
    tdist.broadcast(buf, src=group.ranks[root] if root < len(group.ranks) else root, group=group.process_group)
    if not torch.equal(buf.cpu(), mine.cpu()):
        r = buf.tolist()
        raise CommCheckError(f"{op}: rank {group.rank} passes dtype {tensor.dtype} shape {tuple(tensor.shape)} but root "
                             f"{root} passes dtype {_DTYPES[r[0]] if 0 <= r[0] < len(_DTYPES) else '?'} shape {tuple(r[2:2 + r[1]])}")
