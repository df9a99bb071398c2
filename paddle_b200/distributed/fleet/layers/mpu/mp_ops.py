"""Tensor-parallel glue ops (reference: python/paddle/distributed/fleet/layers/mpu/mp_ops.py:27-925 and the ops
c_identity / c_concat / c_split / mp_allreduce_sum / c_embedding / c_softmax_with_cross_entropy under
paddle/fluid/operators/collective/).

Each op is an autograd Function whose forward/backward pair is the conjugate collective.  The vocab-parallel
cross-entropy moves exactly 3 floats per row between ranks (max, sum-exp, target logit): one fused statistics pass
(K8, csrc/kernels/loss.cu), an exchange - in-kernel over NVLink symmetric memory when available, else one
all-gather - and one backward pass, replacing the reference's three N x 1 all-reduces and >= 7 launches
(c_softmax_with_cross_entropy_op.cu:133-340)."""
from __future__ import annotations

import torch

from paddle_b200 import ops
from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp


def _group(group):
    if group is not None:
        return group
    from paddle_b200.distributed import fleet
    hcg = fleet.get_hybrid_communicate_group()
    return hcg.get_model_parallel_group()


def _nranks(group):
    return group.nranks if group is not None else 1


class _CIdentity(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group, skip):
        ctx.group, ctx.skip = group, skip
        return x.view_as(x)

    @staticmethod
    def backward(ctx, dy):
        # ``skip_c_identity_dynamic`` only drops the forward identity op (a no-op view here); the gradient of the
        # replicated input is ALWAYS summed over the group (reference mp_ops.py c_identity_eager.backward)
        if _nranks(ctx.group) == 1:
            return dy, None, None
        dy = dy.contiguous().clone() if not dy.is_contiguous() else dy.clone()
        comm.all_reduce(dy, ReduceOp.SUM, ctx.group)
        return dy, None, None


def _c_identity(tensor, group=None, skip_c_identity_dynamic=False):
    if group is not None and not group.is_member():
        return tensor
    return _CIdentity.apply(tensor, group, skip_c_identity_dynamic)


class _MpAllReduce(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group, skip):
        ctx.group = group
        # the forward all-reduce is unconditional; ``skip_c_identity_dynamic`` only drops the identity op of the
        # backward pass, which is already a pass-through here (reference mp_ops.py mp_allreduce_eager)
        if _nranks(group) == 1:
            return x.view_as(x)
        out = x.contiguous().clone()
        comm.all_reduce(out, ReduceOp.SUM, group)
        return out

    @staticmethod
    def backward(ctx, dy):
        return dy, None, None


def _mp_allreduce(tensor, op=ReduceOp.SUM, group=None, use_calc_stream=True, use_model_parallel=True,
                  skip_c_identity_dynamic=False):
    if group is not None and not group.is_member():
        return tensor
    assert op == ReduceOp.SUM, f"Unknown parameter: {op}."
    return _MpAllReduce.apply(tensor, group, skip_c_identity_dynamic)


class _CConcat(torch.autograd.Function):
    """all-gather on the last dim (forward) / keep own slice (backward)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        n = _nranks(group)
        if n == 1:
            return x.view_as(x)
        parts = []
        comm.all_gather(parts, x.contiguous(), group)
        return torch.cat(parts, dim=-1)

    @staticmethod
    def backward(ctx, dy):
        g = ctx.group
        n = _nranks(g)
        if n == 1:
            return dy, None
        return dy.chunk(n, dim=-1)[g.rank].contiguous(), None


def _c_concat(tensor, group=None):
    if group is not None and not group.is_member():
        return tensor
    return _CConcat.apply(tensor, group)


class _CSplit(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        n = _nranks(group)
        if n == 1:
            return x.view_as(x)
        return x.chunk(n, dim=-1)[group.rank].contiguous()

    @staticmethod
    def backward(ctx, dy):
        g = ctx.group
        if _nranks(g) == 1:
            return dy, None
        parts = []
        comm.all_gather(parts, dy.contiguous(), g)
        return torch.cat(parts, dim=-1), None


def _c_split(tensor, group=None):
    if group is not None and not group.is_member():
        return tensor
    return _CSplit.apply(tensor, group)


def _c_lookup_table(table, index, start_index=0, vocab_size=-1, name=None):
    """Masked row gather: rows outside [start, start + table.rows) give zeros (c_embedding_kernel.cu:31-62)."""
    rows = table.shape[0]
    local = index.long() - start_index
    inside = (local >= 0) & (local < rows)
    out = torch.nn.functional.embedding(local.clamp(0, rows - 1), table)
    return out * inside.unsqueeze(-1).to(out.dtype)


class _ParallelCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, label, group, ignore_index):
        n = _nranks(group)
        rank = group.rank if group is not None else 0
        V = logits.shape[-1]
        lab = label.squeeze(-1) if (label.dim() == logits.dim() and label.shape[-1] == 1) else label
        vocab_start = rank * V
        stats = ops.ce_local_stats(logits, lab, vocab_start, ignore_index)      # [rows, 3] fp32
        if n > 1:
            from paddle_b200.parallel import symm
            if stats.is_cuda and symm.available(group):
                loss, lse = symm.ce_exchange_combine(stats, lab, group, ignore_index)      # in-kernel NVLink exchange
            else:
                loss, lse = ops.ce_combine(symm.all_gather_small(stats, group), lab, ignore_index)
        else:
            loss, lse = ops.ce_combine(stats.unsqueeze(0), lab, ignore_index)
        ctx.save_for_backward(logits, lab, lse)
        ctx.cfg = (vocab_start, ignore_index)
        return loss.reshape(label.shape).to(torch.float32 if logits.dtype != torch.float64 else logits.dtype)

    @staticmethod
    def backward(ctx, dloss):
        logits, lab, lse = ctx.saved_tensors
        vocab_start, ignore_index = ctx.cfg
        return ops.ce_local_backward(logits, lab, lse, dloss, vocab_start, ignore_index), None, None, None


def _c_softmax_with_cross_entropy(logits, label, group=None, return_softmax=False, ignore_index=-100):
    if group is not None and not group.is_member():
        return
    if logits.dim() != label.dim() and logits.dim() - 1 != label.dim():
        raise ValueError(f"Expected input_dims - 1 = label_dims or input_dims == label_dims "
                         f"(got input_dims{logits.dim()}, label_dims{label.dim()})")
    lab = label.unsqueeze(-1) if logits.dim() - 1 == label.dim() else label
    loss = _ParallelCE.apply(logits, lab, group, ignore_index)
    if return_softmax:
        raise NotImplementedError("return_softmax=True: materialising the [N, V/n] softmax defeats the fused path")
    return loss


def _linear(x, weight, bias=None, name=None):
    from paddle_b200.nn import functional as F
    return F.linear(x, weight, bias)


def _set_var_distributed(var):
    if var is not None:
        var.is_distributed = True


def split(x, size, operation, axis=0, num_partitions=1, gather_out=True, weight_attr=None, bias_attr=None, name=None):
    """paddle.distributed.split (reference: mp_ops.py:700-925): build-and-apply a parallel linear / embedding."""
    from .mp_layers import ColumnParallelLinear, RowParallelLinear, VocabParallelEmbedding
    assert operation in ("linear", "embedding"), "The operation for paddle.distributed.split must be one of " \
                                                 "{'linear', 'embedding'}."
    if operation == "embedding":
        assert axis == 0, "We only support to split the weight of embedding along the first axis now."
        layer = VocabParallelEmbedding(size[0], size[1], weight_attr=weight_attr)
        return layer(x)
    if axis == 0:
        layer = RowParallelLinear(size[0], size[1], weight_attr=weight_attr, has_bias=bias_attr is not False,
                                  input_is_parallel=False)
    else:
        layer = ColumnParallelLinear(size[0], size[1], weight_attr=weight_attr, has_bias=bias_attr is not False,
                                     gather_output=gather_out)
    return layer(x)
