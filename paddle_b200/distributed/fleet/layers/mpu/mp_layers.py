"""Tensor-parallel layers (reference: python/paddle/distributed/fleet/layers/mpu/mp_layers.py:47-790).

Weight shapes / attributes follow the reference: Column ``[in, out/n]`` (``split_axis=1``), Row ``[in/n, out]``
(``split_axis=0``, bias replicated and added after the reduction), VocabParallelEmbedding ``[V/n, h]`` with
``vocab_start_index = rank * V/n``; sharded parameters carry ``is_distributed=True`` and are initialised under the
model-parallel RNG stream so shards differ across ranks.

Fused NVLink paths (FLAGS_use_fused_comm, ``paddle_b200.parallel.symm``): Row-parallel forward and Column-parallel
backward use the GEMM -> all-reduce kernel (K4) when the activation fits the symmetric staging buffer; otherwise
the NCCL-composed ops from mp_ops are used (baseline path).
"""
from __future__ import annotations

import torch

from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.nn import functional as F
from paddle_b200.nn import initializer as I
from paddle_b200.nn.layer.layers import Layer

from . import mp_ops
from .random import get_rng_state_tracker


def _hcg():
    from paddle_b200.distributed import fleet
    return fleet.get_hybrid_communicate_group()


def _mp_cfg(key, default=False):
    from paddle_b200.distributed import fleet
    st = fleet.fleet._user_defined_strategy
    if st is None:
        return default
    return st.hybrid_configs["mp_configs"].get(key, default)


def _resolve_group(mp_group):
    if mp_group is not None:
        return mp_group, mp_group.nranks, mp_group.rank
    hcg = _hcg()
    if hcg is None:
        return None, 1, 0
    return hcg.get_model_parallel_group(), hcg.get_model_parallel_world_size(), hcg.get_model_parallel_rank()


def _sharded_init_ctx(is_mp):
    tracker = get_rng_state_tracker()
    if is_mp and "model_parallel_rng" in tracker.states_:
        return tracker.rng_state()
    import contextlib
    return contextlib.nullcontext()


class VocabParallelEmbedding(Layer):
    def __init__(self, num_embeddings, embedding_dim, weight_attr=None, mp_group=None, name=None):
        super().__init__()
        self.model_parallel_group, self.world_size, self.rank = _resolve_group(mp_group)
        self.origin_num_embeddings = num_embeddings
        self.is_mp = self.world_size > 1
        assert num_embeddings % self.world_size == 0, \
            "The length of the vocabulary must be divisible by the parallelism degree of MP"
        per_part = num_embeddings // self.world_size
        self.vocab_start_index = self.rank * per_part
        self._size = [per_part, embedding_dim]
        with _sharded_init_ctx(self.is_mp):
            self.weight = self.create_parameter(self._size, attr=weight_attr, dtype=self._dtype,
                                                default_initializer=I.Normal(0.0, 1.0) if weight_attr is None else None)
        self.weight.is_distributed = self.is_mp
        if self.is_mp:
            self.weight.split_axis = 0

    def forward(self, x):
        if not self.is_mp:
            return F.embedding(x, self.weight)
        if self.weight.is_cuda and self.weight.shape[1] * self.weight.element_size() % 16 == 0:
            from paddle_b200.parallel import symm
            if symm.available(self.model_parallel_group) and hasattr(symm.lib.kernels.require(), "pdb200_nvl_vocab_embed") \
                    and symm.make_symmetric(self.weight, self.model_parallel_group):
                return _VocabEmbedFused.apply(x, self.weight, self.weight._symm, self.vocab_start_index)
        out = mp_ops._c_lookup_table(self.weight, x, start_index=self.vocab_start_index,
                                     vocab_size=self.origin_num_embeddings)
        return mp_ops._mp_allreduce(out, group=self.model_parallel_group)


class _VocabEmbedFused(torch.autograd.Function):
    """K7: forward gathers each row from the owning rank's shard over NVLink (no [tokens, hidden] all-reduce);
    backward needs no communication: activations are replicated across mp, so each rank accumulates its own rows."""

    @staticmethod
    def forward(ctx, ids, w, where, start):
        from paddle_b200.parallel import symm
        ctx.save_for_backward(ids)
        ctx.start, ctx.rows, ctx.wdtype = start, w.shape[0], w.dtype
        return symm.vocab_embed(ids, w, where)

    @staticmethod
    def backward(ctx, dy):
        (ids,) = ctx.saved_tensors
        local = ids.reshape(-1) - ctx.start
        ok = (local >= 0) & (local < ctx.rows)
        dw = torch.zeros(ctx.rows, dy.shape[-1], dtype=torch.float32, device=dy.device)
        dw.index_add_(0, local.clamp(0, ctx.rows - 1), dy.reshape(-1, dy.shape[-1]).float() * ok[:, None])
        return None, dw.to(ctx.wdtype), None, None


class _FusedRowAR(torch.autograd.Function):
    """K4 forward: GEMM whose epilogue pushes partial tiles to the row owner; owners fold and broadcast the sum."""

    @staticmethod
    def forward(ctx, x, w, group):
        from paddle_b200.parallel import symm
        ctx.save_for_backward(x, w)
        y = symm.gemm_ar(x.reshape(-1, x.shape[-1]), w, group)
        return y.reshape(*x.shape[:-1], w.shape[1])

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        dx = torch.matmul(dy2, w.t()).reshape(x.shape)
        dw = torch.matmul(x.reshape(-1, x.shape[-1]).t(), dy2)
        return dx, dw, None


class _FusedColAR(torch.autograd.Function):
    """ColumnParallelLinear with the dX all-reduce fused into the dX GEMM (K4, transposed weight)."""

    @staticmethod
    def forward(ctx, x, w, group):
        ctx.save_for_backward(x, w)
        ctx.group = group
        return torch.matmul(x, w)

    @staticmethod
    def backward(ctx, dy):
        from paddle_b200.parallel import symm
        x, w = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        dx = symm.gemm_ar(dy2, w, ctx.group, trans_b=True).reshape(x.shape)
        dw = torch.matmul(x.reshape(-1, x.shape[-1]).t(), dy2)
        return dx, dw, None


def _fused_ar_ok(x, w, group, n_out):
    """rows divisible by nranks*128, bf16, arena with room for staging + output."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16):
        return False
    from paddle_b200.parallel import symm
    rows = x.numel() // x.shape[-1]
    if rows % (group.nranks * 128) != 0:
        return False
    if not symm.available(group):
        return False
    return symm.fused_gemm_ok(group, rows // group.nranks, x.shape[-1], n_out, x.dtype) and \
        (symm.get_arena(group).free_bytes() > rows * n_out * 2 + (1 << 20) or
         any(k[1] == "ar_part" for k in symm._stage))


class _AsyncColumnLinear(torch.autograd.Function):
    """``mp_async_allreduce``: in backward the dX all-reduce is launched asynchronously and overlaps the dW GEMM
    (reference InnerOverlapLinear, mp_layers.py:188-331)."""

    @staticmethod
    def forward(ctx, x, w, b, group, fuse_grad_add):
        ctx.save_for_backward(x, w)
        ctx.group, ctx.has_b, ctx.fuse = group, b is not None, fuse_grad_add
        ctx.wref = w
        y = torch.matmul(x, w)
        return y + b if b is not None else y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dx = torch.matmul(dy, w.t())
        task = comm.all_reduce(dx, ReduceOp.SUM, ctx.group, sync_op=False)
        x2, dy2 = x.reshape(-1, x.shape[-1]), dy.reshape(-1, dy.shape[-1])
        mg = getattr(ctx.wref, "main_grad", None)
        if ctx.fuse and mg is not None:
            mg.add_(torch.matmul(x2.t(), dy2).to(mg.dtype))
            dw = None
        else:
            dw = torch.matmul(x2.t(), dy2)
        db = dy2.sum(0) if ctx.has_b else None
        task.wait()
        return dx, dw, db, None, None


class ColumnParallelLinear(Layer):
    def __init__(self, in_features, out_features, weight_attr=None, has_bias=None, gather_output=True,
                 fuse_matmul_bias=False, mp_group=None, name=None):
        super().__init__()
        self.model_parallel_group, self.world_size, _ = _resolve_group(mp_group)
        self._name = name
        self.is_mp = self.world_size > 1
        self.gather_output = gather_output
        assert out_features % self.world_size == 0, (
            f"Number of column of the weight for linear ({out_features}) must be divisible by model parallel size "
            f"({self.world_size})")
        self.output_size_per_partition = out_features // self.world_size
        with _sharded_init_ctx(self.is_mp):
            self.weight = self.create_parameter([in_features, self.output_size_per_partition], attr=weight_attr,
                                                dtype=self._dtype)
        self.weight.is_distributed = self.is_mp
        if self.is_mp:
            self.weight.split_axis = 1
        if has_bias:
            self.bias = self.create_parameter([self.output_size_per_partition], attr=None, dtype=self._dtype,
                                              is_bias=True)
            self.bias.is_distributed = self.is_mp
            if self.is_mp:
                self.bias.split_axis = 0
        else:
            self.bias = None
        self.fuse_matmul_bias = fuse_matmul_bias
        self.mp_async_allreduce = _mp_cfg("mp_async_allreduce") and self.is_mp
        self.mp_skip_c_identity = _mp_cfg("mp_skip_c_identity") and self.is_mp
        self.mp_fused_linear_param_grad_add = _mp_cfg("mp_fused_linear_param_grad_add") and self.is_mp

    def forward(self, x):
        if self.mp_async_allreduce:
            out = _AsyncColumnLinear.apply(x, self.weight, self.bias, self.model_parallel_group,
                                           self.mp_fused_linear_param_grad_add)
        else:
            if self.is_mp and not self.mp_skip_c_identity and x.requires_grad and \
                    _fused_ar_ok(x, self.weight, self.model_parallel_group, x.shape[-1]):
                out = _FusedColAR.apply(x, self.weight, self.model_parallel_group)
                out = out + self.bias if self.bias is not None else out
            else:
                inp = mp_ops._c_identity(x, group=self.model_parallel_group,
                                         skip_c_identity_dynamic=self.mp_skip_c_identity) if self.is_mp else x
                out = F.linear(inp, self.weight, self.bias)
        if self.gather_output and self.is_mp:
            out = mp_ops._c_concat(out, group=self.model_parallel_group)
        return out


class MPScale(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, mp_degree):
        return x / mp_degree

    @staticmethod
    def backward(ctx, dout):
        return dout, None


class RowParallelLinear(Layer):
    def __init__(self, in_features, out_features, weight_attr=None, has_bias=True, input_is_parallel=False,
                 fuse_matmul_bias=False, mp_group=None, name=None):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.input_is_parallel = input_is_parallel
        self._name = name
        self.model_parallel_group, self.world_size, self.rank = _resolve_group(mp_group)
        self.is_mp = self.world_size > 1
        self.mp_skip_c_identity = _mp_cfg("mp_skip_c_identity") and self.is_mp
        assert in_features % self.world_size == 0, (
            f"Number of row of the weight for linear ({in_features}) must be divisible by model parallel size "
            f"({self.world_size})")
        self.input_size_per_partition = in_features // self.world_size
        with _sharded_init_ctx(self.is_mp):
            self.weight = self.create_parameter([self.input_size_per_partition, out_features], attr=weight_attr,
                                                dtype=self._dtype)
        self.weight.is_distributed = self.is_mp
        if self.is_mp:
            self.weight.split_axis = 0
        self.bias = self.create_parameter([out_features], attr=None, dtype=self._dtype, is_bias=True) if has_bias else None
        self.fuse_matmul_bias = fuse_matmul_bias

    def forward(self, x):
        if self.input_is_parallel or not self.is_mp:
            inp = x
        else:
            inp = mp_ops._c_split(x, group=self.model_parallel_group)
        if not self.is_mp:
            return F.linear(inp, self.weight, self.bias)
        if self.fuse_matmul_bias and self.bias is not None:
            # bias scaled by 1/mp before the GEMM so the all-reduce restores it (mp_layers.py:708-719)
            out_parallel = F.linear(inp, self.weight, MPScale.apply(self.bias, self.world_size))
            return mp_ops._mp_allreduce(out_parallel, group=self.model_parallel_group,
                                        skip_c_identity_dynamic=self.mp_skip_c_identity)
        if _fused_ar_ok(inp, self.weight, self.model_parallel_group, self.out_features):
            out = _FusedRowAR.apply(inp, self.weight, self.model_parallel_group)
            return out + self.bias if self.bias is not None else out
        out_parallel = F.linear(inp, self.weight)
        out = mp_ops._mp_allreduce(out_parallel, group=self.model_parallel_group,
                                   skip_c_identity_dynamic=self.mp_skip_c_identity)
        return out + self.bias if self.bias is not None else out


class ParallelCrossEntropy(Layer):
    def __init__(self, mp_group=None, name=None, ignore_index=-100):
        super().__init__()
        self.name = name
        self.model_parallel_group, self.world_size, self.rank = _resolve_group(mp_group)
        self.ignore_index = ignore_index

    def forward(self, input, label):
        return mp_ops._c_softmax_with_cross_entropy(input, label, group=self.model_parallel_group,
                                                    ignore_index=self.ignore_index)
