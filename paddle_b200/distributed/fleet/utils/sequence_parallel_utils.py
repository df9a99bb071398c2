"""Megatron-style sequence parallelism inside the mp group (reference: python/paddle/distributed/fleet/utils/
sequence_parallel_utils.py:42-681).  Activations between TP regions are [s/n, b, h]; ColumnSequenceParallelLinear =
all-gather -> GEMM, RowSequenceParallelLinear = GEMM -> reduce-scatter; parameters that only see a sequence shard
(norm weights, RowSP bias) are marked and their grads all-reduced over mp.

Fused NVLink paths: with FLAGS_use_fused_comm the two linears run K5 ``ag_gemm`` (TMA producer reads A tiles from
peer symmetric buffers) and K6 ``gemm_rs`` (epilogue pushes partial tiles to the owner rank); the NCCL composition
below is the baseline and the oracle."""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.collective import ReduceOp
from paddle_b200.nn import functional as F
from paddle_b200.nn.layer.layers import Layer


def _hcg():
    from paddle_b200.distributed import fleet
    return fleet.get_hybrid_communicate_group()


def _mp_group():
    return _hcg().get_model_parallel_group()


def scatter(input):
    g = _mp_group()
    n, r = g.nranks, g.rank
    seq = input.shape[0]
    assert seq % n == 0, f"Input sequence length {seq} can't be divided exactly by sequence parallelism {n}"
    per = seq // n
    return input[r * per:(r + 1) * per].contiguous()


def all_gather(input):
    g = _mp_group()
    n = g.nranks
    out = torch.empty((input.shape[0] * n,) + tuple(input.shape[1:]), dtype=input.dtype, device=input.device)
    comm.all_gather(out, input.contiguous(), g)
    return out


def reduce_scatter(input):
    g = _mp_group()
    n = g.nranks
    assert input.shape[0] % n == 0, f"Input sequence length {input.shape[0]} can't be divided exactly by sequence parallelism {n}"
    out = torch.empty((input.shape[0] // n,) + tuple(input.shape[1:]), dtype=input.dtype, device=input.device)
    comm.reduce_scatter(out, input.contiguous(), ReduceOp.SUM, g)
    return out


class ScatterOp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input):
        return scatter(input)

    @staticmethod
    def backward(ctx, grad):
        return all_gather(grad)


class GatherOp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input):
        return all_gather(input)

    @staticmethod
    def backward(ctx, grad):
        return scatter(grad)


class AllGatherOp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input):
        return all_gather(input)

    @staticmethod
    def backward(ctx, grad):
        return reduce_scatter(grad)


class ReduceScatterOp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input):
        return reduce_scatter(input)

    @staticmethod
    def backward(ctx, grad):
        return all_gather(grad)


def mark_as_sequence_parallel_parameter(parameter):
    parameter.sequence_parallel = True


def is_sequence_parallel_parameter(parameter):
    return getattr(parameter, "sequence_parallel", False)


def create_fused_allreduce_gradient_hook(parameter_list, accumulation_steps):
    g = _mp_group()
    state = {"step": 0}
    total = accumulation_steps * len(parameter_list)

    @torch.no_grad()
    def hook(p):
        state["step"] += 1
        if state["step"] % total == 0:
            grads = [getattr(q, "main_grad", None) if getattr(q, "main_grad", None) is not None else q.grad for q in parameter_list]
            flat = torch.cat([x.reshape(-1) for x in grads])
            comm.all_reduce(flat, ReduceOp.SUM, g)
            off = 0
            for x in grads:
                x.copy_(flat[off:off + x.numel()].view_as(x))
                off += x.numel()
    return hook


def create_non_fused_allreduce_gradient_hook(param, accumulation_steps):
    g = _mp_group()
    state = {"step": 0}

    @torch.no_grad()
    def hook(p):
        state["step"] += 1
        if state["step"] % accumulation_steps == 0:
            gr = getattr(p, "main_grad", None)
            gr = gr if gr is not None else p.grad
            if gr is not None:
                comm.all_reduce(gr, ReduceOp.SUM, g)
    return hook


def register_sequence_parallel_allreduce_hooks(model, accumulation_steps, fuse_sequence_parallel_allreduce):
    if accumulation_steps <= 0 or not torch.distributed.is_initialized():
        return
    hcg = _hcg()
    if hcg is None or hcg.get_model_parallel_world_size() <= 1:
        return
    params = [p for p in model.parameters() if is_sequence_parallel_parameter(p) and p.requires_grad]
    if fuse_sequence_parallel_allreduce and params:
        hook = create_fused_allreduce_gradient_hook(params, accumulation_steps)
        for p in params:
            p.register_post_accumulate_grad_hook(hook)
    else:
        for p in params:
            p.register_post_accumulate_grad_hook(create_non_fused_allreduce_gradient_hook(p, accumulation_steps))


class _SPColumnFn(torch.autograd.Function):
    """all-gather -> GEMM forward; dX GEMM -> reduce-scatter overlapped with the dW GEMM backward
    (reference SPInnerOverlapLinear, :255-424; ``recompute_allgather`` drops the gathered input)."""

    @staticmethod
    def forward(ctx, x, w, b, group, recompute_allgather, fuse_grad_add):
        from paddle_b200.parallel import symm
        full = symm.all_gather_seq(x, group)
        y = torch.matmul(full, w)
        if b is not None:
            y = y + b
        ctx.save_for_backward(x if recompute_allgather else full, w)
        ctx.cfg = (group, recompute_allgather, b is not None, fuse_grad_add)
        ctx.wref = w
        return y

    @staticmethod
    def backward(ctx, dy):
        from paddle_b200.parallel import symm
        saved, w = ctx.saved_tensors
        group, re_ag, has_b, fuse = ctx.cfg
        full = symm.all_gather_seq(saved, group) if re_ag else saved
        dfull = torch.matmul(dy, w.t())
        n = group.nranks
        dx = torch.empty((dfull.shape[0] // n,) + tuple(dfull.shape[1:]), dtype=dfull.dtype, device=dfull.device)
        task = comm.reduce_scatter(dx, dfull, ReduceOp.SUM, group, sync_op=False)
        x2, dy2 = full.reshape(-1, full.shape[-1]), dy.reshape(-1, dy.shape[-1])
        mg = getattr(ctx.wref, "main_grad", None)
        if fuse and mg is not None:
            mg.add_(torch.matmul(x2.t(), dy2).to(mg.dtype))
            dw = None
        else:
            dw = torch.matmul(x2.t(), dy2)
        db = dy2.sum(0) if has_b else None
        task.wait()
        return dx, dw, db, None, None, None


class _FusedColSP(torch.autograd.Function):
    """K5 forward (all-gather fused into the GEMM's TMA producer), K6 for dX in backward."""

    @staticmethod
    def forward(ctx, x, w, group):
        from paddle_b200.parallel import symm
        shp = x.shape
        y = symm.ag_gemm(x.reshape(-1, shp[-1]), w, group)
        ctx.save_for_backward(x, w)
        ctx.group = group
        return y.reshape(shp[0] * group.nranks, *shp[1:-1], w.shape[1])

    @staticmethod
    def backward(ctx, dy):
        from paddle_b200.parallel import symm
        x, w = ctx.saved_tensors
        g = ctx.group
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        dx = symm.gemm_rs(dy2, w, g, trans_b=True).reshape(x.shape)            # dY W^T -> reduce-scatter, fused
        full = symm.all_gather_seq(x, g).reshape(-1, x.shape[-1])
        mg = getattr(w, "main_grad", None)
        dw = torch.matmul(full.t(), dy2)
        return dx, dw, None


class _FusedRowSP(torch.autograd.Function):
    """K6 forward (partial tiles pushed to the owner as they leave TMEM), K5 for dX in backward."""

    @staticmethod
    def forward(ctx, x, w, group):
        from paddle_b200.parallel import symm
        shp = x.shape
        y = symm.gemm_rs(x.reshape(-1, shp[-1]), w, group)
        ctx.save_for_backward(x, w)
        ctx.group = group
        return y.reshape(shp[0] // group.nranks, *shp[1:-1], w.shape[1])

    @staticmethod
    def backward(ctx, dy):
        from paddle_b200.parallel import symm
        x, w = ctx.saved_tensors
        g = ctx.group
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        dx = symm.ag_gemm(dy2, w, g, trans_b=True).reshape(x.shape)           # all-gather(dY) W^T, fused
        dfull = symm.all_gather_seq(dy, g).reshape(-1, dy.shape[-1])
        dw = torch.matmul(x.reshape(-1, x.shape[-1]).t(), dfull)
        return dx, dw, None


def _fused_sp_ok(x, w, group, gathered_rows_from_x):
    from paddle_b200.parallel import symm
    if not (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16):
        return False
    rows = x.numel() // x.shape[-1]
    rows_per_rank = rows if gathered_rows_from_x else rows // group.nranks
    return symm.fused_gemm_ok(group, rows_per_rank, w.shape[0], w.shape[1], x.dtype)


class ColumnSequenceParallelLinear(Layer):
    def __init__(self, in_features, out_features, weight_attr=None, has_bias=None, gather_output=True,
                 fuse_matmul_bias=False, mp_group=None, name=None):
        super().__init__()
        from paddle_b200.distributed.fleet.layers.mpu.mp_layers import _mp_cfg, _resolve_group, _sharded_init_ctx
        self.model_parallel_group, self.world_size, _ = _resolve_group(mp_group)
        self.is_mp = self.world_size > 1
        assert gather_output is False, "If sequence_parallel is True, gather_output is False"
        assert out_features % self.world_size == 0
        self.output_size_per_partition = out_features // self.world_size
        with _sharded_init_ctx(self.is_mp):
            self.weight = self.create_parameter([in_features, self.output_size_per_partition], attr=weight_attr)
        self.weight.is_distributed = self.is_mp
        if self.is_mp:
            self.weight.split_axis = 1
        self.bias = None
        if has_bias:
            self.bias = self.create_parameter([self.output_size_per_partition], is_bias=True)
            self.bias.is_distributed = self.is_mp
            if self.is_mp:
                self.bias.split_axis = 0
        self.recompute_allgather = bool(_mp_cfg("recompute_allgather"))
        self.fuse_grad_add = bool(_mp_cfg("mp_fused_linear_param_grad_add"))
        self.overlap = bool(_mp_cfg("mp_async_allreduce"))

    def forward(self, x):
        if not self.is_mp:
            return F.linear(x, self.weight, self.bias)
        if _fused_sp_ok(x, self.weight, self.model_parallel_group, True):
            y = _FusedColSP.apply(x, self.weight, self.model_parallel_group)
            return y + self.bias if self.bias is not None else y
        if self.overlap:
            return _SPColumnFn.apply(x, self.weight, self.bias, self.model_parallel_group, self.recompute_allgather,
                                     self.fuse_grad_add)
        return F.linear(AllGatherOp.apply(x), self.weight, self.bias)


class RowSequenceParallelLinear(Layer):
    def __init__(self, in_features, out_features, weight_attr=None, has_bias=True, input_is_parallel=False,
                 fuse_matmul_bias=False, mp_group=None, name=None):
        super().__init__()
        from paddle_b200.distributed.fleet.layers.mpu.mp_layers import _resolve_group, _sharded_init_ctx
        self.model_parallel_group, self.world_size, self.rank = _resolve_group(mp_group)
        self.is_mp = self.world_size > 1
        assert input_is_parallel is True, "If sequence_parallel is True, input_is_parallel should be true."
        assert in_features % self.world_size == 0
        self.input_size_per_partition = in_features // self.world_size
        with _sharded_init_ctx(self.is_mp):
            self.weight = self.create_parameter([self.input_size_per_partition, out_features], attr=weight_attr)
        self.weight.is_distributed = self.is_mp
        if self.is_mp:
            self.weight.split_axis = 0
        self.bias = None
        if has_bias:
            self.bias = self.create_parameter([out_features], is_bias=True)
            if self.is_mp:
                mark_as_sequence_parallel_parameter(self.bias)

    def forward(self, x):
        if not self.is_mp:
            return F.linear(x, self.weight, self.bias)
        if _fused_sp_ok(x, self.weight, self.model_parallel_group, False):
            out = _FusedRowSP.apply(x, self.weight, self.model_parallel_group)
        else:
            out = ReduceScatterOp.apply(F.linear(x, self.weight))
        return out + self.bias if self.bias is not None else out
