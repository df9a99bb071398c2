"""MoELayer (reference: python/paddle/incubate/distributed/models/moe/moe_layer.py:99-503).

Same dataflow as the reference - gate -> count_by_gate -> local scatter by position -> global_scatter (token
all-to-all) -> experts -> global_gather -> weighted combine - with two changes: the per-expert loop slices by counts
read ONCE per layer (the reference syncs D2H three times), and when the experts are plain
``ExpertFFN`` stacks the grouped-GEMM path runs all local experts in one launch."""
from __future__ import annotations

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.utils.moe_utils import global_gather, global_scatter
from paddle_b200.nn.layer.layers import Layer, LayerList

from .gate import BaseGate, GShardGate, NaiveGate, SwitchGate
from .utils import count_by_gate


def _local_scatter(inp, pos):
    return inp.index_select(0, pos)


def _local_gather(inp, pos, out_batch_size, maybe_overlap=True):
    out = torch.zeros((out_batch_size, inp.shape[-1]), dtype=inp.dtype, device=inp.device)
    out.index_add_(0, pos, inp) if maybe_overlap else out.index_copy_(0, pos, inp)
    return out


class _AllGatherMP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, rank, world_size, group):
        ctx.cfg = (rank, world_size, inp.shape[0])
        out = torch.empty((inp.shape[0] * world_size,) + tuple(inp.shape[1:]), dtype=inp.dtype, device=inp.device)
        comm.all_gather(out, inp.contiguous(), group)
        return out

    @staticmethod
    def backward(ctx, g):
        rank, world, n = ctx.cfg
        return g[rank * n:(rank + 1) * n].contiguous(), None, None, None


class _SliceMP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, rank, world_size, group):
        n = inp.shape[0] // world_size
        ctx.cfg = (rank, world_size, group)
        return inp[rank * n:(rank + 1) * n].contiguous()

    @staticmethod
    def backward(ctx, g):
        rank, world, group = ctx.cfg
        out = torch.empty((g.shape[0] * world,) + tuple(g.shape[1:]), dtype=g.dtype, device=g.device)
        comm.all_gather(out, g.contiguous(), group)
        return out, None, None, None


def prepare_forward(gate, num_expert, world_size, moe_group):
    pos, lec, gec = count_by_gate(gate, num_expert, world_size, group=moe_group)
    with torch.no_grad():
        fwd_expert_count = gec.reshape(world_size, num_expert).sum(0)
    return pos, lec, gec, fwd_expert_count


class MoELayer(Layer):
    def __init__(self, d_model, experts, gate=None, moe_group=None, mp_group=None, recompute_interval=0,
                 recompute_ctx=None):
        super().__init__()
        self.recompute_ctx = recompute_ctx
        if gate is None:
            gate = {}
        assert isinstance(gate, (dict, BaseGate)), "gate config' type must be dict or an instance of BaseGate"
        self.group = moe_group
        self.world_size = moe_group.nranks if moe_group is not None else 1
        from .grouped_experts import GroupedExpertsFFN
        self._grouped = isinstance(experts, GroupedExpertsFFN)
        self.num_expert = len(experts)
        self.recompute_interval = recompute_interval
        assert experts is not None
        self.experts = experts if isinstance(experts, torch.nn.Module) else LayerList(experts)
        self.mp_group = mp_group
        self.d_model = d_model
        if isinstance(gate, dict):
            self.top_k = gate.get("top_k", 2)
            kind = gate.get("type", "gshard")
            if kind == "naive" or kind is None:
                gate = NaiveGate(d_model, num_expert=self.num_expert, world_size=self.world_size, topk=self.top_k)
            elif kind == "gshard":
                gate = GShardGate(d_model, num_expert=self.num_expert, world_size=self.world_size, topk=self.top_k,
                                  group=self.group)
            elif kind == "switch":
                gate = SwitchGate(d_model, num_expert=self.num_expert, world_size=self.world_size, topk=self.top_k,
                                  group=self.group)
            else:
                raise AssertionError(f"We only support naive gate, gshard gate and switch gate, but you choose {kind} gate.")
        elif isinstance(gate, NaiveGate):
            self.top_k = gate.top_k
        elif isinstance(gate, BaseGate):
            raise TypeError("Unimplemented gate type: ", type(gate))
        self.gate = gate
        for p in self.experts.parameters():
            p.no_sync = True     # experts differ per rank: skipped by sync_params_buffers (Appendix B.12)

    def _forward_fused(self, inp, value, gate, origin_shape, mp_rank, mp_size):
        """K9 / K10: dispatch and combine through NVLink symmetric memory with device-side routing; the only host
        read is the per-expert row counts that size the expert GEMMs."""
        from paddle_b200 import ops
        from paddle_b200.parallel import moe_fused
        topk = gate.shape[1] if gate.dim() == 2 else 1
        assert topk == self.top_k
        with torch.no_grad():
            lec = ops.number_count(gate, self.num_expert * self.world_size)
            pos = ops.assign_pos(gate, torch.cumsum(lec, 0), padded=self._grouped)   # padded: no host read of the total
        if self._grouped:
            # device-scheduled grouped GEMMs: the slab is used at its provisioned capacity and the row counts never
            # leave the device - zero host synchronisations in this layer's forward and backward
            x, counts_dev, meta = moe_fused.dispatch(inp, pos, lec, self.group, self.num_expert, topk, sync=False)
            x = self.experts(x, counts_dev)
            x = moe_fused.combine(x, meta).reshape(-1, self.top_k, self.d_model)
            value = value.reshape(x.shape[0], 1, self.top_k).to(x.dtype)
            x = torch.bmm(value, x).reshape(-1, self.d_model)
            if mp_size > 1:
                x = _AllGatherMP.apply(x, mp_rank, mp_size, self.mp_group)
            return x.reshape(origin_shape)
        x, counts, meta = moe_fused.dispatch(inp, pos, lec, self.group, self.num_expert, topk)

        def experts_fwd(x):
            if x.shape[0] == 0:
                return x
            ys, last = [], 0
            for idx, c in enumerate(counts):
                if c <= 0:
                    continue
                ys.append(self.experts[idx](x[last:last + c]))
                last += c
            return torch.cat(ys, 0)
        if self.recompute_interval <= 0 or x.shape[0] == 0:
            x = experts_fwd(x)
        else:
            from paddle_b200.distributed.fleet.recompute import recompute
            x = recompute(experts_fwd, x)
        x = moe_fused.combine(x, meta).reshape(-1, self.top_k, self.d_model)
        value = value.reshape(x.shape[0], 1, self.top_k).to(x.dtype)
        x = torch.bmm(value, x).reshape(-1, self.d_model)
        if mp_size > 1:
            x = _AllGatherMP.apply(x, mp_rank, mp_size, self.mp_group)
        return x.reshape(origin_shape)

    def forward(self, inp):
        assert inp.dim() == 3
        origin_shape = inp.shape
        inp = inp.reshape(-1, origin_shape[2])
        mp_rank, mp_size = 0, 1
        if self.mp_group is not None:
            mp_rank, mp_size = self.mp_group.rank, self.mp_group.nranks
        if mp_size > 1:
            inp = _SliceMP.apply(inp, mp_rank, mp_size, self.mp_group)
        value, gate = self.gate(inp)
        from paddle_b200.parallel import moe_fused
        if self.world_size > 1 and moe_fused.available(self.group, inp):
            return self._forward_fused(inp, value, gate, origin_shape, mp_rank, mp_size)
        pos, lec, gec, fwd_expert_count = prepare_forward(gate, self.num_expert, self.world_size, self.group)
        topk = gate.shape[1] if gate.dim() == 2 else 1
        assert topk == self.top_k
        temp_pos = pos // topk if pos.numel() else pos
        # ---- scatter
        x = _local_scatter(inp, temp_pos)
        if self.world_size > 1:
            x = global_scatter(x, lec, gec, group=self.group)
        # ---- experts
        if self._grouped:
            x = self.experts(x, fwd_expert_count) if x.shape[0] else x
            counts = None
        else:
            counts = fwd_expert_count.tolist()      # one host read of the counts per layer

        def experts_fwd(x):
            if x.shape[0] == 0 or counts is None:
                return x
            ys, last = [], 0
            for idx, c in enumerate(counts):
                if c <= 0:
                    continue
                ys.append(self.experts[idx](x[last:last + c]))
                last += c
            return torch.cat(ys, 0)
        if self.recompute_interval <= 0 or x.shape[0] == 0:
            x = experts_fwd(x)
        else:
            from paddle_b200.distributed.fleet.recompute import recompute
            x = recompute(experts_fwd, x)
        # ---- gather
        out_batch_size = inp.shape[0] * (gate.shape[1] if gate.dim() == 2 else 1)
        if self.world_size > 1:
            x = global_gather(x, lec, gec, group=self.group)
        x = _local_gather(x, pos, out_batch_size, maybe_overlap=False)
        x = x.reshape(-1, self.top_k, self.d_model)
        value = value.reshape(x.shape[0], 1, self.top_k).to(x.dtype)
        x = torch.bmm(value, x).reshape(-1, self.d_model)
        if mp_size > 1:
            x = _AllGatherMP.apply(x, mp_rank, mp_size, self.mp_group)
        return x.reshape(origin_shape)
