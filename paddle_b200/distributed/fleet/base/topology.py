"""Hybrid-parallel topology (reference: python/paddle/distributed/fleet/base/topology.py:37-579).

A cartesian rank grid over the axes [data, pipe, sharding, sep, model] (outer -> inner, order user-overridable);
one communicator per axis line plus the check groups, fused dp x sep / pp x mp groups and pipeline neighbours.
With ``mp`` innermost, tensor-parallel peers are adjacent GPUs; on NVSwitch every pair has full bandwidth, so the
order only matters for rank naming, not for link locality.
"""
from __future__ import annotations

import collections
import itertools
import os
from functools import reduce as _reduce

from ... import collective as C
from ... import communication as comm

_HYBRID_PARALLEL_GROUP = None
_use_four_directions = os.environ.get("PADDLE_USE_FOUR_DIRECTIONS_P2P", "0") in ("1", "true", "True")


class ParallelMode:
    DATA_PARALLEL = 0
    TENSOR_PARALLEL = 1
    PIPELINE_PARALLEL = 2
    SHARDING_PARALLEL = 3
    SEGMENT_PARALLEL = 4


class CommunicateTopology:
    def __init__(self, hybrid_group_names=("data", "pipe", "sharding", "sep", "model"), dims=(1, 1, 1, 1, 1)):
        self._parallel_names = list(hybrid_group_names)
        self._dims = list(dims)
        self.coordinate = collections.namedtuple("Coordinate", self._parallel_names)
        self._world_size = _reduce(lambda x, y: x * y, self._dims, 1)
        coords = [self.coordinate(*x) for x in itertools.product(*[range(d) for d in self._dims])]
        self._coord2rank = {c: i for i, c in enumerate(coords)}
        self._rank2coord = {i: c for c, i in self._coord2rank.items()}

    def get_hybrid_group_names(self):
        return self._parallel_names

    def get_dim(self, axis_name):
        return self._dims[self._parallel_names.index(axis_name)]

    get_dim_size = get_dim

    def world_size(self):
        return self._world_size

    def get_rank(self, **args):
        assert len(args) == len(self._dims)
        return self._coord2rank[self.coordinate(**args)]

    def get_coord(self, rank):
        return self._rank2coord[rank]

    def get_axis_list(self, axis_name, index):
        axis = self._parallel_names.index(axis_name)
        return sorted(r for c, r in self._coord2rank.items() if c[axis] == index)

    def get_comm_list(self, axis_name):
        """All rank lists that vary only along ``axis_name``."""
        axis = self._parallel_names.index(axis_name)
        others = [range(d) for i, d in enumerate(self._dims) if i != axis]
        out = []
        for fixed in itertools.product(*others):
            ranks = []
            for k in range(self._dims[axis]):
                full = list(fixed[:axis]) + [k] + list(fixed[axis:])
                ranks.append(self._coord2rank[self.coordinate(*full)])
            out.append(ranks)
        return out

    def get_fused_ranks(self, fused_axis):
        non_fused = [n for n in self._parallel_names if n not in fused_axis]
        out = []
        for nf in itertools.product(*[range(self.get_dim(n)) for n in non_fused]):
            base = dict(zip(non_fused, nf))
            ranks = []
            for f in itertools.product(*[range(self.get_dim(n)) for n in fused_axis]):
                d = dict(base)
                d.update(zip(fused_axis, f))
                ranks.append(self._coord2rank[self.coordinate(**d)])
            out.append(sorted(ranks))
        return out

    def get_rank_from_stage(self, global_rank, **kwargs):
        coord = self.get_coord(global_rank)._replace(**kwargs)
        return self._coord2rank[coord]


class HybridCommunicateGroup:
    def __init__(self, topology):
        global _HYBRID_PARALLEL_GROUP
        self.nranks = C.get_world_size()
        self.global_rank = C.get_rank()
        self._topo = topology
        self._dp_degree = topology.get_dim("data")
        self._mp_degree = topology.get_dim("model")
        self._pp_degree = topology.get_dim("pipe")
        self._sharding_degree = topology.get_dim("sharding")
        self._sep_degree = topology.get_dim("sep") if "sep" in topology.get_hybrid_group_names() else 1
        coord = topology.get_coord(self.global_rank)
        self._data_parallel_id = coord.data
        self._model_parallel_id = coord.model
        self._sharding_parallel_id = coord.sharding
        self._sep_parallel_id = getattr(coord, "sep", 0)
        self.stage_id = coord.pipe
        assert self._check_valid_topo(), (
            f"nranks: {self.nranks}, mp_num: {self._mp_degree}, sharding_num: {self._sharding_degree}, "
            f"pp_num: {self._pp_degree}, dp_num: {self._dp_degree}, sep_num: {self._sep_degree}")

        self._pp_group, self._pp_comm_group = self._set_comm_group("pipe")
        if self.nranks > 1:
            import torch
            dev = "cuda" if C._state.backend == "nccl" else "cpu"
            comm.all_reduce(torch.zeros(1, dtype=torch.int32, device=dev), group=self._pp_comm_group)
        self._dp_group, self._dp_comm_group = self._set_comm_group("data")
        self._mp_group, self._mp_comm_group = self._set_comm_group("model")
        self._sharding_group, self._sharding_comm_group = self._set_comm_group("sharding")
        self._sep_group, self._sep_comm_group = None, None
        if self._sep_degree > 1:
            self._sep_group, self._sep_comm_group = self._set_comm_group("sep")
        self._check_group, self._check_comm_group = self._set_check_group("data")
        if self._sharding_degree > 1:
            self.sharding_check_group, self.sharding_check_comm_group = self._set_check_group("sharding")
        if self._sep_degree > 1:
            self._dp_sep_group, self._dp_sep_comm_group = self.create_fuse_group(["data", "sep"])
            self._pp_mp_group, self._pp_mp_comm_group = self.create_fuse_group(["pipe", "model"])
        self.sharding_check_group, self.sharding_check_comm_group = self._set_check_group("sharding")

        self.is_first_stage = self.stage_id == 0
        self.is_last_stage = self.stage_id == self._pp_degree - 1
        if self._pp_degree > 1:
            self._set_p2p_prev_next()
            if _use_four_directions:
                self._set_four_directions_p2p_group()
        _HYBRID_PARALLEL_GROUP = self

    # ------------------------------------------------------------------ construction helpers
    def _check_valid_topo(self):
        return self._dp_degree * self._mp_degree * self._pp_degree * self._sharding_degree * self._sep_degree == self.nranks

    def _set_comm_group(self, parallel_method="data"):
        mine, mine_comm = [], None
        for ranks in self._topo.get_comm_list(parallel_method):
            g = C.new_group(ranks=ranks)
            if self.global_rank in ranks:
                mine, mine_comm = ranks, g
        assert mine and mine_comm is not None
        return mine, mine_comm

    def _set_check_group(self, parallel_method="data"):
        mine, mine_comm = [], None
        for idx in range(self._topo.get_dim(parallel_method)):
            ranks = self._topo.get_axis_list(parallel_method, idx)
            g = C.new_group(ranks=ranks)
            if self.global_rank in ranks:
                mine, mine_comm = ranks, g
        assert mine and mine_comm is not None
        return mine, mine_comm

    def create_fuse_group(self, fused_strategy_list):
        assert len(fused_strategy_list) > 0
        mine, mine_comm = [], None
        for ranks in self._topo.get_fused_ranks(fused_strategy_list):
            g = C.new_group(ranks=ranks)
            if self.global_rank in ranks:
                mine, mine_comm = ranks, g
        return mine, mine_comm

    def _set_p2p_prev_next(self):
        for ranks in self._topo.get_comm_list("pipe"):
            for idx, r in enumerate(ranks):
                if r == self.global_rank:
                    self.next_rank = ranks[(idx + 1) % self._pp_degree]
                    self.prev_rank = ranks[(idx - 1) % self._pp_degree]

    def _set_four_directions_p2p_group(self):
        self.send_next_group = self.send_prev_group = self.recv_next_group = self.recv_prev_group = None
        for ranks in self._topo.get_comm_list("pipe"):
            for idx, cur in enumerate(ranks):
                nxt, prv = ranks[(idx + 1) % self._pp_degree], ranks[(idx - 1) % self._pp_degree]
                g = C.new_group(ranks=[cur, nxt])
                if self.global_rank == cur:
                    self.send_next_group = g
                elif self.global_rank == nxt:
                    self.recv_prev_group = g
                g2 = C.new_group(ranks=[prv, cur])
                if self.global_rank == cur:
                    self.send_prev_group = g2
                elif self.global_rank == prv:
                    self.recv_next_group = g2

    # ------------------------------------------------------------------ queries
    def get_parallel_mode(self):
        if self._pp_degree == 1 and self._mp_degree == 1 and self._sep_degree == 1 and self._sharding_degree == 1 \
                and self._dp_degree > 1:
            return ParallelMode.DATA_PARALLEL
        if self._pp_degree == 1 and self._mp_degree == 1 and self._sep_degree == 1 and self._sharding_degree > 1:
            return ParallelMode.SHARDING_PARALLEL
        if self._pp_degree == 1 and self._mp_degree == 1 and self._sep_degree > 1:
            return ParallelMode.SEGMENT_PARALLEL
        if self._pp_degree == 1 and self._mp_degree > 1:
            return ParallelMode.TENSOR_PARALLEL
        if self._pp_degree > 1:
            return ParallelMode.PIPELINE_PARALLEL
        return ParallelMode.DATA_PARALLEL

    def topology(self):
        return self._topo

    def get_global_rank(self):
        return self.global_rank

    def get_data_parallel_rank(self):
        return self._data_parallel_id

    def get_data_parallel_world_size(self):
        return self._dp_degree

    def get_data_parallel_group(self):
        return self._dp_comm_group

    def get_data_parallel_group_src_rank(self):
        return self._dp_comm_group.ranks[0]

    def get_model_parallel_rank(self):
        return self._model_parallel_id

    def get_model_parallel_world_size(self):
        return self._mp_degree

    def get_model_parallel_group(self):
        return self._mp_comm_group

    def get_model_parallel_group_src_rank(self):
        return self._mp_comm_group.ranks[0]

    def get_stage_id(self):
        return self.stage_id

    def get_pipe_parallel_rank(self):
        return self.stage_id

    def get_pipe_parallel_world_size(self):
        return self._pp_degree

    def get_pipe_parallel_group(self):
        return self._pp_comm_group

    def get_p2p_groups(self):
        assert _use_four_directions, "If you want to use four directions p2p group, set the environment variable " \
                                     "PADDLE_USE_FOUR_DIRECTIONS_P2P to True."
        return self.send_next_group, self.send_prev_group, self.recv_next_group, self.recv_prev_group

    def _get_p2p_next_rank(self):
        return self.next_rank

    def _get_p2p_prev_rank(self):
        return self.prev_rank

    def get_sharding_parallel_rank(self):
        return self._sharding_parallel_id

    def get_sharding_parallel_world_size(self):
        return self._sharding_degree

    def get_sharding_parallel_group(self):
        return self._sharding_comm_group

    def get_sharding_parallel_group_src_rank(self):
        return self._sharding_comm_group.ranks[0]

    def get_sep_parallel_rank(self):
        return self._sep_parallel_id

    def get_sep_parallel_world_size(self):
        return self._sep_degree

    def get_sep_parallel_group(self):
        assert self._sep_degree > 1, "sep not exist"
        return self._sep_comm_group

    def get_sep_parallel_group_src_rank(self):
        assert self._sep_degree > 1, "sep not exist"
        return self._sep_comm_group.ranks[0]

    def get_dp_sep_parallel_group(self):
        assert self._sep_degree > 1, "sep not exist"
        return self._dp_sep_comm_group

    def get_pp_mp_parallel_group(self):
        assert self._sep_degree > 1, "sep not exist"
        return self._pp_mp_comm_group

    def get_check_parallel_group(self, sharding=False):
        return self.sharding_check_comm_group if sharding else self._check_comm_group

    def get_rank_from_stage(self, stage_id, **kwargs):
        return self._topo.get_rank_from_stage(self.global_rank, pipe=stage_id, **kwargs)


def _get_global_hcg():
    return _HYBRID_PARALLEL_GROUP
