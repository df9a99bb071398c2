"""paddle.distributed.fleet (reference: python/paddle/distributed/fleet/__init__.py)."""
from . import base, layers, meta_optimizers, meta_parallel, recompute as _recompute_pkg, utils  # noqa: F401
from .base.distributed_strategy import DistributedStrategy  # noqa: F401
from .base.role_maker import PaddleCloudRoleMaker, Role, UserDefinedRoleMaker  # noqa: F401
from .base.topology import CommunicateTopology, HybridCommunicateGroup, ParallelMode  # noqa: F401
from .fleet import Fleet, fleet  # noqa: F401
from .layers.mpu.random import get_rng_state_tracker  # noqa: F401
from .recompute import recompute, recompute_hybrid, recompute_sequential  # noqa: F401

init = fleet.init
distributed_model = fleet.distributed_model
distributed_optimizer = fleet.distributed_optimizer
distributed_scaler = fleet.distributed_scaler
get_hybrid_communicate_group = fleet.get_hybrid_communicate_group
worker_index = fleet.worker_index
worker_num = fleet.worker_num
node_num = fleet.node_num
local_rank = fleet.local_rank
is_first_worker = fleet.is_first_worker
is_worker = fleet.is_worker
is_server = fleet.is_server
worker_endpoints = fleet.worker_endpoints
barrier_worker = fleet.barrier_worker
init_worker = fleet.init_worker
stop_worker = fleet.stop_worker
collective_perf = fleet.collective_perf


class UtilBase:
    """Cross-worker helpers (reference: fleet/base/util_factory.py): all_reduce / barrier / all_gather of python numbers
    over the collective group, and file sharding."""

    def all_reduce(self, input, mode="sum", comm_world="worker"):
        import numpy as np
        import torch

        from .. import collective as C
        from .. import communication as comm
        t = torch.as_tensor(np.asarray(input, dtype=np.float64))
        if C._state.backend == "nccl":
            t = t.cuda()
        comm.all_reduce(t, {"sum": C.ReduceOp.SUM, "max": C.ReduceOp.MAX, "min": C.ReduceOp.MIN}[mode])
        return t.cpu().numpy()

    def barrier(self, comm_world="worker"):
        from .. import collective as C
        C.barrier()

    def all_gather(self, input, comm_world="worker"):
        from .. import communication as comm
        out = []
        comm.all_gather_object(out, input)
        return out

    def get_file_shard(self, files):
        from .. import collective as C
        n, r = C.get_world_size(), C.get_rank()
        per, rem = divmod(len(files), n)
        start = r * per + min(r, rem)
        return files[start:start + per + (1 if r < rem else 0)]

    def print_on_rank(self, message, rank_id):
        from .. import collective as C
        if C.get_rank() == rank_id:
            print(message)


util = UtilBase()


class MultiSlotDataGenerator:
    """Text-slot sample generator of the PS data pipeline (reference: fleet/data_generator): ``generate_sample`` yields
    [(slot_name, [values]), ...]; ``run_from_stdin`` prints the slot-count-prefixed line format."""

    def generate_sample(self, line):
        raise NotImplementedError

    def _gen_str(self, sample):
        return " ".join(f"{len(v)} " + " ".join(str(x) for x in v) for _, v in sample) + "\n"

    def run_from_memory(self, lines=()):
        return [self._gen_str(s) for line in lines for s in self.generate_sample(line)()]

    def run_from_stdin(self):
        import sys
        for line in sys.stdin:
            for s in self.generate_sample(line)():
                sys.stdout.write(self._gen_str(s))


class MultiSlotStringDataGenerator(MultiSlotDataGenerator):
    def _gen_str(self, sample):
        return " ".join(f"{len(v)} " + " ".join(v) for _, v in sample) + "\n"
