"""Fleet facade (reference: python/paddle/distributed/fleet/fleet.py:167-662)."""
from __future__ import annotations

import os
import time

import torch

from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.parallel import ParallelEnv, init_parallel_env

from .base import topology as tp
from .base.distributed_strategy import DistributedStrategy
from .base.role_maker import PaddleCloudRoleMaker, RoleMakerBase
from .utils.log_util import logger, set_log_level

_ORDER_NAME = {"dp": "data", "pp": "pipe", "sharding": "sharding", "sep": "sep", "mp": "model"}


class Fleet:
    def __init__(self):
        self._role_maker = None
        self._user_defined_strategy = None
        self._is_collective = False
        self._hcg = None
        self._topology = None
        self.strategy_compiler = None
        self._runtime_handle = None

    def init(self, role_maker=None, is_collective=False, strategy=None, log_level="INFO"):
        set_log_level(log_level)
        if strategy is None:
            strategy = DistributedStrategy()
        self._user_defined_strategy = strategy
        if role_maker is None:
            if not isinstance(is_collective, bool):
                raise ValueError(f"`is_collective` should be instance of `bool`, but got {type(is_collective)}")
            self._is_collective = is_collective
            self._role_maker = PaddleCloudRoleMaker(is_collective=is_collective)
        elif isinstance(role_maker, RoleMakerBase):
            self._role_maker = role_maker
            self._is_collective = role_maker._is_collective()
        else:
            raise ValueError(f"`role_maker` should be subclass of `RoleMakerBase`, but got {type(role_maker)}")
        self._role_maker._generate_role() if hasattr(self._role_maker, "_generate_role") else None
        if not self._is_collective:
            raise NotImplementedError("parameter-server mode (is_collective=False) is out of scope (SURVEY §2.2 N15)")
        os.environ["FLAGS_nccl_nrings"] = str(strategy.nccl_comm_num)
        init_parallel_env()
        self._init_hybrid_parallel_env()
        return self

    def _init_hybrid_parallel_env(self):
        st = self._user_defined_strategy
        hc = st.hybrid_configs
        world = C.get_world_size()
        mp, pp, sh, sep = int(hc["mp_degree"]), int(hc["pp_degree"]), int(hc["sharding_degree"]), int(hc["sep_degree"])
        dp = int(hc["dp_degree"])
        assert mp >= 0 and pp >= 0 and sh >= 0 and sep >= 0
        mp, pp, sh, sep = max(mp, 1), max(pp, 1), max(sh, 1), max(sep, 1)
        if dp < 0:
            assert world % (mp * pp * sh * sep) == 0, \
                f"world size {world} is not divisible by mp*pp*sharding*sep = {mp * pp * sh * sep}"
            dp = world // (mp * pp * sh * sep)
        hc["dp_degree"] = dp
        degree = {"dp": dp, "pp": pp, "sharding": sh, "sep": sep, "mp": mp}
        order = list(st.hybrid_parallel_order)
        names = [_ORDER_NAME[o] for o in order]
        dims = [degree[o] for o in order]
        self._topology = tp.CommunicateTopology(hybrid_group_names=names, dims=dims)
        self._hcg = tp.HybridCommunicateGroup(self._topology)
        if mp > 1:
            from .layers.mpu.random import model_parallel_random_seed
            seed = st.tensor_parallel_configs["tensor_init_seed"]
            model_parallel_random_seed(None if seed == -1 else seed)

    # ------------------------------------------------------------- info
    def get_hybrid_communicate_group(self):
        return self._hcg

    def get_hybrid_parallel_topology(self):
        return self._topology

    def is_first_worker(self):
        return self.worker_index() == 0

    def worker_index(self):
        return C.get_rank()

    def worker_num(self):
        return C.get_world_size()

    def node_num(self):
        return int(os.environ.get("PADDLE_NNODES", "1"))

    def local_rank(self):
        return ParallelEnv()._local_rank

    def local_device_ids(self):
        return [ParallelEnv().device_id]

    def world_device_ids(self):
        return list(range(C.get_world_size()))

    def is_worker(self):
        return True

    def is_server(self):
        return False

    def worker_endpoints(self, to_string=False):
        eps = ParallelEnv().trainer_endpoints
        return ",".join(eps) if to_string else eps

    def barrier_worker(self):
        C.barrier()

    def stop_worker(self):
        pass

    def init_worker(self, scopes=None):
        pass

    # ------------------------------------------------------------- wrapping
    def distributed_model(self, model):
        from .model import distributed_model
        self._model = distributed_model(model)
        return self._model

    def distributed_optimizer(self, optimizer, strategy=None):
        from .optimizer import distributed_optimizer
        if strategy is not None:
            self._user_defined_strategy = strategy
        self._optimizer = distributed_optimizer(optimizer, strategy)
        return self._optimizer

    def distributed_scaler(self, scaler):
        from .scaler import distributed_scaler
        self._scaler = distributed_scaler(scaler)
        return self._scaler

    # ------------------------------------------------------------- self benchmark
    def collective_perf(self, comm_type, round=50, size_and_time=None):
        """Built-in collective micro-benchmark (reference: fleet.py:363-597).  Device-timed with CUDA events,
        10 warm-up rounds; ``size_and_time`` = {bytes: max_seconds} thresholds, default sweep 1 MB..1 GB."""
        size_and_time = size_and_time or {}
        sizes = sorted(size_and_time) if size_and_time else [1 << p for p in range(20, 31)]
        dev = "cuda" if C._state.backend == "nccl" else "cpu"
        n = C.get_world_size()
        results = {}
        for nbytes in sizes:
            numel = nbytes // 4
            x = torch.zeros(max(numel // (n if comm_type in ("allgather",) else 1), 1), dtype=torch.float32, device=dev)
            out = torch.zeros(max(numel, n), dtype=torch.float32, device=dev)

            def run():
                if comm_type == "allreduce":
                    comm.all_reduce(out)
                elif comm_type == "reduce":
                    comm.reduce(out, 0)
                elif comm_type == "broadcast":
                    comm.broadcast(out, 0)
                elif comm_type == "allgather":
                    comm.all_gather(out[: x.numel() * n], x)
                elif comm_type == "reduce_scatter":
                    comm.reduce_scatter(out[: out.numel() // n], out[: out.numel() // n * n])
                else:
                    raise ValueError(f"unknown comm_type {comm_type}")
            for _ in range(10):
                run()
            if dev == "cuda":
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(round):
                    run()
                e1.record()
                torch.cuda.synchronize()
                sec = e0.elapsed_time(e1) / 1e3 / round
            else:
                t0 = time.perf_counter()
                for _ in range(round):
                    run()
                sec = (time.perf_counter() - t0) / round
            results[nbytes] = sec
            thr = size_and_time.get(nbytes)
            if thr is not None and thr > 0 and sec > thr:
                logger.warning(f"[Perf Warning] {comm_type} Test Timeout! {sec} > {thr}")
            logger.info(f"[{comm_type}] {nbytes / (1 << 20):.0f} MB: {sec * 1e3:.3f} ms")
        return results

    # the per-collective entry points ``collective_perf`` dispatches to in the reference (fleet.py:363-560)
    def allreduce_perf(self, iteration, x=None, y=None, pred_time=None, size=None):
        return self._one_perf("allreduce", iteration, x, pred_time, size)

    def reduce_perf(self, iteration, x=None, pred_time=None, size=None):
        return self._one_perf("reduce", iteration, x, pred_time, size)

    def broadcast_perf(self, iteration, x=None, pred_time=None, size=None):
        return self._one_perf("broadcast", iteration, x, pred_time, size)

    def allgather_perf(self, iteration, x=None, y=None, pred_time=None, size=None):
        return self._one_perf("allgather", iteration, x, pred_time, size)

    def reduce_scatter_perf(self, iteration, x=None, y=None, pred_time=None, size=None):
        return self._one_perf("reduce_scatter", iteration, x, pred_time, size)

    def _one_perf(self, comm_type, iteration, x, pred_time, size):
        nbytes = int(size) if size is not None else (x.numel() * x.element_size() if x is not None else 1 << 20)
        return self.collective_perf(comm_type, round=int(iteration), size_and_time={nbytes: pred_time if pred_time else 0})[nbytes]

    def all_reduce(self, input, mode="sum"):
        """All-reduce a tensor / ndarray / number over the workers and return the result (reference: fleet.py all_reduce,
        used for metrics)."""
        import numpy as np
        t = input.clone() if isinstance(input, torch.Tensor) else torch.as_tensor(np.asarray(input))
        if C.is_initialized() and C.get_world_size() > 1:
            op = {"sum": comm.ReduceOp.SUM, "max": comm.ReduceOp.MAX, "min": comm.ReduceOp.MIN}[mode]
            dev = "cuda" if C._state.backend == "nccl" else "cpu"
            moved = t.to(dev)
            comm.all_reduce(moved, op)
            t = moved.to(t.device)
        return t if isinstance(input, torch.Tensor) else t.numpy()

    def get_loss_scaling(self):
        """Current loss scale of the scaler handed to ``distributed_scaler`` (or the strategy's initial value)."""
        scaler = getattr(self, "_scaler", None)
        if scaler is not None:
            return scaler.get_loss_scaling()
        cfg = getattr(self._strategy, "amp_configs", {}) if getattr(self, "_strategy", None) is not None else {}
        return torch.tensor([float(cfg.get("init_loss_scaling", 32768.0))])

    def amp_init(self, place=None, scope=None, test_program=None, use_fp16_test=False):
        """Static-graph AMP casts parameters here; in dygraph ``paddle.amp.decorate`` already did, so nothing is left."""
        return None

    def qat_init(self, place=None, scope=None, test_program=None):
        return None

    def save(self, dirname, feed=(), fetch=(), **configs):
        """Collective-mode checkpoint of the wrapped model (and optimizer) from the first worker:
        ``dirname/model.pdparams`` (+ ``model.pdopt``)."""
        import os
        from ...framework.io import save as _save
        model = getattr(self, "_model", None)
        if model is None:
            raise RuntimeError("fleet.save needs a model wrapped by fleet.distributed_model first")
        if self.is_first_worker():
            os.makedirs(dirname, exist_ok=True)
            _save(model.state_dict(), os.path.join(dirname, "model.pdparams"))
            opt = getattr(self, "_optimizer", None)
            if opt is not None:
                _save(opt.state_dict(), os.path.join(dirname, "model.pdopt"))
        self.barrier_worker()

    def load_model(self, path, mode=0):
        import os
        from ...framework.io import load as _load
        model = getattr(self, "_model", None)
        if model is None:
            raise RuntimeError("fleet.load_model needs a model wrapped by fleet.distributed_model first")
        model.set_state_dict(_load(os.path.join(path, "model.pdparams")))

    # static-graph entry point
    def minimize(self, *a, **k):
        raise NotImplementedError("static-graph fleet.minimize is out of scope (SURVEY P37)")

    def save_persistables(self, *a, **k):
        raise NotImplementedError("parameter-server persistence is out of scope")

    # parameter-server / federated-learning mode: not part of this framework
    def is_coordinator(self):
        return False

    def server_num(self):
        return 0

    def server_index(self):
        return -1

    def server_endpoints(self, to_string=False):
        return "" if to_string else []

    def _ps_only(self, *a, **k):
        raise NotImplementedError("this entry point belongs to the parameter-server mode, which this framework does not "
                                  "implement (collective training only)")

    init_coordinator = make_fl_strategy = get_fl_client = init_server = load_one_table = load_inference_model = _ps_only
    run_server = save_inference_model = save_cache_model = check_save_pre_patch_done = save_cache_table = _ps_only
    save_one_table = save_dense_params = set_date = shrink = _ps_only


fleet = Fleet()
