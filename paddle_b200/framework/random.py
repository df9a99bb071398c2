"""Seeding and RNG state (reference: python/paddle/framework/random.py; phi::Generator indexable states used by
fleet/layers/mpu/random.py).  States are torch generator states per device; named states for the model-parallel
RNG tracker are kept in that module."""
from __future__ import annotations

import torch

_seed = 0


def seed(s: int):
    global _seed
    _seed = int(s)
    torch.manual_seed(_seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(_seed)
    return _seed


def get_seed():
    return _seed


def get_rng_state(device=None):
    if torch.cuda.is_available() and (device is None or str(device).startswith("gpu")):
        return [torch.cuda.get_rng_state(i) for i in range(torch.cuda.device_count())]
    return [torch.get_rng_state()]


def set_rng_state(state_list, device=None):
    if torch.cuda.is_available() and (device is None or str(device).startswith("gpu")):
        for i, s in enumerate(state_list):
            torch.cuda.set_rng_state(s, i)
    else:
        torch.set_rng_state(state_list[0])


def get_cuda_rng_state():
    return [torch.cuda.get_rng_state(i) for i in range(torch.cuda.device_count())] if torch.cuda.is_available() else []


def set_cuda_rng_state(state_list):
    for i, s in enumerate(state_list):
        torch.cuda.set_rng_state(s, i)


def current_rng_state():
    """RNG state of the device ops currently draw from (cuda if a GPU is the current place)."""
    from .place import current_place
    p = current_place()
    if p.is_gpu_place():
        return torch.cuda.get_rng_state(p.get_device_id())
    return torch.get_rng_state()


def set_current_rng_state(state):
    from .place import current_place
    p = current_place()
    if p.is_gpu_place():
        torch.cuda.set_rng_state(state, p.get_device_id())
    else:
        torch.set_rng_state(state)
