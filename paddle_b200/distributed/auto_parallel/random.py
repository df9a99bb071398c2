"""Randomness of distributed tensors (reference: python/paddle/distributed/auto_parallel/random.py:36-160).

Two requirements on a random op (dropout) applied to a DistTensor:
  * independence - along a mesh axis that SHARDS the tensor, ranks must draw different numbers (they hold different
    elements of one tensor);
  * consistency  - along a mesh axis that replicates it, ranks must draw the same numbers (they hold the same elements).
``parallel_manual_seed(seed)`` switches this control on.  The generator seed of an op is then derived from the basic
seed, the mesh, and the rank's coordinates along the sharding axes only - so it is equal exactly across the ranks that
must agree - plus a per-seed call counter, so that successive calls differ."""
from __future__ import annotations

import contextlib

import torch

from .. import collective as C

_state = {"enabled": False, "seed": 42, "name": "", "calls": {}}
_MESH_OFFSET = 173
_DIM_OFFSETS = [11, 23, 37, 73, 101, 131]


def is_enable_auto_rand_ctrl():
    return _state["enabled"]


def enable_auto_rand_ctrl():
    _state["enabled"] = True


def parallel_manual_seed(seed, name=""):
    """Enable the control and set the basic seed (call once, before training; the seed may depend on the step)."""
    _state.update(enabled=True, seed=int(seed), name=name, calls={})


def determinate_seed(mesh, placements):
    """Seed shared by exactly the ranks that hold the same elements of a tensor placed like this."""
    seed = _state["seed"] + _MESH_OFFSET * (hash(tuple(mesh.process_ids)) % 9973)
    coord = mesh.coord_of(C.get_rank()) or tuple([0] * mesh.ndim)
    for ax, pl in enumerate(placements):
        if pl.is_shard() or pl.is_partial():
            seed += _DIM_OFFSETS[ax % len(_DIM_OFFSETS)] * (coord[ax] + 1) * (31 ** ax)
    return seed % (2 ** 31 - 1)


@contextlib.contextmanager
def rng_for(mesh, placements, device):
    """Run the body with the default generator of ``device`` seeded for this placement; the ambient generator state is
    restored afterwards (the op does not consume from the user's stream of random numbers)."""
    if not _state["enabled"]:
        yield
        return
    seed = determinate_seed(mesh, placements)
    n = _state["calls"].get(seed, 0)
    _state["calls"][seed] = n + 1
    is_cuda = getattr(device, "type", str(device)) == "cuda"
    saved = torch.cuda.get_rng_state(device) if is_cuda else torch.get_rng_state()
    try:
        if is_cuda:
            g = torch.cuda.default_generators[device.index if device.index is not None else torch.cuda.current_device()]
            g.manual_seed(seed + 7919 * n)
        else:
            torch.manual_seed(seed + 7919 * n)
        yield
    finally:
        if is_cuda:
            torch.cuda.set_rng_state(saved, device)
        else:
            torch.set_rng_state(saved)
