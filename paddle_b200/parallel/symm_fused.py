"""Fused NVLink paths (filled in as the kernels land; see csrc/kernels/nvl_*.cu)."""
from __future__ import annotations


def available(group, t) -> bool:
    return False


def all_gather_small(t, group, out):
    raise NotImplementedError
