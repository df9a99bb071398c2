"""Recompute (reference: python/paddle/distributed/fleet/recompute/{recompute,recompute_hybrid}.py)."""
from .recompute import recompute, recompute_sequential  # noqa: F401
from .recompute_hybrid import recompute_hybrid  # noqa: F401
