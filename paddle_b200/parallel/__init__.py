"""B200-specific parallel runtime: symmetric memory over NVLink 5 / NVSwitch and the fused compute+collective
kernels built on it (``symm``), plus the Paddle-API parallel strategies re-exported for convenience."""
from . import symm  # noqa: F401
