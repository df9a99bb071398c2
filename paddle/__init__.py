"""Alias package: ``import paddle`` == ``import paddle_b200`` (same module objects, see paddle_b200/_alias.py)."""
import sys

import paddle_b200 as _impl

sys.modules[__name__] = _impl
