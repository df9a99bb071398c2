from . import distributed_strategy, role_maker, topology  # noqa: F401
