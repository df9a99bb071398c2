from .group_sharded import group_sharded_parallel, save_group_sharded_model  # noqa: F401
