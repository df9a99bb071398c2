"""Functions started by paddle.distributed.spawn in tests/test_spawn.py (they must be importable by the children)."""
import os


def allreduce_rank(scale):
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    import torch

    import paddle
    import paddle.distributed as dist
    dist.init_parallel_env()
    t = torch.tensor([float(dist.get_rank() + 1) * scale])
    dist.all_reduce(t)
    return dist.get_rank(), dist.get_world_size(), float(t)


def fails_on_rank_one():
    if os.environ["PADDLE_TRAINER_ID"] == "1":
        raise ValueError("boom from rank 1")
    import time
    time.sleep(30)
