"""Worker (2 processes): paddle.distributed.rpc round trips (sync, async, remote exception, worker infos)."""
import os
import sys

import paddle_b200.distributed.rpc as rpc


def add(a, b):
    return a + b


def whoami():
    return rpc.get_current_worker_info().name


def boom():
    raise ValueError("remote failure")


rank = int(os.environ["PADDLE_TRAINER_ID"])
rpc.init_rpc(f"worker{rank}", rank=rank, world_size=2, master_endpoint=os.environ["RPC_MASTER"])
other = f"worker{1 - rank}"
assert rpc.rpc_sync(other, add, args=(2, 3)) == 5
assert rpc.rpc_sync(other, whoami) == other
futs = [rpc.rpc_async(other, add, args=(i, i)) for i in range(8)]
assert [f.wait() for f in futs] == [2 * i for i in range(8)]
try:
    rpc.rpc_sync(other, boom)
    raise AssertionError("the remote exception must be re-raised")
except ValueError as e:
    assert "remote failure" in str(e)
infos = rpc.get_all_worker_infos()
assert [w.name for w in infos] == ["worker0", "worker1"] and rpc.get_worker_info(other).rank == 1 - rank
rpc.shutdown()
print("rpc ok", rank)
