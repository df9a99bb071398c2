"""Worker for the fault-injection tests: a tiny data-parallel loop; FAULT_INJECT decides who dies / hangs and when."""
import os
import sys

import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist

paddle.set_flags({"FLAGS_enable_async_trace": True, "FLAGS_comm_watchdog_timeout_ms": 1500, "FLAGS_comm_watchdog_poll_ms": 200,
                  "FLAGS_comm_watchdog_abort": True})
dist.init_parallel_env()
r = dist.get_rank()
paddle.seed(3)
m = paddle.nn.Linear(8, 8)
dp = paddle.DataParallel(m)
opt = paddle.optimizer.SGD(0.1, parameters=m.parameters())
for s in range(4):
    x = torch.randn(4, 8)
    dp(x).sum().backward()
    opt.step()
    opt.clear_grad()
print(f"fault_worker done rank={r} attempt={os.environ.get('PADDLE_RESTART_COUNT', '0')}", flush=True)
