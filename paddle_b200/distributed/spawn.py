"""paddle.distributed.spawn (reference: python/paddle/distributed/spawn.py:146-618): multiprocessing 'spawn'
start method, one env dict per rank (same PADDLE_* contract as the launcher), error queue with traceback
propagation to the parent."""
from __future__ import annotations

import multiprocessing
import os
import socket
import sys
import traceback


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _env_for(rank, nprocs, endpoints, options):
    env = {
        "PADDLE_TRAINER_ID": str(rank), "PADDLE_TRAINERS_NUM": str(nprocs), "PADDLE_GLOBAL_RANK": str(rank),
        "PADDLE_GLOBAL_SIZE": str(nprocs), "PADDLE_LOCAL_RANK": str(rank), "PADDLE_LOCAL_SIZE": str(nprocs),
        "PADDLE_RANK_IN_NODE": str(rank), "PADDLE_NNODES": "1", "PADDLE_CURRENT_ENDPOINT": endpoints[rank],
        "PADDLE_TRAINER_ENDPOINTS": ",".join(endpoints), "PADDLE_MASTER": endpoints[0],
        "FLAGS_selected_gpus": str(options["gpus"][rank]) if options.get("gpus") else str(rank),
    }
    if options.get("backend"):
        env["PADDLE_DISTRI_BACKEND"] = options["backend"]
    return env


def _func_wrapper(func, args, error_queue, return_queue, env):
    os.environ.update(env)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        os.environ.pop(k, None)
    try:
        result = func(*args)
        return_queue.put((int(env["PADDLE_TRAINER_ID"]), result))
    except KeyboardInterrupt:
        pass
    except Exception:  # noqa: BLE001
        error_queue.put(traceback.format_exc())
        sys.exit(1)


class MultiprocessContext:
    def __init__(self, processes, error_queues, return_queues):
        self.processes, self.error_queues, self.return_queues = processes, error_queues, return_queues
        self.sentinels = {p.sentinel: i for i, p in enumerate(processes)}

    def join(self, timeout=None):
        if not self.sentinels:
            return True
        ready = multiprocessing.connection.wait(self.sentinels.keys(), timeout=timeout)
        failed = None
        for s in ready:
            i = self.sentinels.pop(s)
            p = self.processes[i]
            p.join()
            if p.exitcode != 0:
                failed = i
                break
        if failed is None:
            return len(self.sentinels) == 0
        for p in self.processes:
            if p.is_alive():
                p.terminate()
            p.join()
        q = self.error_queues[failed]
        if q.empty():
            raise Exception(f"Process {failed} terminated with exit code {self.processes[failed].exitcode}.")
        raise Exception(f"\n\n----------------------------------------------\nProcess {failed} terminated with the "
                        f"following error:\n----------------------------------------------\n\n{q.get()}")


def spawn(func, args=(), nprocs=-1, join=True, daemon=False, **options):
    import multiprocessing.connection  # noqa: F401
    import torch
    if nprocs == -1:
        nprocs = torch.cuda.device_count() if torch.cuda.is_available() else 1
    if "backend" not in options or options["backend"] in (None, "auto"):
        options["backend"] = "nccl" if torch.cuda.is_available() and nprocs <= max(torch.cuda.device_count(), 0) else "gloo"
    if options.get("ips") not in (None, "127.0.0.1", "localhost"):
        raise NotImplementedError("spawn starts the processes of one node; use paddle.distributed.launch for multi-node")
    start_port = options.get("start_port") or _free_port()
    endpoints = [f"127.0.0.1:{start_port}"] + [f"127.0.0.1:{_free_port()}" for _ in range(nprocs - 1)]
    ctx = multiprocessing.get_context("spawn")
    procs, eqs, rqs = [], [], []
    for r in range(nprocs):
        eq, rq = ctx.SimpleQueue(), ctx.SimpleQueue()
        p = ctx.Process(target=_func_wrapper, args=(func, args, eq, rq, _env_for(r, nprocs, endpoints, options)), daemon=daemon)
        p.start()
        procs.append(p)
        eqs.append(eq)
        rqs.append(rq)
    context = MultiprocessContext(procs, eqs, rqs)
    if not join:
        return context
    while not context.join():
        pass
    return context
