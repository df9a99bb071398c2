"""Kernel timeline of one stage-3 training step (torch.profiler / CUPTI), summarised on the box:
per-stream busy time, idle gaps on the compute stream, duration of every transport kernel, and the total time of the
GEMM / attention kernels - with and without the transport kernels (PDB200_STAGE3_COMM_NOOP) so a slowdown of the
compute kernels under communication is visible.  Launch with torchrun like bench.py; rank 0 prints one JSON line per mode.
    python -m torch.distributed.run --nproc-per-node 2 ... tools/trace_stage3.py [--layers 8]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_b200 as paddle  # noqa: E402
from paddle_b200.models import LlamaConfig, LlamaForCausalLM  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=8)
ap.add_argument("--batch", type=int, default=2)
args = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
paddle.set_device(f"gpu:{local}")
dev = torch.device("cuda", local)
if world > 1:
    paddle.distributed.init_parallel_env()
cfg = LlamaConfig.llama2_7b()
cfg.num_hidden_layers = args.layers
paddle.seed(1)
paddle.set_default_dtype("bfloat16")
model = LlamaForCausalLM(cfg)
paddle.set_default_dtype("float32")
opt = paddle.optimizer.AdamW(1e-5, parameters=model.parameters(), multi_precision=True, weight_decay=0.1,
                             grad_clip=paddle.nn.ClipGradByGlobalNorm(1.0))
if world > 1:
    from paddle_b200.distributed.sharding import group_sharded_parallel
    model, opt, _ = group_sharded_parallel(model, opt, level="p_g_os")
tok = torch.randint(0, cfg.vocab_size, (args.batch, 4097), device=dev)


def step():
    loss = model(tok[:, :-1], labels=tok[:, 1:])
    loss.backward()
    opt.step()
    opt.clear_grad(set_to_zero=False)


def summarise(prof, tag):
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time_total > 0]
    by_name, by_stream = {}, {}
    rows = []
    for e in evs:
        t0 = e.time_range.start
        dur = e.device_time_total
        n = e.name[:60]
        a = by_name.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += dur
        rows.append((t0, dur, n, getattr(e, "device_resource_id", 0)))
    rows.sort()
    for t0, dur, n, sid in rows:
        s = by_stream.setdefault(sid, {"busy_us": 0.0, "n": 0, "first": t0, "last": t0 + dur})
        s["busy_us"] += dur
        s["n"] += 1
        s["last"] = max(s["last"], t0 + dur)
    main = max(by_stream, key=lambda k: by_stream[k]["busy_us"]) if by_stream else None
    gaps, prev_end = [], None
    for t0, dur, n, sid in rows:
        if sid != main:
            continue
        if prev_end is not None and t0 - prev_end > 15:
            gaps.append((round(t0 - prev_end, 1), n))
        prev_end = max(prev_end or 0, t0 + dur)
    top = sorted(by_name.items(), key=lambda kv: -kv[1][1])[:14]
    comm = [(n, round(d, 1)) for t0, d, n, sid in rows if "multi_kernel" in n or "nccl" in n.lower()]
    out = {"mode": tag, "streams": {str(k): {"busy_ms": round(v["busy_us"] / 1e3, 2), "n": v["n"], "span_ms": round((v["last"] - v["first"]) / 1e3, 2)} for k, v in by_stream.items()},
           "compute_stream_gaps_ms": round(sum(g for g, _ in gaps) / 1e3, 2), "n_gaps": len(gaps),
           "biggest_gaps": sorted(gaps, reverse=True)[:8],
           "top_kernels_ms": [(n, c, round(t / 1e3, 2)) for n, (c, t) in top],
           "comm_kernels_us": comm[:40], "comm_total_ms": round(sum(d for _, d in comm) / 1e3, 2)}
    if rank == 0:
        print(json.dumps(out), flush=True)


for mode in ("comm", "noop"):
    nvl = getattr(model, "_nvl", None)
    if nvl is not None:
        nvl.noop = "" if mode == "comm" else "all"
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        paddle.distributed.barrier()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    summarise(prof, mode)
if world > 1:
    paddle.distributed.destroy_process_group()
