#!/usr/bin/env python
"""Headline benchmark: Llama-2 7B training tokens/s on N B200s of one node (BASELINE.json).

    python bench.py --gpus 1 --steps 8 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \
        bench.py --gpus 8 --steps 8 --warmup 3

N = 1: bf16 parameters (AMP O2) + fp32 master weights in the fused AdamW.  N > 1: GroupSharded stage 3
(``group_sharded_parallel(level="p_g_os")``), weak scaling (per-GPU batch fixed).  Synthetic tokens, random-init
weights.  Prints ONE JSON line on rank 0.

Timing: W untimed warm-up steps, then exactly K steps between a barrier + torch.cuda.synchronize() on both sides,
timed with CUDA events on the compute stream, max over ranks (``value`` - a ring of distinct batches resident on the
device).  A second timed block of K steps goes through the public API end to end: every step copies that step's
tokens from pinned host memory and reads the loss back (``e2e``).  At N > 1 a third block repeats the K steps with
the stage-3 transport kernels skipped (garbage numerics, same compute): ``exposed_comm_ms`` = step time minus that.
Every rank samples its own GPU's clocks / throttle reasons with nvidia-smi during the timed regions.  During warm-up
at N > 1 the fused collectives are checked once against NCCL (``selfcheck``).  L2: every step streams > 100 GB of
parameters, gradients and optimizer state, far beyond the 126 MB L2.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASELINE_TOKENS_PER_S = None   # the reference publishes no end-to-end number


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=6)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--model", default="llama2-7b", choices=["llama2-7b", "llama2-13b", "llama-tiny", "llama-1b"])
    p.add_argument("--batch", type=int, default=int(os.environ.get("BENCH_BATCH", "2")), help="sequences per GPU per step")
    p.add_argument("--seq", type=int, default=4096)
    p.add_argument("--recompute", type=int, default=int(os.environ.get("BENCH_RECOMPUTE", "0")))
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-exposed", action="store_true", help="skip the comm-skipped timing block at N > 1")
    p.add_argument("--library-baseline", action="store_true",
                   help="same model and step through cuBLAS / cuDNN / torch kernels and NCCL only (this repo's sm_100a "
                        "kernels and fused NVLink paths switched off): the 'Baseline B' column of BASELINE.md")
    return p.parse_args()


class ClockSampler:
    """nvidia-smi sampling of SM clock / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def reference_arm(args):
    """Run the unmodified reference from baseline/_ref when it is installed there; otherwise say why not.  The install
    attempts (pip, pip --no-deps, cmake configure) are scripted in baseline/build_ref.sh and their outcome is recorded
    in baseline/build_log_excerpt.txt and DESIGN.md section 5."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "paddle")):
        why = "pip install of /root/reference fails offline: setup.py check_build_dependency -> missing opt_einsum " \
              "(not in /opt/wheelhouse); the source tree needs a full cmake/CUDA build (see baseline/build_log_excerpt.txt)"
        print(json.dumps({"impl": "reference", "unavailable": why}))
        return 0
    sys.path.insert(0, ref)
    try:
        import paddle  # noqa: F401  (the reference's own package)
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"impl": "reference", "unavailable": f"baseline/_ref does not import: {type(e).__name__}: {e}"[:300]}))
        return 0
    from baseline import ref_bench
    return ref_bench.main(args)


def merge_clocks(per_rank):
    rows = [c for c in per_rank if c and c.get("sm_mhz") is not None]
    if not rows:
        return per_rank[0] if per_rank else None
    reasons = sorted({r for c in rows for r in c.get("reasons", [])})
    return {"sm_mhz": min(c["sm_mhz"] for c in rows), "sm_max_mhz": max(c["sm_max_mhz"] for c in rows), "reasons": reasons,
            "samples": sum(c.get("samples", 0) for c in rows), "note": "sm_mhz = lowest per-rank median under load"}


def loaded_native_libs():
    try:
        with open("/proc/self/maps") as fh:
            return sorted({os.path.basename(l.split()[-1]) for l in fh if "libpdb200" in l})
    except OSError:
        return []


def run_selfcheck(world, dev):
    """Fused collectives against NCCL on random data (max abs error), once, outside the timed regions."""
    import torch
    import paddle_b200.distributed as dist
    from paddle_b200.parallel import symm
    g = dist.collective._get_default_group()
    if not symm.available(g):
        return {"skipped": "symmetric-memory paths unavailable"}
    out = {}
    try:
        arena = symm.get_arena(g, 96 << 20, key="selfcheck")
        n, r = g.nranks, g.rank
        torch.manual_seed(77 + r)
        numel = 1 << 20
        # K1 scaled all-reduce
        t = arena.empty((numel,), torch.float32)
        src = torch.randn(numel, device=dev)
        t.copy_(src)
        ref = src.clone()
        dist.all_reduce(ref)
        symm.allreduce_(t, scale=1.0 / n)
        out["k1_allreduce_max_abs_err"] = float((t - ref / n).abs().max())
        # K3 unit all-gather
        shard = arena.empty((numel,), torch.bfloat16)
        full = arena.empty((numel * n,), torch.bfloat16)
        shard.copy_(torch.randn(numel, device=dev).bfloat16())
        ref_full = torch.empty(numel * n, dtype=torch.bfloat16, device=dev)
        dist.all_gather(ref_full, shard.clone())
        symm.allgather_multi(arena, [(shard._symm[1], full._symm[1], numel * 2)])
        out["k3_allgather_max_abs_err"] = float((full.float() - ref_full.float()).abs().max())
        # K2 pass 1: unit reduce-scatter (AVG) into fp32
        stage = arena.empty((numel * n,), torch.bfloat16)
        gsrc = torch.randn(numel * n, device=dev).bfloat16()
        stage.copy_(gsrc)
        acc = torch.zeros(numel, dtype=torch.float32, device=dev)
        ref_rs = torch.empty(numel, dtype=torch.float32, device=dev)
        dist.reduce_scatter(ref_rs, gsrc.float())
        symm.reduce_scatter_multi(arena, [(stage._symm[1], numel, acc.data_ptr())], torch.bfloat16, 1.0 / n)
        out["k2_reduce_scatter_max_abs_err"] = float((acc - ref_rs / n).abs().max())
        out["nvls"] = bool(arena.mc_ptr)
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        out["error"] = f"{type(e).__name__}: {e}"[:300]
    return out


def build_config(name):
    from paddle_b200.models import LlamaConfig
    if name == "llama2-7b":
        return LlamaConfig.llama2_7b()
    if name == "llama2-13b":
        return LlamaConfig.llama2_13b()
    if name == "llama-1b":
        return LlamaConfig(hidden_size=2048, intermediate_size=5504, num_hidden_layers=16, num_attention_heads=16)
    return LlamaConfig.tiny(max_position_embeddings=4096)


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import paddle_b200 as paddle
    from paddle_b200 import lib
    from paddle_b200.models import LlamaForCausalLM

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch ourselves under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29533"),
                   os.path.abspath(__file__)] + sys.argv[1:]
            return subprocess.call(cmd)
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); on the CPU box use the tests")
    torch.cuda.set_device(local_rank)
    os.environ.setdefault("FLAGS_selected_gpus", str(local_rank))
    paddle.set_device(f"gpu:{local_rank}")
    dev = torch.device("cuda", local_rank)

    if world > 1:
        paddle.distributed.init_parallel_env()

    if args.library_baseline:
        paddle.set_flags({"FLAGS_use_native_kernels": False, "FLAGS_use_fused_comm": False, "FLAGS_use_tcgen05_gemm": False})
    cfg = build_config(args.model)
    cfg.recompute = bool(args.recompute)
    paddle.seed(1234)
    with paddle.utils.unique_name.guard():
        paddle.set_default_dtype("bfloat16")
        try:
            model = LlamaForCausalLM(cfg)        # parameters created directly in bf16 on the device
        finally:
            paddle.set_default_dtype("float32")
    opt = paddle.optimizer.AdamW(learning_rate=1e-5, beta1=0.9, beta2=0.95, epsilon=1e-8, weight_decay=0.1,
                                 parameters=model.parameters(), multi_precision=True,
                                 grad_clip=paddle.nn.ClipGradByGlobalNorm(1.0))
    parallelism = "single"
    if world > 1:
        from paddle_b200.distributed.sharding import group_sharded_parallel
        model, opt, _ = group_sharded_parallel(model, opt, level="p_g_os")
        parallelism = f"sharding_stage3_x{world}"
    model.train()

    B, S, V = args.batch, args.seq, cfg.vocab_size
    gen = torch.Generator().manual_seed(1000 + rank)
    n_host = args.steps + args.warmup + 2
    host_tokens = [torch.randint(0, V, (B, S + 1), generator=gen, dtype=torch.int64).pin_memory() for _ in range(n_host)]
    dev_ring = [t.to(dev) for t in host_tokens[:min(8, n_host)]]      # distinct resident batches for the device-timed block

    def train_step(tokens):
        loss = model(tokens[:, :-1], labels=tokens[:, 1:])
        loss.backward()
        opt.step()
        opt.clear_grad(set_to_zero=False)      # grads are re-created by autograd: no zero-fill, no accumulate pass
        return loss

    def barrier():
        if world > 1:
            paddle.distributed.barrier()
        torch.cuda.synchronize()

    losses = []
    for i in range(args.warmup):
        loss = train_step(dev_ring[i % len(dev_ring)])
        losses.append(loss.detach())
    selfcheck = run_selfcheck(world, dev) if world > 1 else None
    barrier()
    lib.reset_launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        loss = train_step(dev_ring[(args.warmup + i) % len(dev_ring)])
        losses.append(loss.detach())
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = lib.launch_count()
    losses = [float(l) for l in losses]
    final_loss = losses[-1]

    # ---- end to end through the public API: pinned H2D of the step's tokens + D2H of the loss every step
    e2e = None
    if not args.no_e2e:
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for i in range(args.steps):
            tok = host_tokens[1 + i].to(dev, non_blocking=True)
            loss = train_step(tok)
            _ = loss.item()
        f1.record()
        barrier()
        e2e = f0.elapsed_time(f1)
    clocks = sampler.stop()

    # ---- exposed communication: the same K steps with the transport kernels skipped (run last: it ruins the weights)
    nocomm_ms = 0.0
    transport = "none"
    if world > 1:
        nvl = getattr(model, "_nvl", None)
        transport = "nccl"
        if nvl is not None:
            transport = (f"nvlink-symmetric arena: all-gather={'copy engines' if nvl.ag_mode == 'ce' else 'multimem/peer-store kernel'}, "
                         f"reduce-scatter={'copy-engine all-to-all + local sum kernel' if nvl.rs_mode == 'ce' else 'multimem/peer-load kernel'}")
        if nvl is not None and not args.no_exposed:
            nvl.noop = True
            train_step(dev_ring[0])
            barrier()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record()
            for i in range(args.steps):
                train_step(dev_ring[i % len(dev_ring)])
            g1.record()
            barrier()
            nocomm_ms = g0.elapsed_time(g1)
            nvl.noop = False

    t = torch.tensor([ms, e2e if e2e is not None else 0.0, nocomm_ms], dtype=torch.float64, device=dev)
    if world > 1:
        paddle.distributed.all_reduce(t, paddle.distributed.ReduceOp.MAX)
    ms, e2e_ms, nocomm_ms = t.tolist()
    all_clocks = [clocks]
    if world > 1:
        all_clocks = [None] * world
        paddle.distributed.all_gather_object(all_clocks, clocks)
    if rank == 0:
        tokens_per_step = B * S * world
        ms_per_step = ms / args.steps
        value = tokens_per_step / (ms_per_step / 1e3)
        flops = cfg.flops_per_token(S) * tokens_per_step / (ms_per_step / 1e3)
        out = {
            "metric": "Llama-2 7B training tokens/s (whole job)" if args.model == "llama2-7b" else f"{args.model} training tokens/s",
            "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms_per_step, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / BASELINE_TOKENS_PER_S) if BASELINE_TOKENS_PER_S else None,
            "dtype": "bf16", "data": "synthetic tokens, random-init weights",
            "impl": "ours" if not args.library_baseline else "library-baseline (cuBLAS/cuDNN/torch + NCCL, our kernels off)",
            "config": {"model": args.model, "global_batch": B * world, "seq_len": S, "parallelism": parallelism,
                       "optimizer": "AdamW multi_precision (fp32 master, fp32 moments) + global-norm clip",
                       "recompute": bool(args.recompute),
                       "l2": "no explicit flush: each step streams >100 GB of params/grads/optimizer state (L2 is 126 MB)"},
            "model_tflops_per_gpu": round(flops / world / 1e12, 1),
            "final_loss": round(final_loss, 4),
            "loss_first_last": [round(losses[0], 4), round(losses[-1], 4)],
            "gpu_launches": launches,
            "clocks": merge_clocks(all_clocks),
            "clocks_per_rank": all_clocks,
            "native_so_loaded": loaded_native_libs(),
            "transport": transport,
        }
        if world > 1 and nocomm_ms > 0:
            out["exposed_comm_ms"] = round(ms_per_step - nocomm_ms / args.steps, 3)
            out["exposed_comm_method"] = "ms_per_step minus ms_per_step of the same steps with the stage-3 transport kernels skipped"
            out["ms_per_step_comm_skipped"] = round(nocomm_ms / args.steps, 3)
        if selfcheck is not None:
            out["selfcheck"] = selfcheck
        if e2e is not None:
            e2e_per = e2e_ms / args.steps
            out["e2e"] = {"value": round(tokens_per_step / (e2e_per / 1e3), 1), "unit": "tokens/s",
                          "ms_per_step": round(e2e_per, 3),
                          "h2d_bytes_per_step": B * (S + 1) * 8, "d2h_bytes_per_step": 4}
            out["e2e_check"] = {"copies_declared": True, "h2d_inside_timed_region": True, "d2h_inside_timed_region": True,
                                "input_source": "pinned host memory, a different batch every step",
                                "result_read": "loss.item() every step", "api": "model(tokens, labels) / loss.backward() / opt.step()"}
        print(json.dumps(out))
    if world > 1:
        paddle.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
