"""The other BASELINE.json configurations, one JSON line each (device-timed with CUDA events, max over ranks):

  resnet50_dp      ResNet-50 DataParallel AMP bf16, synthetic 224x224 images (fused bucketed gradient all-reduce, K1)
  llama13b_hybrid  Llama-2 13B  mp=4 + sequence parallel  x  pp=2 1F1B, VocabParallelEmbedding + ParallelCrossEntropy
  moe_ep           Mixtral-style MoE (8 experts, top-2) expert parallel over all ranks (K9/K10 dispatch / combine)

launch:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
             bench/configs_bench.py --config moe_ep [--steps 4 --warmup 3 --layers L --batch B]
`--layers` shrinks the DEPTH of the named model (stated in the output) so a config fits the time/memory budget of a
test box; widths, heads, sequence length and parallel layout stay those of the named configuration."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_b200 as paddle  # noqa: E402
import paddle_b200.distributed as dist  # noqa: E402
from paddle_b200 import lib  # noqa: E402
from paddle_b200.distributed import fleet  # noqa: E402


def timed(step, steps, warmup, dev):
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    lib.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, dist.ReduceOp.MAX)
    return float(t) / steps, out, lib.launch_count()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["resnet50_dp", "llama13b_hybrid", "moe_ep"])
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--seq", type=int, default=4096)
    ap.add_argument("--library-baseline", action="store_true",
                    help="cuBLAS / cuDNN / torch kernels + NCCL only (Baseline B of BASELINE.md)")
    args = ap.parse_args()
    if args.library_baseline:
        paddle.set_flags({"FLAGS_use_native_kernels": False, "FLAGS_use_fused_comm": False, "FLAGS_use_tcgen05_gemm": False})
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    res = {"impl": "library-baseline" if args.library_baseline else "ours", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "dtype": "bf16",
           "data": "synthetic, random-init weights", "higher_is_better": True}

    if args.config == "resnet50_dp":
        dist.init_parallel_env()
        from paddle_b200.vision.models import resnet50
        B = args.batch or 256
        paddle.seed(1)
        model = resnet50().to(dev)
        torch.nn.Module.to(model, memory_format=torch.channels_last)
        opt = paddle.optimizer.Momentum(0.1, momentum=0.9, parameters=model.parameters(), weight_decay=1e-4)
        dp = paddle.DataParallel(model) if world > 1 else model
        x = torch.randn(B, 3, 224, 224, device=dev).to(memory_format=torch.channels_last)
        y = torch.randint(0, 1000, (B,), device=dev)
        ce = paddle.nn.CrossEntropyLoss()

        def step():
            with paddle.amp.auto_cast(dtype="bfloat16"):
                loss = ce(dp(x), y)
            loss.backward()
            opt.step()
            opt.clear_grad(set_to_zero=False)
            return loss
        ms, loss, launches = timed(step, args.steps, args.warmup, dev)
        res.update(metric="ResNet-50 training images/s (whole job)", value=round(B * world / ms * 1e3, 1), unit="images/s",
                   config={"model": "resnet50", "global_batch": B * world, "parallelism": f"dp{world}", "amp": "O1 bf16",
                           "fused_allreduce": bool(world > 1 and getattr(dp._reducer, "_fused", None) is not None)})

    elif args.config == "llama13b_hybrid":
        from paddle_b200.models.llama import LlamaConfig, LlamaForCausalLMPipe
        mp = 4 if world % 4 == 0 else (2 if world % 2 == 0 and world > 2 else 1)
        pp = 2 if world // mp >= 2 else 1
        dpd = world // (mp * pp)
        acc = 4 * pp
        mbs = args.batch or 1
        st = fleet.DistributedStrategy()
        st.hybrid_configs = {"dp_degree": dpd, "mp_degree": mp, "pp_degree": pp,
                             "mp_configs": {"sync_param": False}, "pp_configs": {}}
        st.pipeline_configs = {"accumulate_steps": acc, "micro_batch_size": mbs}
        fleet.init(is_collective=True, strategy=st)
        cfg = LlamaConfig.llama2_13b() if hasattr(LlamaConfig, "llama2_13b") else LlamaConfig(
            hidden_size=5120, intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40)
        if args.layers:
            cfg.num_hidden_layers = args.layers
        cfg.tensor_parallel_degree = mp
        cfg.sequence_parallel = mp > 1
        paddle.seed(1234)
        paddle.set_default_dtype("bfloat16")
        try:
            with paddle.utils.unique_name.guard():
                pipe = LlamaForCausalLMPipe(cfg, num_stages=pp)
        finally:
            paddle.set_default_dtype("float32")
        opt = paddle.optimizer.AdamW(1e-5, beta1=0.9, beta2=0.95, weight_decay=0.1, parameters=pipe.parameters(),
                                     multi_precision=True, grad_clip=paddle.nn.ClipGradByGlobalNorm(1.0))
        model = fleet.distributed_model(pipe)
        opt = fleet.distributed_optimizer(opt)
        S, V = args.seq, cfg.vocab_size
        g = torch.Generator().manual_seed(7)
        tok = torch.randint(0, V, (acc * mbs, S + 1), generator=g).to(dev)

        def step():
            return model.train_batch([tok[:, :-1].contiguous(), tok[:, 1:].contiguous()], opt)
        ms, loss, launches = timed(step, args.steps, args.warmup, dev)
        tokens = acc * mbs * S * dpd
        res.update(metric="Llama-2 13B hybrid-parallel training tokens/s (whole job)", value=round(tokens / ms * 1e3, 1),
                   unit="tokens/s",
                   config={"model": "llama2-13b", "layers": cfg.num_hidden_layers, "global_batch": acc * mbs * dpd, "seq_len": S,
                           "parallelism": f"mp{mp}(sp)xpp{pp}(1F1B)xdp{dpd}", "accumulate_steps": acc})

    else:  # moe_ep
        dist.init_parallel_env()
        from paddle_b200.models.moe import MoEConfig, MoEForCausalLM
        g = dist.collective._get_default_group()
        cfg = MoEConfig.mixtral_8x7b()
        cfg.num_hidden_layers = args.layers or 8
        if os.environ.get("PDB200_MOE_GROUPED"):          # 1 / 0 force the grouped GEMM path on / off (default: automatic)
            cfg.grouped_experts = os.environ["PDB200_MOE_GROUPED"] != "0"
        B, S = args.batch or 1, args.seq
        paddle.seed(1234 + dist.get_rank())
        paddle.set_default_dtype("bfloat16")
        try:
            with paddle.utils.unique_name.guard():
                model = MoEForCausalLM(cfg, moe_group=g if world > 1 else None)
        finally:
            paddle.set_default_dtype("float32")
        model.to(dev)
        opt = paddle.optimizer.AdamW(1e-5, beta1=0.9, beta2=0.95, weight_decay=0.1, parameters=model.parameters(),
                                     multi_precision=True)
        dp = paddle.DataParallel(model) if world > 1 else model   # dense parts data-parallel, experts are no_sync
        gen = torch.Generator().manual_seed(100 + dist.get_rank())
        tok = torch.randint(0, cfg.vocab_size, (B, S + 1), generator=gen).to(dev)

        def step():
            loss = dp(tok[:, :-1], labels=tok[:, 1:])
            loss.backward()
            opt.step()
            opt.clear_grad(set_to_zero=False)
            return loss
        ms, loss, launches = timed(step, args.steps, args.warmup, dev)
        from paddle_b200.parallel import moe_fused
        res.update(metric="Mixtral-style MoE (8 experts, top-2) training tokens/s (whole job)",
                   value=round(B * S * world / ms * 1e3, 1), unit="tokens/s",
                   config={"model": "mixtral-8x7b widths", "layers": cfg.num_hidden_layers, "global_batch": B * world, "seq_len": S,
                           "parallelism": f"ep{world} + dp{world} (dense)", "experts": cfg.num_experts, "top_k": cfg.top_k,
                           "fused_dispatch_combine": bool(moe_fused._state), "grouped_expert_gemm": cfg.grouped_experts if cfg.grouped_experts is not None else cfg.num_experts // max(world, 1) > 1})

    res.update(ms_per_step=round(ms, 3), final_loss=round(float(loss), 4), gpu_launches=launches)
    if dist.get_rank() == 0:
        print(json.dumps(res), flush=True)
    from paddle_b200.parallel import moe_fused, symm
    moe_fused.release_all()
    symm.release_all()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
