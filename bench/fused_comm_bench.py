"""Fused NVLink paths vs their NCCL(+cuBLAS) compositions, device-timed (CUDA events, max over ranks), with the
achieved fraction of each path's roofline = slower of (FLOPs / measured GEMM peak) and (bytes that must cross NVLink per
GPU / 770 GB/s measured peer bandwidth per direction).

launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
            bench/fused_comm_bench.py > gpurun_out/fused_comm_nN.json
Rows: K1 scaled all-reduce (DP bucket), K2+K3 reduce-scatter + AdamW + all-gather, K5 all-gather -> GEMM,
K6 GEMM -> reduce-scatter, K4 GEMM -> all-reduce, K8 vocab-parallel CE statistics exchange."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_b200 as paddle  # noqa: E402
import paddle_b200.distributed as dist  # noqa: E402
from paddle_b200.parallel import symm  # noqa: E402

NVLINK_GBS = 770.0
PEAKS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "hbm_gbs": 6650.0}


def timeit(fn, dev, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
    dist.all_reduce(t, dist.ReduceOp.MAX)
    return float(t)


def main():
    dist.init_parallel_env()
    r, n = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device())
    g = dist.collective._get_default_group()
    assert symm.available(g)
    arena = symm.get_arena(g, 3 << 30, key="bench")
    rows = []

    def add(name, fused_ms, nccl_ms, flops, link_bytes, hbm_bytes=0.0, note=""):
        roof_ms = max(flops / (PEAKS["bf16_tflops"] * 1e12), link_bytes / (NVLINK_GBS * 1e9),
                      hbm_bytes / (PEAKS["hbm_gbs"] * 1e9)) * 1e3
        rows.append(dict(path=name, n_gpus=n, fused_ms=round(fused_ms, 4), nccl_ms=round(nccl_ms, 4),
                         speedup_vs_nccl=round(nccl_ms / fused_ms, 2), roofline_ms=round(roof_ms, 4),
                         fraction_of_roofline=round(roof_ms / fused_ms, 3), note=note))
        if r == 0:
            print(json.dumps(rows[-1]), flush=True)

    # ---- K1: DP bucket all-reduce with the 1/n scale folded in
    for dtype, numel, tag in ((torch.float32, 25 * (1 << 20) // 4, "25 MB fp32 bucket"), (torch.bfloat16, 128 << 20, "256 MB bf16")):
        t = arena.empty((numel,), dtype)
        t.normal_()
        ref = torch.randn(numel, device=dev).to(dtype)

        def nccl():
            ref.mul_(1.0 / n)
            dist.all_reduce(ref)
        nbytes = numel * t.element_size()
        add(f"K1 all-reduce+scale {tag}", timeit(lambda: symm.allreduce_(t, scale=1.0 / n), dev), timeit(nccl, dev), 0.0,
            nbytes * (n - 1) / n * (1.0 if arena.mc_ptr else 2.0), note="NVLS: 1x slice in + 1x slice out per GPU" if arena.mc_ptr else "P2P")

    # ---- K2 + K3: reduce-scatter + AdamW + all-gather on a flat 256 MB bf16 buffer
    total = (128 << 20) // n // 8 * 8 * n
    S = total // n
    gbuf, pbuf = arena.empty((total,), torch.bfloat16), arena.empty((total,), torch.bfloat16)
    gbuf.normal_()
    pbuf.normal_()
    master, m, v = (torch.zeros(S, device=dev) for _ in range(3))
    pows = torch.tensor([0.9, 0.95], device=dev)
    g2, p2 = gbuf.clone(), pbuf.clone()
    shard_g = torch.empty(S, dtype=torch.bfloat16, device=dev)
    p_shard = torch.nn.Parameter(torch.zeros(S, device=dev))
    ref_opt = torch.optim.AdamW([p_shard], lr=1e-3, fused=True)

    def nccl_zero():
        dist.reduce_scatter(shard_g, g2, dist.ReduceOp.AVG)
        p_shard.grad = shard_g.float()
        ref_opt.step()
        dist.all_gather(p2, p_shard.detach().bfloat16())
    add("K2+K3 reduce-scatter+AdamW+all-gather 256 MB bf16",
        timeit(lambda: symm.rs_adamw_ag(gbuf, pbuf, r * S, S, master, m, v, 1e-3, 0.9, 0.95, 1e-8, 0.1, pows), dev),
        timeit(nccl_zero, dev), 0.0, 2.0 * total * 2 * (n - 1) / n / (2.0 if arena.mc_ptr else 1.0) * 1.0,
        hbm_bytes=S * (2 + 2 + 6 * 4), note="one kernel; torch fused AdamW on the NCCL side")

    # ---- K5 / K6 / K4 at the Llama-2 13B sequence-parallel shapes (seq 4096, micro-batch 1, mp = n)
    tokens, H, FF = 4096, 5120, 13824
    rows_per = tokens // n
    xs = (torch.randn(rows_per, H, device=dev) * 0.1).bfloat16()
    w_up = (torch.randn(H, 2 * FF // n, device=dev) * 0.02).bfloat16()
    w_down = (torch.randn(FF // n, H, device=dev) * 0.02).bfloat16()
    xf = (torch.randn(tokens, FF // n, device=dev) * 0.1).bfloat16()
    gathered = torch.empty(tokens, H, dtype=torch.bfloat16, device=dev)
    rs_out = torch.empty(rows_per, H, dtype=torch.bfloat16, device=dev)

    def nccl_ag_gemm():
        dist.all_gather(gathered, xs)
        return torch.matmul(gathered, w_up)

    def nccl_gemm_rs():
        dist.reduce_scatter(rs_out, torch.matmul(xf, w_down))

    def nccl_gemm_ar():
        y = torch.matmul(xf, w_down)
        dist.all_reduce(y)
        return y
    fl_up = 2.0 * tokens * H * (2 * FF // n)
    fl_down = 2.0 * tokens * (FF // n) * H
    ag_bytes = tokens * H * 2 * (n - 1) / n
    t_push = timeit(lambda: symm.ag_gemm(xs, w_up, g), dev)
    t_n = timeit(nccl_ag_gemm, dev)
    add("K5 all-gather->GEMM gate_up (NVLS multimem.st all-gather into local memory + CTA-pair GEMM)", t_push, t_n, fl_up, ag_bytes)
    t_rs_n, t_ar_n = timeit(nccl_gemm_rs, dev), timeit(nccl_gemm_ar, dev)
    add("K6 GEMM->reduce-scatter down_proj (CTA-pair GEMM -> local symmetric buffer, NVLS pull-reduce)",
        timeit(lambda: symm.gemm_rs(xf, w_down, g), dev), t_rs_n, fl_down, tokens * H * 2 * (n - 1) / n)
    add("K4 GEMM->all-reduce down_proj (CTA-pair GEMM -> local symmetric buffer, NVLS two-shot)",
        timeit(lambda: symm.gemm_ar(xf, w_down, g), dev), t_ar_n, fl_down, 2.0 * tokens * H * 2 * (n - 1) / n)

    # ---- K8 vocab-parallel CE (13B: V = 32000, N = 4096 rows)
    from paddle_b200.distributed.fleet.layers.mpu import mp_ops
    V = 32000 // n // 8 * 8
    logits = (torch.randn(tokens, V, device=dev)).bfloat16().requires_grad_(True)
    labels = torch.randint(0, V * n, (tokens, 1), device=dev)

    def ce():
        logits.grad = None
        mp_ops._c_softmax_with_cross_entropy(logits, labels, group=g).sum().backward()
    paddle.set_flags({"FLAGS_use_fused_comm": True})
    t_f = timeit(ce, dev)
    paddle.set_flags({"FLAGS_use_fused_comm": False})
    t_nc = timeit(ce, dev)
    paddle.set_flags({"FLAGS_use_fused_comm": True})
    add("K8 vocab-parallel CE fwd+bwd (3 floats/row exchanged in-kernel)", t_f, t_nc, 0.0, tokens * 12.0 * (n - 1),
        hbm_bytes=tokens * V * 2 * 3.0, note="NCCL side: same local kernels, statistics via all_gather")
    if r == 0:
        json.dump(rows, open(os.path.join(os.path.dirname(__file__), "..", "gpurun_out", f"fused_comm_n{n}.json"), "w"), indent=1)
    symm.release_all()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
