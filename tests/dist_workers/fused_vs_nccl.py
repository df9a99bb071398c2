"""Worker (>= 2 GPUs): one training config, transport chosen by FLAGS_use_fused_comm (env); final parameters are
written to <out>.<rank>.pt so the driver can compare the fused NVLink run with the NCCL run."""
import sys

import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200 import nn
from paddle_b200.distributed import fleet
from paddle_b200.framework.flags import flag

mode, out = sys.argv[1], sys.argv[2]
H = 512


def build():
    paddle.seed(11)
    paddle.set_default_dtype("bfloat16")
    try:
        with paddle.utils.unique_name.guard():
            return nn.Sequential(nn.Linear(H, 4 * H, bias_attr=False), nn.GELU(), nn.Linear(4 * H, H, bias_attr=False))
    finally:
        paddle.set_default_dtype("float32")


def data(step, rank, dev):
    g = torch.Generator().manual_seed(100 * step + rank)
    return torch.randn(64, H, generator=g).to(dev).bfloat16(), torch.randn(64, H, generator=g).to(dev).bfloat16()


if mode in ("tp", "sp"):
    # K4 / K5 / K6 / K7 / K8 inside the TP and sequence-parallel layers
    from paddle_b200.distributed.fleet.layers.mpu import mp_layers
    from paddle_b200.distributed.fleet.utils import sequence_parallel_utils as spu
    n = dist.ParallelEnv().world_size
    st = fleet.DistributedStrategy()
    st.hybrid_configs = {"dp_degree": 1, "mp_degree": n, "pp_degree": 1, "sharding_degree": 1}
    fleet.init(is_collective=True, strategy=st)
    r = dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    V = 1024

    class Block(nn.Layer):
        def __init__(self):
            super().__init__()
            self.emb = mp_layers.VocabParallelEmbedding(V, H)
            if mode == "tp":
                self.up = mp_layers.ColumnParallelLinear(H, 4 * H, has_bias=False, gather_output=False)
                self.down = mp_layers.RowParallelLinear(4 * H, H, has_bias=False, input_is_parallel=True)
            else:
                self.up = spu.ColumnSequenceParallelLinear(H, 4 * H, has_bias=False, gather_output=False)
                self.down = spu.RowSequenceParallelLinear(4 * H, H, has_bias=False, input_is_parallel=True)
            self.head = mp_layers.ColumnParallelLinear(H, V, has_bias=False, gather_output=False)
            self.ce = mp_layers.ParallelCrossEntropy()

        def forward(self, ids, labels):
            h = self.emb(ids)                                  # [b, s, H]
            if mode == "sp":
                h = spu.ScatterOp.apply(h.transpose([1, 0, 2]).contiguous())   # [s/n, b, H]
            h = h + self.down(torch.nn.functional.gelu(self.up(h)))
            if mode == "sp":
                h = spu.GatherOp.apply(h).transpose([1, 0, 2])
            return self.ce(self.head(h), labels).mean()

    paddle.seed(11)
    paddle.set_default_dtype("bfloat16")
    with paddle.utils.unique_name.guard():
        m = Block()
    paddle.set_default_dtype("float32")
    opt = paddle.optimizer.AdamW(1e-3, parameters=m.parameters(), multi_precision=True)
    model = fleet.distributed_model(m)
    opt = fleet.distributed_optimizer(opt)
    for s in range(4):
        gen = torch.Generator().manual_seed(500 + s)
        ids = torch.randint(0, V, (4, 128), generator=gen).to(dev)
        labels = torch.randint(0, V, (4, 128, 1), generator=gen).to(dev)
        loss = model(ids, labels)
        loss.backward()
        opt.step()
        opt.clear_grad()
    torch.cuda.synchronize()
    if flag("use_fused_comm"):
        assert getattr(m.emb.weight, "_symm", None) is not None, "K7 path not taken"
    torch.save({k: v.detach().float().cpu() for k, v in m.state_dict().items()}, f"{out}.{r}.pt")
    print(f"fused_vs_nccl {mode} loss={float(loss):.4f} ok", r, flush=True)
    from paddle_b200.parallel import symm
    symm.release_all()
    dist.destroy_process_group()
    sys.exit(0)

if mode == "stage3":
    # GroupSharded stage 3: K3 push all-gather + pull reduce-scatter over symmetric memory vs NCCL
    from paddle_b200.distributed.sharding import group_sharded_parallel
    dist.init_parallel_env()
    r = dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    m = build()
    opt = paddle.optimizer.AdamW(1e-3, parameters=m.parameters(), multi_precision=True,
                                 grad_clip=nn.ClipGradByGlobalNorm(0.05))
    paddle.set_flags({"FLAGS_stage3_nvl_transport": bool(flag("use_fused_comm"))})
    model, opt, _ = group_sharded_parallel(m, opt, level="p_g_os")
    used_fused = model._nvl is not None
    assert used_fused == bool(flag("use_fused_comm")), (used_fused, flag("use_fused_comm"))
    for s in range(4):
        x, y = data(s, r, dev)
        ((model(x) - y).float() ** 2).mean().backward()
        opt.step()
        opt.clear_grad()
    torch.cuda.synchronize()
    model.get_all_parameters()
    torch.save({k: v.detach().float().cpu() for k, v in m.state_dict().items()}, f"{out}.{r}.pt")
    print(f"fused_vs_nccl stage3 fused={used_fused} ok", r, flush=True)
    from paddle_b200.parallel import symm
    if model._nvl is not None:
        model._nvl.close()
    symm.release_all()
    dist.destroy_process_group()
    sys.exit(0)

if mode == "dp":
    dist.init_parallel_env()
    r = dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    m = build()
    model = paddle.DataParallel(m, comm_buffer_size=2, last_comm_buffer_size=1)
    used_fused = model._reducer._fused is not None
    opt = paddle.optimizer.AdamW(1e-3, parameters=m.parameters(), multi_precision=True)
else:
    st = fleet.DistributedStrategy()
    st.hybrid_configs = {"dp_degree": 1, "mp_degree": 1, "pp_degree": 1, "sharding_degree": dist.ParallelEnv().world_size,
                         "sharding_configs": {"split_param": True}}
    fleet.init(is_collective=True, strategy=st)
    r = dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    m = build()
    clip = nn.ClipGradByGlobalNorm(0.05) if mode in ("v2clip", "v2amp") else None
    opt = paddle.optimizer.AdamW(1e-3, parameters=m.parameters(), multi_precision=True, grad_clip=clip)
    model = fleet.distributed_model(m)
    opt = fleet.distributed_optimizer(opt)
    used_fused = opt._inner_opt._arena is not None
assert used_fused == bool(flag("use_fused_comm")), (used_fused, flag("use_fused_comm"))
if mode == "v2amp":
    # loss scaling through the fused K2+K3 path: 1/scale rides in the kernel's gradient multiplier, the squared norm of
    # the reduced shard is the inf check; step 2 overflows on rank 1 and must be skipped everywhere
    scaler = fleet.distributed_scaler(paddle.amp.GradScaler(init_loss_scaling=256.0, decr_every_n_nan_or_inf=1))
    for s in range(5):
        x, y = data(s if s < 2 else s - 1, r, dev)
        loss = ((model(x) - y).float() ** 2).mean()
        if s == 2 and r == 1:
            loss = loss * float("inf")
        scaler.scale(loss).backward()
        scaler.step(opt)
        scaler.update()
        opt.clear_grad()
    torch.cuda.synchronize()
    assert float(scaler._scale) == 128.0, float(scaler._scale)
else:
    for s in range(4):
        x, y = data(s, r, dev)
        ((model(x) - y).float() ** 2).mean().backward()
        opt.step()
        opt.clear_grad()
torch.cuda.synchronize()
torch.save({k: v.detach().float().cpu() for k, v in m.state_dict().items()}, f"{out}.{r}.pt")
print(f"fused_vs_nccl {mode} fused={used_fused} ok", r, flush=True)
from paddle_b200.parallel import symm
symm.release_all()
dist.destroy_process_group()
