"""Worker: ZeRO variants == plain data-parallel training (parameter equality after N steps):
DygraphShardingOptimizer V1 / V2 through fleet (sharding_degree=2) and GroupSharded stage 2 / stage 3."""
import sys

import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200 import nn
from paddle_b200.distributed import fleet

mode = sys.argv[1]


def build():
    paddle.seed(11)
    with paddle.utils.unique_name.guard():
        return nn.Sequential(nn.Linear(16, 64), nn.GELU(), nn.Linear(64, 64), nn.GELU(), nn.Linear(64, 8))


def data(step, rank):
    g = torch.Generator().manual_seed(100 * step + rank)
    return torch.randn(4, 16, generator=g), torch.randn(4, 8, generator=g)


def reference(steps, n, clip):
    m = build()
    opt = paddle.optimizer.AdamW(1e-2, parameters=m.parameters(), weight_decay=0.01,
                                 grad_clip=nn.ClipGradByGlobalNorm(0.5) if clip else None)
    for s in range(steps):
        loss = 0
        for r in range(n):
            x, y = data(s, r)
            loss = loss + ((m(x) - y) ** 2).mean() / n
        loss.backward()
        opt.step()
        opt.clear_grad()
    return m


STEPS = 4
if mode in ("v1amp", "v2amp"):
    # sharding V2 + GradScaler: the unscale / inf check must see the REDUCED shard; one step carries an inf
    st = fleet.DistributedStrategy()
    st.hybrid_configs = {"dp_degree": 1, "mp_degree": 1, "pp_degree": 1, "sharding_degree": 2,
                         "sharding_configs": {"split_param": mode == "v2amp"}}
    fleet.init(is_collective=True, strategy=st)
    r, n = dist.get_rank(), dist.get_world_size()
    m = build()
    opt = paddle.optimizer.AdamW(1e-2, parameters=m.parameters(), weight_decay=0.01, grad_clip=nn.ClipGradByGlobalNorm(0.5))
    model = fleet.distributed_model(m)
    opt = fleet.distributed_optimizer(opt)
    scaler = fleet.distributed_scaler(paddle.amp.GradScaler(init_loss_scaling=1024.0, decr_every_n_nan_or_inf=1))
    for s in range(STEPS + 1):
        x, y = data(s if s < 2 else s - 1, r)
        loss = ((model(x) - y) ** 2).mean()
        if s == 2 and r == 1:
            loss = loss * float("inf")          # rank 1 overflows in step 2: every rank must skip it
        scaler.scale(loss).backward()
        scaler.step(opt)
        scaler.update()
        opt.clear_grad()
    assert float(scaler._scale) == 512.0, float(scaler._scale)      # one decrease, on every rank
    ref = reference(STEPS, n, clip=True)
elif mode in ("v1", "v2", "v1fuse", "v1overlap"):
    st = fleet.DistributedStrategy()
    st.hybrid_configs = {"dp_degree": 1, "mp_degree": 1, "pp_degree": 1, "sharding_degree": 2,
                         "sharding_configs": {"split_param": mode == "v2", "tensor_fusion": mode == "v1fuse",
                                              "comm_overlap": mode == "v1overlap", "comm_buffer_size_MB": 1}}
    fleet.init(is_collective=True, strategy=st)
    r, n = dist.get_rank(), dist.get_world_size()
    m = build()
    opt = paddle.optimizer.AdamW(1e-2, parameters=m.parameters(), weight_decay=0.01, grad_clip=nn.ClipGradByGlobalNorm(0.5))
    model = fleet.distributed_model(m)
    opt = fleet.distributed_optimizer(opt)
    if mode in ("v1fuse", "v1overlap"):
        inner = opt._inner_opt
        assert inner._comm_buffers and inner.tensor_fusion == (mode == "v1fuse") and inner.comm_overlap == (mode == "v1overlap")
        assert sum(len(b.params) for b in inner._comm_buffers) == len([p for p in m.parameters() if p.requires_grad])
    for s in range(STEPS):
        x, y = data(s, r)
        ((model(x) - y) ** 2).mean().backward()
        opt.step()
        opt.clear_grad()
    ref = reference(STEPS, n, clip=True)
else:
    dist.init_parallel_env()
    r, n = dist.get_rank(), dist.get_world_size()
    m = build()
    opt = paddle.optimizer.AdamW(1e-2, parameters=m.parameters(), weight_decay=0.01, grad_clip=nn.ClipGradByGlobalNorm(0.5))
    level = {"stage2": "os_g", "stage3": "p_g_os"}[mode.split("_")[0]]
    offload = mode.endswith("_offload")
    model, opt, _ = dist.group_sharded_parallel(m, opt, level, segment_size=256, offload=offload,
                                                buffer_max_size=2 ** 12 if mode.startswith("stage2") else 2 ** 23)
    if mode.startswith("stage2"):
        # small gradients ride in per-destination buckets: grads are views of them, owners' params are packed flat
        sts = model._grad_storage_list
        assert sts and all(st._params for st in sts), "no GradStorage bucket was filled"
        bucketed = [p for p in m.parameters() if getattr(p, "_gs_grad_bucket", None) is not None]
        assert bucketed and all(p.grad is not None and p.grad.data_ptr() >= p._gs_grad_bucket.buffer.data_ptr() for p in bucketed)
        assert any(getattr(p, "_gs_grad_bucket", None) is None for p in m.parameters()), "expected one oversized parameter"
        assert opt._param_storages, "owners' parameters were not packed into ParamStorage"
    if offload:
        inner = opt._optim if hasattr(opt, "_optim") else opt
        assert inner._offload_pairs and all(sh.device.type == "cpu" and sh.dtype == torch.float32 for _, sh in inner._offload_pairs)
    for s in range(STEPS):
        x, y = data(s, r)
        ((model(x) - y) ** 2).mean().backward()
        opt.step()
        opt.clear_grad()
    if mode.startswith("stage3"):
        assert m[0].weight.numel() == 0, "stage 3 must release full parameters between steps"
        model.get_all_parameters()
    ref = reference(STEPS, n, clip=True)
for (k, a), (_, b) in zip(m.state_dict().items(), ref.state_dict().items()):
    torch.testing.assert_close(a.detach(), b.detach(), rtol=1e-4, atol=1e-5, msg=lambda s: f"{mode} {k}: {s}")
if mode.startswith("stage3"):
    import tempfile, os
    d = tempfile.mkdtemp() if r == 0 else None
    box = [d]
    dist.broadcast_object_list(box, 0)
    if r == 0:
        dist.save_group_sharded_model(model, box[0], opt)
        assert os.path.exists(os.path.join(box[0], "model.pdmodel")) and os.path.exists(os.path.join(box[0], "model.pdopt"))
    else:
        model.get_all_parameters()      # collective inside save on rank 0
        model.release_all_parameters()
print(f"sharding {mode} ok", r)
