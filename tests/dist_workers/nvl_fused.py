"""Worker (>= 2 GPUs, NCCL): fused NVLink kernels against their NCCL compositions.
K1 scaled all-reduce, K2+K3 reduce-scatter + AdamW + all-gather, K8 vocab-parallel CE exchange, native arena."""
import os

os.environ.setdefault("PDB200_OVERLAP_MIN_GFLOP", "0")     # the small test shapes must take the overlapped paths

import torch

import paddle_b200 as paddle
import paddle_b200.distributed as dist
from paddle_b200.parallel import symm

dist.init_parallel_env()
r, n = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", torch.cuda.current_device())
g = dist.collective._get_default_group()
assert symm.available(g), "fused NVLink path unavailable"
arena = symm.get_arena(g, 256 << 20, key="test")
print(f"rank {r}: arena {arena.nbytes >> 20} MiB, multicast={'yes' if arena.mc_ptr else 'no'}", flush=True)

# ---- peer mapping sanity: write my rank into my segment, read peers' through their mappings
mine = arena.empty((1024,), torch.float32)
mine.fill_(float(r + 1))
symm.barrier(arena)
torch.cuda.synchronize()

# ---- K1: scaled all-reduce, fp32 and bf16, odd sizes (multiple of 16 B)
for dtype, numel in ((torch.float32, 1 << 20), (torch.bfloat16, (1 << 21) + 8), (torch.float32, 4)):
    torch.manual_seed(100 + r)
    t = arena.empty((numel,), dtype)
    src = torch.randn(numel, device=dev).to(dtype)
    t.copy_(src)
    ref = src.clone().float()
    dist.all_reduce(ref)
    ref = ref / n
    symm.allreduce_(t, scale=1.0 / n)
    torch.cuda.synchronize()
    tol = dict(rtol=1e-5, atol=1e-5) if dtype == torch.float32 else dict(rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(t.float(), ref, **tol)

# ---- K2 + K3: flat buffers, shard per rank
torch.manual_seed(7)
total = 8 * 4096 * n
pbuf = arena.empty((total,), torch.bfloat16)
gbuf = arena.empty((total,), torch.bfloat16)
master_full = torch.randn(total, device=dev)
dist.broadcast(master_full, 0)
pbuf.copy_(master_full.bfloat16())
S = total // n
master = master_full[r * S:(r + 1) * S].clone()
m = torch.zeros(S, device=dev)
v = torch.zeros(S, device=dev)
ref_p = master_full.clone().requires_grad_(True)
ref_opt = torch.optim.AdamW([ref_p], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
pows = torch.tensor([0.9, 0.95], device=dev)
for step in range(3):
    torch.manual_seed(1000 * step + r)
    grad = torch.randn(total, device=dev).bfloat16()
    gbuf.copy_(grad)
    gsum = grad.float().clone()
    dist.all_reduce(gsum)
    # the switch reduces bf16 inputs with fp32 accumulation and rounds once
    ref_p.grad = (gsum / n)
    symm.rs_adamw_ag(gbuf, pbuf, r * S, S, master, m, v, 1e-2, 0.9, 0.95, 1e-8, 0.1, pows)
    pows.mul_(torch.tensor([0.9, 0.95], device=dev))
    ref_opt.step()
    torch.cuda.synchronize()
torch.testing.assert_close(master, ref_p.detach()[r * S:(r + 1) * S], rtol=2e-2, atol=2e-3)
torch.testing.assert_close(pbuf.float(), ref_p.detach().bfloat16().float(), rtol=2e-2, atol=2e-2)

# ---- K8 through ParallelCrossEntropy (mp group = world)
from paddle_b200.distributed.fleet.layers.mpu import mp_ops
torch.manual_seed(3)
V, N = 1024 * n, 300
full = torch.randn(N, V, device=dev)
dist.broadcast(full, 0)
labels = torch.randint(0, V, (N, 1), device=dev)
labels[5] = -100
dist.broadcast(labels, 0)
shard = full[:, r * (V // n):(r + 1) * (V // n)].bfloat16().contiguous().requires_grad_(True)
loss = mp_ops._c_softmax_with_cross_entropy(shard, labels, group=g, ignore_index=-100)
loss.sum().backward()
ref = torch.nn.functional.cross_entropy(full.bfloat16().float(), labels.squeeze(-1), ignore_index=-100, reduction="none")
torch.testing.assert_close(loss.squeeze(-1), ref, rtol=1e-3, atol=2e-3)
# ---- K5 all-gather -> GEMM and K6 GEMM -> reduce-scatter on the tcgen05 kernel
torch.manual_seed(50 + r)
rows, K, N = 256, 512, 768
xs = (torch.randn(rows, K, device=dev) * 0.5).bfloat16()
w = (torch.randn(K, N, device=dev) * 0.1).bfloat16()
dist.broadcast(w, 0)
parts = []
dist.all_gather(parts, xs)
ref = torch.cat(parts, 0).float() @ w.float()
for _ in range(2):                       # twice: staging reuse + epoch handling
    got = symm.ag_gemm(xs, w, g)
    torch.cuda.synchronize()
    torch.testing.assert_close(got.float(), ref, rtol=2e-2, atol=5e-2)
xf = (torch.randn(rows * n, K, device=dev) * 0.5).bfloat16()
partial = xf.float() @ w.float()
dist.all_reduce(partial)
ref_rs = partial[r * rows:(r + 1) * rows]
for _ in range(2):
    got = symm.gemm_rs(xf, w, g)
    torch.cuda.synchronize()
    torch.testing.assert_close(got.float(), ref_rs, rtol=2e-2, atol=1e-1)
# overlapped K5 on a shape with several chunks per rank; 5 calls with fresh data walk both gathered buffers and the
# flag epochs; oracle = fp32 matmul of the NCCL-gathered activation
rows2, K2, N2 = 768, 1024, 1536
w2 = (torch.randn(K2, N2, device=dev) * 0.05).bfloat16()
dist.broadcast(w2, 0)
for it in range(5):
    torch.manual_seed(900 + 10 * it + r)
    xs2 = (torch.randn(rows2, K2, device=dev) * 0.5).bfloat16()
    parts = []
    dist.all_gather(parts, xs2)
    ref2 = torch.cat(parts, 0).float() @ w2.float()
    got = symm.ag_gemm(xs2, w2, g)
    torch.testing.assert_close(got.float(), ref2, rtol=2e-2, atol=1e-1, msg=lambda m: f"overlapped ag_gemm call {it}: {m}")
assert any(v["calls"] >= 5 for v in symm._ag_state.values()), "the gated (overlapped) path did not run"
# overlapped K6: blocks leave for their owners while the GEMM still computes the later ones; 5 calls, fresh data
w3 = (torch.randn(K2, N2, device=dev) * 0.05).bfloat16()
dist.broadcast(w3, 0)
for it in range(5):
    torch.manual_seed(700 + 10 * it + r)
    xf2 = (torch.randn(rows2 * n, K2, device=dev) * 0.5).bfloat16()
    partial2 = xf2.float() @ w3.float()
    dist.all_reduce(partial2)
    got = symm.gemm_rs(xf2, w3, g)
    torch.testing.assert_close(got.float(), partial2[r * rows2:(r + 1) * rows2], rtol=2e-2, atol=2e-1,
                               msg=lambda m: f"overlapped gemm_rs call {it}: {m}")
assert any(v["calls"] >= 5 for v in symm._rs_state.values()), "the overlapped reduce-scatter path did not run"
for it in range(4):                       # overlapped K4 = overlapped K6 + copy-engine all-gather of the reduced blocks
    torch.manual_seed(300 + 10 * it + r)
    xf2 = (torch.randn(rows2 * n, K2, device=dev) * 0.5).bfloat16()
    partial2 = xf2.float() @ w3.float()
    dist.all_reduce(partial2)
    got = symm.gemm_ar(xf2, w3, g)
    torch.testing.assert_close(got.float(), partial2, rtol=2e-2, atol=2e-1, msg=lambda m: f"overlapped gemm_ar call {it}: {m}")
assert any(v["calls"] >= 4 for v in symm._ar_state.values()), "the overlapped all-reduce path did not run"
wt = w.t().contiguous()                   # trans_b variants (backward shapes)
got = symm.ag_gemm(xs, wt, g, trans_b=True)
torch.testing.assert_close(got.float(), ref, rtol=2e-2, atol=5e-2)
# ---- K4 GEMM -> all-reduce (RowParallel forward / Column dX) incl. the layers that route through it
for _ in range(2):
    got = symm.gemm_ar(xf, w, g)
    torch.cuda.synchronize()
    torch.testing.assert_close(got.float(), partial, rtol=2e-2, atol=1e-1)
got = symm.gemm_ar(xf, wt, g, trans_b=True)
torch.testing.assert_close(got.float(), partial, rtol=2e-2, atol=1e-1)

# ---- K7 vocab-parallel embedding: gather from the owner's shard == masked lookup + all-reduce
from paddle_b200.distributed.fleet.layers.mpu import mp_layers
torch.manual_seed(11)
Vp, H = 512, 256
table_full = torch.randn(Vp * n, H, device=dev).bfloat16()
dist.broadcast(table_full, 0)
emb = mp_layers.VocabParallelEmbedding(Vp * n, H, mp_group=g)
emb.to(dtype="bfloat16")
with torch.no_grad():
    emb.weight.copy_(table_full[r * Vp:(r + 1) * Vp])
ids = torch.randint(0, Vp * n, (4, 33), device=dev)
dist.broadcast(ids, 0)
out = emb(ids)
assert getattr(emb.weight, "_symm", None) is not None, "K7 path was not taken"
torch.testing.assert_close(out, table_full[ids], rtol=0, atol=0)
gout = torch.randn_like(out)
dist.broadcast(gout, 0)
out.backward(gout)
ref_dw = torch.zeros(Vp * n, H, device=dev)
ref_dw.index_add_(0, ids.reshape(-1), gout.reshape(-1, H).float())
torch.testing.assert_close(emb.weight.grad.float(), ref_dw[r * Vp:(r + 1) * Vp], rtol=2e-2, atol=2e-2)

# ---- K14 SyncBatchNorm statistics through the NVLS all-reduce
torch.manual_seed(20 + r)
bn = paddle.nn.SyncBatchNorm(24).to(dev)
xb = torch.randn(6, 24, 5, 5, device=dev, requires_grad=True)
yb = bn(xb)
allx = []
dist.all_gather(allx, xb.detach())
allx = torch.cat(allx, 0)
ref_bn = torch.nn.functional.batch_norm(allx, None, None, bn.weight.detach(), bn.bias.detach(), True, 0.1, 1e-5)
torch.testing.assert_close(yb.detach(), ref_bn[r * 6:(r + 1) * 6], rtol=1e-4, atol=1e-4)
yb.square().sum().backward()
assert torch.isfinite(xb.grad).all()
# ---- arena heap: free gives the bytes back at the same offset on every rank; exhaustion evicts cached staging buffers
small = symm.SymmArena(g, 8 << 20)
t1 = small.empty((3 << 20,), torch.uint8)
t2 = small.empty((1 << 20,), torch.uint8)
off1 = t1._symm[1]
small.free(t1)
t3 = small.empty((2 << 20,), torch.uint8)
assert t3._symm[1] == off1, "freed block was not reused"
offs = [None] * n
dist.all_gather_object(offs, (t2._symm[1], t3._symm[1]))
assert len(set(offs)) == 1, f"ranks disagree on offsets: {offs}"
small.free(t2)
small.free(t3)
cap = small.free_bytes()                      # the driver rounds the arena up to its mapping granularity
chunk = cap // 8 * 3 // 4096 * 4096           # two of these fit, three do not
s1 = symm._staging(small, "evict_a", (chunk,), torch.uint8)
s2 = symm._staging(small, "evict_b", (chunk,), torch.uint8)
s3 = symm._staging(small, "evict_c", (chunk,), torch.uint8)            # does not fit next to a and b: a (least recent) goes
assert (id(small), "evict_a", (chunk,), torch.uint8) not in symm._stage
assert (id(small), "evict_b", (chunk,), torch.uint8) in symm._stage
st = small.stats()
assert st["live_blocks"] == 2 and st["in_use"] == 2 * chunk, st
s3.fill_(r + 1)
try:
    small.empty((cap + (1 << 20),), torch.uint8)
    raise AssertionError("oversize allocation must fail")
except MemoryError:
    pass
small.close()
print("nvl_fused ok", r, flush=True)
symm.release_all()
dist.destroy_process_group()
