"""The fused NVLink collectives on ONE GPU: the "ranks" are segments of this GPU's memory handed to the real entry
points as peer pointers (no NVLS address, so the peer-access code paths run), one call per emulated rank, one after
the other on one stream, with the cross-rank barriers switched off (pdb200_nvl_set_emulation).  Every result is
checked against a plain fp32 PyTorch statement of the collective.  The NVLS (multimem) paths and the barriers
themselves need >= 2 GPUs: tests/test_nvl_gpu.py and bench.py's selfcheck block."""
import ctypes

import pytest
import torch

from paddle_b200 import lib

pytestmark = pytest.mark.gpu
DEV = "cuda"
N = 4                      # emulated ranks
SEG = 64 << 20             # bytes per emulated segment


@pytest.fixture()
def world():
    k = lib.kernels.require()
    base = torch.zeros(N * SEG, dtype=torch.uint8, device=DEV)
    ptrs = (ctypes.c_uint64 * N)(*[base.data_ptr() + r * SEG for r in range(N)])
    k.pdb200_nvl_set_emulation(1)

    class W:
        kern = lib.kernels
        peers = ctypes.cast(ptrs, ctypes.c_void_p)
        stream = torch.cuda.current_stream().cuda_stream

        @staticmethod
        def seg(r, off, shape, dtype):
            n = 1
            for s in shape:
                n *= s
            nb = n * torch.empty(0, dtype=dtype).element_size()
            return base[r * SEG + off: r * SEG + off + nb].view(dtype).view(*shape)
    yield W
    torch.cuda.synchronize()
    k.pdb200_nvl_set_emulation(0)


def _check(rc, what):
    lib.check(rc, what)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_k1_allreduce_scale(world, dtype):
    numel = (1 << 18) + 8
    torch.manual_seed(1)
    src = [torch.randn(numel, device=DEV).to(dtype) for _ in range(N)]
    for r in range(N):
        world.seg(r, 4096, (numel,), dtype).copy_(src[r])
    ref = sum(s.float() for s in src) / N
    code = 0 if dtype == torch.float32 else 1
    for r in range(N):
        _check(world.kern.pdb200_nvl_allreduce(world.peers, 0, N, r, 4096, numel, code, 1.0 / N, 0, 1, 0, world.stream), "allreduce")
    tol = dict(rtol=1e-5, atol=1e-5) if dtype == torch.float32 else dict(rtol=2e-2, atol=2e-2)
    for r in range(N):
        torch.testing.assert_close(world.seg(r, 4096, (numel,), dtype).float(), ref, **tol)


def test_k3_unit_allgather_multi(world):
    sizes = [8192, 40960, 1024]                              # bf16 elements per rank shard, three "parameters"
    torch.manual_seed(2)
    shards = [[torch.randn(n, device=DEV).bfloat16() for n in sizes] for _ in range(N)]
    src_off, dst_off, segs, o = [], [], [], 4096
    for n in sizes:
        src_off.append(o)
        o += n * 2
    o = (o + 255) // 256 * 256
    for n in sizes:
        dst_off.append(o)
        o += n * 2 * N
    for r in range(N):
        for i, n in enumerate(sizes):
            world.seg(r, src_off[i], (n,), torch.bfloat16).copy_(shards[r][i])
    flat = (ctypes.c_int64 * (3 * len(sizes)))(*[v for i, n in enumerate(sizes) for v in (src_off[i], dst_off[i], n * 2)])
    for r in range(N):
        _check(world.kern.pdb200_nvl_allgather_multi(world.peers, 0, N, r, ctypes.cast(flat, ctypes.c_void_p), len(sizes), 0, 1, 8,
                                                      world.stream), "allgather_multi")
    for r in range(N):
        for i, n in enumerate(sizes):
            full = world.seg(r, dst_off[i], (N * n,), torch.bfloat16)
            ref = torch.cat([shards[q][i] for q in range(N)])
            assert torch.equal(full, ref), f"rank {r} param {i}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_k2_unit_reduce_scatter_multi(world, dtype):
    per = [4096, 24576]                                      # shard elements per rank, two "parameters"
    esz = 2 if dtype == torch.bfloat16 else 4
    torch.manual_seed(3)
    grads = [[torch.randn(n * N, device=DEV).to(dtype) for n in per] for _ in range(N)]
    offs, o = [], 4096
    for n in per:
        offs.append(o)
        o += (n * N * esz + 255) // 256 * 256
    for r in range(N):
        for i, n in enumerate(per):
            world.seg(r, offs[i], (n * N,), dtype).copy_(grads[r][i])
    outs = [[torch.full((n,), 3.0, dtype=torch.float32, device=DEV) for n in per] for _ in range(N)]
    code = 0 if dtype == torch.float32 else 1
    for accumulate in (0, 1):
        for r in range(N):
            flat = (ctypes.c_int64 * (3 * len(per)))(*[v for i, n in enumerate(per) for v in (offs[i], n, outs[r][i].data_ptr())])
            _check(world.kern.pdb200_nvl_reduce_scatter_multi(world.peers, 0, N, r, ctypes.cast(flat, ctypes.c_void_p), len(per), code,
                                                               1.0 / N, accumulate, 0, 1, 8, world.stream), "reduce_scatter_multi")
    for r in range(N):
        for i, n in enumerate(per):
            avg = sum(g[i].float()[r * n:(r + 1) * n] for g in grads) / N
            torch.testing.assert_close(outs[r][i], 2 * avg, rtol=1e-5, atol=1e-5)   # written once, accumulated once


def test_k2_k3_rs_adamw_ag_matches_fp32_adamw(world):
    S = 8 * 2048                                             # shard elements per rank
    total = S * N
    goff, poff = 4096, 4096 + total * 2 + 4096
    torch.manual_seed(4)
    master_full = torch.randn(total, device=DEV)
    for r in range(N):
        world.seg(r, poff, (total,), torch.bfloat16).copy_(master_full.bfloat16())
    masters = [master_full[r * S:(r + 1) * S].clone() for r in range(N)]
    ms = [torch.zeros(S, device=DEV) for _ in range(N)]
    vs = [torch.zeros(S, device=DEV) for _ in range(N)]
    ref_p = master_full.clone().requires_grad_(True)
    ref_opt = torch.optim.AdamW([ref_p], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    pows = torch.tensor([0.9, 0.95], device=DEV)
    for step in range(3):
        gs = [torch.randn(total, device=DEV).bfloat16() for _ in range(N)]
        for r in range(N):
            world.seg(r, goff, (total,), torch.bfloat16).copy_(gs[r])
        ref_p.grad = sum(g.float() for g in gs) / N
        ref_opt.step()
        for r in range(N):
            _check(world.kern.pdb200_nvl_rs_adamw_ag(world.peers, 0, N, r, goff, 1, poff, r * S, S, masters[r].data_ptr(),
                                                      ms[r].data_ptr(), vs[r].data_ptr(), 1e-2, 0.9, 0.95, 1e-8, 0.1, pows.data_ptr(),
                                                      None, None, None, 0, 1, world.stream), "rs_adamw_ag")
        pows.mul_(torch.tensor([0.9, 0.95], device=DEV))
    got = torch.cat(masters)
    torch.testing.assert_close(got, ref_p.detach(), rtol=1e-4, atol=1e-5)
    for r in range(N):      # every rank's bf16 parameter buffer holds all updated shards
        torch.testing.assert_close(world.seg(r, poff, (total,), torch.bfloat16).float(), got.bfloat16().float(), rtol=0, atol=0)


def test_k6_rs_sum_and_k8_ce_combine_and_k7_embed(world):
    # K6 tail: sum reduce-scatter of bf16 partial products
    rows, cols = 64, 256
    torch.manual_seed(5)
    parts = [torch.randn(N * rows, cols, device=DEV).bfloat16() for _ in range(N)]
    for r in range(N):
        world.seg(r, 4096, (N * rows, cols), torch.bfloat16).copy_(parts[r])
    for r in range(N):
        out = torch.empty(rows, cols, dtype=torch.bfloat16, device=DEV)
        _check(world.kern.pdb200_nvl_rs_sum_bf16(world.peers, 0, N, r, 4096, r * rows * cols, rows * cols, out.data_ptr(), 0, 1,
                                                  world.stream), "rs_sum")
        ref = sum(p.float()[r * rows:(r + 1) * rows] for p in parts)
        torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=5e-2)
    # K8: combine per-rank (max, sumexp, target) statistics == log-softmax over the concatenated vocabulary
    R, Vp = 96, 50
    logits = torch.randn(R, N * Vp, device=DEV)
    labels = torch.randint(0, N * Vp, (R,), device=DEV)
    labels[3] = -100
    for r in range(N):
        sh = logits[:, r * Vp:(r + 1) * Vp]
        mx = sh.max(1).values
        se = torch.exp(sh - mx[:, None]).sum(1)
        tgt = torch.zeros(R, device=DEV)
        own = (labels >= r * Vp) & (labels < (r + 1) * Vp)
        tgt[own] = sh[own, labels[own] - r * Vp]
        world.seg(r, 8 << 20, (R, 3), torch.float32).copy_(torch.stack([mx, se, tgt], 1))
    loss = torch.empty(R, device=DEV)
    lse = torch.empty(R, device=DEV)
    _check(world.kern.pdb200_nvl_ce_combine(world.peers, N, 0, 8 << 20, labels.data_ptr(), loss.data_ptr(), lse.data_ptr(), R, -100, 0, 1,
                                             world.stream), "ce_combine")
    ref = torch.nn.functional.cross_entropy(logits, labels, ignore_index=-100, reduction="none")
    torch.testing.assert_close(loss, ref, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(lse, torch.logsumexp(logits, 1), rtol=1e-5, atol=1e-5)
    # K7: vocab-parallel embedding as a gather from the owner's shard
    vpr, hid = 128, 64
    tables = [torch.randn(vpr, hid, device=DEV).bfloat16() for _ in range(N)]
    for r in range(N):
        world.seg(r, 16 << 20, (vpr, hid), torch.bfloat16).copy_(tables[r])
    ids = torch.randint(0, N * vpr, (300,), device=DEV)
    out = torch.empty(300, hid, dtype=torch.bfloat16, device=DEV)
    _check(world.kern.pdb200_nvl_vocab_embed(world.peers, N, 1, 16 << 20, ids.data_ptr(), out.data_ptr(), 300, hid, vpr, 1, 0, 1,
                                              world.stream), "vocab_embed")
    assert torch.equal(out, torch.cat(tables)[ids])
