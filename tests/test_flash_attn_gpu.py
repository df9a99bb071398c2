"""tcgen05 flash attention (K16/K17) against fp32 naive attention (the reference's own oracle, test_flash_attention.py)."""
import math

import pytest
import torch

from paddle_b200 import lib, ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def naive(q, k, v, causal):
    b, sq, hq, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    g = hq // hk
    kf = k.float().repeat_interleave(g, dim=2)
    vf = v.float().repeat_interleave(g, dim=2)
    s = torch.einsum("bqhd,bkhd->bhqk", q.float(), kf) / math.sqrt(d)
    if causal:
        m = torch.ones(sq, sk, dtype=torch.bool, device=q.device).tril(sk - sq)
        s = s.masked_fill(~m, float("-inf"))
    p = torch.softmax(s, -1)
    return torch.einsum("bhqk,bkhd->bqhd", p, vf), torch.logsumexp(s, -1)


@pytest.mark.parametrize("d", [128, 64])
@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("b,s,hq,hk", [(2, 256, 4, 4), (1, 512, 8, 2), (1, 1024, 2, 2)])
def test_flash_fwd_matches_naive(d, causal, b, s, hq, hk):
    assert hasattr(lib.kernels.require(), "pdb200_flash_attn_fwd")
    torch.manual_seed(s + d)
    q = torch.randn(b, s, hq, d, device=DEV).bfloat16()
    k = torch.randn(b, s, hk, d, device=DEV).bfloat16()
    v = torch.randn(b, s, hk, d, device=DEV).bfloat16()
    o, lse = ops.flash_attention(q, k, v, causal=causal, return_lse=True)
    assert lse is not None, "native flash attention kernel was not used"
    ro, rlse = naive(q, k, v, causal)
    torch.testing.assert_close(o.float(), ro, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(lse, rlse, rtol=1e-3, atol=2e-3)


def test_flash_fwd_cross_lengths():
    torch.manual_seed(9)
    q = torch.randn(1, 128, 4, 128, device=DEV).bfloat16()
    k = torch.randn(1, 384, 4, 128, device=DEV).bfloat16()
    v = torch.randn(1, 384, 4, 128, device=DEV).bfloat16()
    for causal in (False, True):
        o, lse = ops.flash_attention(q, k, v, causal=causal, return_lse=True)
        ro, rlse = naive(q, k, v, causal)
        torch.testing.assert_close(o.float(), ro, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("d", [128, 64])
@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("b,sq,sk,hq,hk", [(1, 256, 256, 4, 4), (2, 384, 384, 4, 2), (1, 128, 384, 2, 1), (1, 1024, 1024, 2, 2)])
def test_flash_backward_matches_naive(d, causal, b, sq, sk, hq, hk):
    assert hasattr(lib.kernels.require(), "pdb200_flash_attn_bwd"), "native backward missing"
    torch.manual_seed(4 + sq + d)
    q = torch.randn(b, sq, hq, d, device=DEV).bfloat16().requires_grad_(True)
    k = torch.randn(b, sk, hk, d, device=DEV).bfloat16().requires_grad_(True)
    v = torch.randn(b, sk, hk, d, device=DEV).bfloat16().requires_grad_(True)
    o = ops.flash_attention(q, k, v, causal=causal)
    g = torch.randn_like(o)
    o.backward(g)
    qr, kr, vr = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    ro, _ = naive(qr, kr, vr, causal)
    ro.backward(g.float())
    for name, a, r in (("dq", q, qr), ("dk", k, kr), ("dv", v, vr)):
        torch.testing.assert_close(a.grad.float(), r.grad, rtol=3e-2, atol=3e-2, msg=lambda m, n=name: f"{n}: {m}")


def test_flash_backward_is_deterministic():
    torch.manual_seed(0)
    q, k, v = (torch.randn(1, 512, 4, 128, device=DEV).bfloat16().requires_grad_(True) for _ in range(3))
    g = torch.randn(1, 512, 4, 128, device=DEV).bfloat16()
    outs = []
    for _ in range(2):
        for t in (q, k, v):
            t.grad = None
        ops.flash_attention(q, k, v, causal=True).backward(g)
        outs.append([t.grad.clone() for t in (q, k, v)])
    for a, b_ in zip(*outs):
        assert torch.equal(a, b_)


# ---------------------------------------------------------------------------------------------------------------------
# tails (lengths that are not multiples of 128), fp16, packed variable-length batches, dropout

def naive_masked(q, k, v, causal, keep=None, inv_keep=1.0):
    """fp32 oracle with an optional dropout keep-mask [b, hq, sq, sk] applied after the softmax."""
    b, sq, hq, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    g = hq // hk
    kf = k.float().repeat_interleave(g, dim=2)
    vf = v.float().repeat_interleave(g, dim=2)
    s = torch.einsum("bqhd,bkhd->bhqk", q.float(), kf) / math.sqrt(d)
    if causal:
        m = torch.ones(sq, sk, dtype=torch.bool, device=q.device).tril(sk - sq)
        sees = m.any(-1, keepdim=True)                     # rows that see no key: output 0, no gradient
        s = torch.where(sees, s.masked_fill(~m, float("-inf")), torch.zeros_like(s))
        p = torch.softmax(s, -1) * sees
    else:
        p = torch.softmax(s, -1)
    if keep is not None:
        p = p * keep * inv_keep
    return torch.einsum("bhqk,bkhd->bqhd", p, vf)


def _grads_close(pairs, tol=3e-2):
    for name, a, r in pairs:
        torch.testing.assert_close(a.float(), r, rtol=tol, atol=tol, msg=lambda m, n=name: f"{n}: {m}")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("b,sq,sk,hq,hk,d", [(2, 200, 200, 4, 2, 128), (1, 77, 333, 2, 2, 64), (3, 129, 129, 2, 1, 128),
                                             (1, 1, 1, 2, 2, 64), (2, 300, 45, 2, 2, 128)])
def test_flash_tails_and_fp16_match_naive(dtype, causal, b, sq, sk, hq, hk, d):
    torch.manual_seed(sq * 7 + sk)
    q = torch.randn(b, sq, hq, d, device=DEV).to(dtype).requires_grad_(True)
    k = torch.randn(b, sk, hk, d, device=DEV).to(dtype).requires_grad_(True)
    v = torch.randn(b, sk, hk, d, device=DEV).to(dtype).requires_grad_(True)
    o, lse = ops.flash_attention(q, k, v, causal=causal, return_lse=True)
    assert lse is not None, "native flash attention kernel was not used"
    g = torch.randn_like(o)
    o.backward(g)
    qr, kr, vr = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    ro = naive_masked(qr, kr, vr, causal)
    ro.backward(g.float())
    torch.testing.assert_close(o.float(), ro, rtol=2e-2, atol=2e-2)
    assert torch.isfinite(q.grad).all() and torch.isfinite(k.grad).all() and torch.isfinite(v.grad).all()
    _grads_close((("dq", q.grad, qr.grad), ("dk", k.grad, kr.grad), ("dv", v.grad, vr.grad)))


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("d,hq,hk", [(128, 4, 2), (64, 2, 2)])
def test_flash_varlen_matches_per_sequence_naive(causal, d, hq, hk):
    """Packed batch, one launch, cu_seqlens on the device: every sequence equals the fp32 attention of that sequence
    alone, and no host synchronisation happens inside the op."""
    torch.manual_seed(31 + d)
    lens_q = [5, 128, 300, 1, 257]
    lens_k = lens_q if causal else [9, 200, 128, 77, 400]
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32, device=DEV)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32, device=DEV)
    tq, tk = sum(lens_q), sum(lens_k)
    q = torch.randn(tq, hq, d, device=DEV).bfloat16().requires_grad_(True)
    k = torch.randn(tk, hk, d, device=DEV).bfloat16().requires_grad_(True)
    v = torch.randn(tk, hk, d, device=DEV).bfloat16().requires_grad_(True)
    g = torch.randn(tq, hq, d, device=DEV).bfloat16()
    torch.cuda.synchronize()
    torch.cuda.set_sync_debug_mode("error")
    try:
        o = ops.flash_attn_varlen(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), 1.0 / math.sqrt(d), causal=causal)
        o.backward(g)
    finally:
        torch.cuda.set_sync_debug_mode("default")
    oq, ok = 0, 0
    for lq, lk in zip(lens_q, lens_k):
        qr, kr, vr = (t.detach()[a:a + n][None].float().requires_grad_(True)
                      for t, a, n in ((q, oq, lq), (k, ok, lk), (v, ok, lk)))
        ro = naive_masked(qr, kr, vr, causal)
        ro.backward(g[oq:oq + lq][None].float())
        torch.testing.assert_close(o[oq:oq + lq][None].float(), ro, rtol=2e-2, atol=2e-2)
        _grads_close((("dq", q.grad[oq:oq + lq][None], qr.grad), ("dk", k.grad[ok:ok + lk][None], kr.grad),
                      ("dv", v.grad[ok:ok + lk][None], vr.grad)))
        oq, ok = oq + lq, ok + lk


def _mix32(x):
    m = 0xFFFFFFFF
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & m
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & m
    return x ^ (x >> 16)


def _keep_mask(p_drop, seed, offset, b, hq, sq, sk):
    """Host restatement of csrc/kernels/flash_common.cuh (make_dropout / dropout_patch / dropout_keep)."""
    m64 = (1 << 64) - 1
    thr = min(255, max(1, int((1.0 - p_drop) * 256.0 + 0.5)))
    z = (seed * 0x9E3779B97F4A7C15 + offset * 0xBF58476D1CE4E5B9 + 0x94D049BB133111EB) & m64
    lo, hi = z & 0xFFFFFFFF, z >> 32
    m = 0xFFFFFFFF
    bh = torch.arange(b * hq, dtype=torch.int64, device=DEV).view(b, hq, 1, 1)
    qi = torch.arange(sq, dtype=torch.int64, device=DEV).view(1, 1, sq, 1)
    ki = torch.arange(sk, dtype=torch.int64, device=DEV).view(1, 1, 1, sk)
    a = _mix32((((qi >> 1) * 0x9E3779B9) + lo) & m)
    w = _mix32(a ^ ((((ki >> 1) * 0x85EBCA6B) + bh * 0xC2B2AE35 + hi) & m))
    byte = (w >> (((qi & 1) * 2 + (ki & 1)) * 8)) & 0xFF
    return (byte < thr), 256.0 / thr


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("b,sq,sk,hq,hk,d", [(2, 256, 256, 4, 2, 128), (1, 200, 333, 2, 2, 64)])
def test_flash_dropout_matches_naive_with_the_same_mask(causal, b, sq, sk, hq, hk, d):
    p_drop, seed, offset = 0.2, 1234, 8
    torch.manual_seed(3)
    q = torch.randn(b, sq, hq, d, device=DEV).bfloat16().requires_grad_(True)
    k = torch.randn(b, sk, hk, d, device=DEV).bfloat16().requires_grad_(True)
    v = torch.randn(b, sk, hk, d, device=DEV).bfloat16().requires_grad_(True)
    o = ops.flash_attention(q, k, v, dropout=p_drop, causal=causal, fixed_seed_offset=(seed, offset))
    g = torch.randn_like(o)
    o.backward(g)
    keep, inv_keep = _keep_mask(p_drop, seed, offset, b, hq, sq, sk)
    frac = keep.float().mean().item()
    assert abs(frac - (1 - p_drop)) < 0.02, f"keep fraction {frac}"
    qr, kr, vr = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    ro = naive_masked(qr, kr, vr, causal, keep.float(), inv_keep)
    ro.backward(g.float())
    torch.testing.assert_close(o.float(), ro, rtol=3e-2, atol=3e-2)
    _grads_close((("dq", q.grad, qr.grad), ("dk", k.grad, kr.grad), ("dv", v.grad, vr.grad)), tol=4e-2)
    # a different offset draws a different mask; eval mode drops nothing
    o2 = ops.flash_attention(q, k, v, dropout=p_drop, causal=causal, fixed_seed_offset=(seed, offset + 4))
    assert (o2 - o).abs().max() > 1e-3
    o3 = ops.flash_attention(q, k, v, dropout=p_drop, causal=causal, training=False)
    torch.testing.assert_close(o3.float(), naive_masked(qr, kr, vr, causal), rtol=2e-2, atol=2e-2)


def test_flash_dropout_advances_the_generator():
    torch.manual_seed(5)
    q, k, v = (torch.randn(1, 256, 2, 64, device=DEV).bfloat16() for _ in range(3))
    a = ops.flash_attention(q, k, v, dropout=0.3)
    b_ = ops.flash_attention(q, k, v, dropout=0.3)
    assert (a - b_).abs().max() > 1e-3, "two calls drew the same dropout mask"
    torch.manual_seed(5)                # same generator history -> same mask
    q2, k2, v2 = (torch.randn(1, 256, 2, 64, device=DEV).bfloat16() for _ in range(3))
    a2 = ops.flash_attention(q2, k2, v2, dropout=0.3)
    torch.testing.assert_close(a2, a, rtol=0, atol=0)
