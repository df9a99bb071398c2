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
