"""tcgen05/TMA GEMM (K23/K13) against torch fp32 matmul, all operand majors, tails, epilogues."""
import pytest
import torch

from paddle_b200 import ops
from paddle_b200.ops import gemm as G

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref(a, b, ta, tb):
    A = a.float().t() if ta else a.float()
    B = b.float().t() if tb else b.float()
    return A @ B


@pytest.mark.parametrize("ta,tb", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (512, 1024, 512), (384, 640, 320), (1024, 4096, 1024), (264, 328, 200)])
def test_gemm_majors_and_tails(ta, tb, M, N, K):
    torch.manual_seed(M + N + K)
    a = (torch.randn((K, M) if ta else (M, K), device=DEV) * 0.5).bfloat16()
    b = (torch.randn((N, K) if tb else (K, N), device=DEV) * 0.5).bfloat16()
    assert G.supported(a, b, ta, tb)
    out = G.gemm(a, b, ta, tb)
    ref = _ref(a, b, ta, tb)
    assert out.dtype == torch.bfloat16 and out.shape == (M, N)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2 * (K ** 0.5) * 0.25)


def test_gemm_fp32_out_and_accumulate():
    torch.manual_seed(1)
    M, N, K = 512, 768, 1024
    a = torch.randn(K, M, device=DEV).bfloat16()     # wgrad shape: X^T dY
    b = torch.randn(K, N, device=DEV).bfloat16()
    acc = torch.randn(M, N, device=DEV)
    ref = acc + _ref(a, b, True, False)
    G.gemm(a, b, True, False, accum_into=acc)
    torch.testing.assert_close(acc, ref, rtol=1e-3, atol=1e-2)
    out = G.gemm(a, b, True, False, out_dtype=torch.float32)
    torch.testing.assert_close(out, _ref(a, b, True, False), rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("act", ["none", "gelu", "relu"])
def test_gemm_bias_act_epilogue(act):
    torch.manual_seed(2)
    M, N, K = 256, 512, 256
    a = (torch.randn(M, K, device=DEV) * 0.3).bfloat16()
    b = (torch.randn(K, N, device=DEV) * 0.3).bfloat16()
    bias = torch.randn(N, device=DEV).bfloat16()
    out = G.gemm(a, b, bias=bias, act=act)
    ref = _ref(a, b, False, False) + bias.float()
    ref = {"none": lambda x: x, "gelu": torch.nn.functional.gelu, "relu": torch.relu}[act](ref)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=5e-2)


def test_linear_autograd_through_tc_gemm():
    """F.linear forward/backward routed through the tcgen05 kernel == torch matmul autograd."""
    import paddle_b200 as paddle
    torch.manual_seed(3)
    x = (torch.randn(4, 128, 512, device=DEV) * 0.5).bfloat16().requires_grad_(True)
    w = paddle.framework.Parameter((torch.randn(512, 1024, device=DEV) * 0.05).bfloat16())
    y = paddle.nn.functional.linear(x, w)
    g = torch.randn_like(y)
    y.backward(g)
    xr, wr = x.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    (xr @ wr).backward(g.float())
    torch.testing.assert_close(y.float(), xr @ wr, rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(x.grad.float(), xr.grad, rtol=2e-2, atol=1e-1)
    torch.testing.assert_close(w.grad.float(), wr.grad, rtol=2e-2, atol=5e-1)
    # fp32 main_grad accumulation in the wgrad epilogue (K13)
    w2 = paddle.framework.Parameter(w.detach().clone())
    w2.main_grad = torch.zeros(512, 1024, device=DEV)
    w2._fuse_main_grad = True
    x2 = x.detach().clone().requires_grad_(True)
    for _ in range(2):
        paddle.nn.functional.linear(x2, w2).backward(g)
    assert w2.grad is None
    torch.testing.assert_close(w2.main_grad, 2 * wr.grad, rtol=1e-2, atol=5e-1)


@pytest.mark.parametrize("ta,tb", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(512, 512, 64), (640, 768, 320), (520, 1000, 200), (2048, 1536, 4096)])
def test_gemm_cta_pair_kernel(ta, tb, M, N, K):
    """256x256 tiles on CTA pairs (cta_group::2), incl. partial pair-tiles in M and N, vs the single-CTA kernel."""
    from paddle_b200 import lib
    from paddle_b200.framework import flags
    assert hasattr(lib.kernels.require(), "pdb200_gemm_bf16_pair")
    torch.manual_seed(M + N + K + 1)
    a = (torch.randn((K, M) if ta else (M, K), device=DEV) * 0.5).bfloat16()
    b = (torch.randn((N, K) if tb else (K, N), device=DEV) * 0.5).bfloat16()
    flags.set_flags({"gemm_cta_pair": True})
    out = G.gemm(a, b, ta, tb)
    flags.set_flags({"gemm_cta_pair": False})
    single = G.gemm(a, b, ta, tb)
    flags.set_flags({"gemm_cta_pair": True})
    ref = _ref(a, b, ta, tb)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2 * (K ** 0.5) * 0.25)
    torch.testing.assert_close(out.float(), single.float(), rtol=1e-2, atol=1e-2 * (K ** 0.5) * 0.25)


def test_gemm_cta_pair_accumulate_and_epilogue():
    torch.manual_seed(5)
    M, N, K = 768, 1024, 512
    a = torch.randn(K, M, device=DEV).bfloat16()
    b = torch.randn(K, N, device=DEV).bfloat16()
    acc = torch.randn(M, N, device=DEV)
    ref = acc + _ref(a, b, True, False)
    G.gemm(a, b, True, False, accum_into=acc)
    torch.testing.assert_close(acc, ref, rtol=1e-3, atol=1e-2)
    x = (torch.randn(M, K, device=DEV) * 0.3).bfloat16()
    w = (torch.randn(K, N, device=DEV) * 0.3).bfloat16()
    bias = torch.randn(N, device=DEV).bfloat16()
    out = G.gemm(x, w, bias=bias, act="gelu")
    torch.testing.assert_close(out.float(), torch.nn.functional.gelu(_ref(x, w, False, False) + bias.float()), rtol=2e-2, atol=5e-2)


@pytest.mark.parametrize("ta,tb", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(512, 512, 64), (512, 1024, 128), (640, 768, 320), (520, 1000, 200), (2048, 1536, 4096),
                                   (4096, 2816, 512), (8192, 4096, 256)])
def test_gemm_macro_tile_kernel(ta, tb, M, N, K):
    """256x512 blocks per CTA pair with the TMA-store epilogue: full rounds, half-block tail rounds, partial blocks in
    M and N (the TMA clips them), every operand major; against fp32 matmul."""
    from paddle_b200.framework import flags
    torch.manual_seed(M + N + K + 7)
    a = (torch.randn((K, M) if ta else (M, K), device=DEV) * 0.5).bfloat16()
    b = (torch.randn((N, K) if tb else (K, N), device=DEV) * 0.5).bfloat16()
    flags.set_flags({"gemm_kernel": "macro"})
    try:
        out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
        G.gemm(a, b, ta, tb, out=out)
    finally:
        flags.set_flags({"gemm_kernel": ""})
    ref = _ref(a, b, ta, tb)
    assert not torch.isnan(out.float()).any(), "part of C was never written"
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2 * (K ** 0.5) * 0.25)


def test_gemm_macro_tile_accumulate_bias_act():
    from paddle_b200.framework import flags
    torch.manual_seed(11)
    M, N, K = 768, 1280, 512
    a = torch.randn(K, M, device=DEV).bfloat16()
    b = torch.randn(K, N, device=DEV).bfloat16()
    acc = torch.randn(M, N, device=DEV)
    ref = acc + _ref(a, b, True, False)
    flags.set_flags({"gemm_kernel": "macro"})
    try:
        G.gemm(a, b, True, False, accum_into=acc)            # C += through the TMA reduce-add
        G.gemm(a, b, True, False, accum_into=acc)
        ref = ref + _ref(a, b, True, False)
        torch.testing.assert_close(acc, ref, rtol=1e-3, atol=2e-2)
        f32 = G.gemm(a, b, True, False, out_dtype=torch.float32)
        torch.testing.assert_close(f32, _ref(a, b, True, False), rtol=1e-3, atol=1e-2)
        x = (torch.randn(M, K, device=DEV) * 0.3).bfloat16()
        w = (torch.randn(K, N, device=DEV) * 0.3).bfloat16()
        bias = torch.randn(N, device=DEV).bfloat16()
        out = G.gemm(x, w, bias=bias, act="gelu")
    finally:
        flags.set_flags({"gemm_kernel": ""})
    torch.testing.assert_close(out.float(), torch.nn.functional.gelu(_ref(x, w, False, False) + bias.float()), rtol=2e-2, atol=5e-2)


def _grouped_case(E, cap, counts, K, N, seed):
    torch.manual_seed(seed)
    x = (torch.randn(cap, K, device=DEV) * 0.5).bfloat16()
    w = (torch.randn(E, K, N, device=DEV) * 0.1).bfloat16()
    cnt = torch.tensor(counts, dtype=torch.int64, device=DEV)
    total = sum(counts)
    x[total:] = float("nan")                     # rows nobody owns must never leak into owned rows or into dW
    return x, w, cnt, total


@pytest.mark.parametrize("counts", [[300, 0, 129, 64], [1, 1, 1, 1], [0, 0, 700, 0], [513, 255, 128, 127]])
def test_grouped_gemm_forward_dgrad_wgrad(counts):
    """Device-scheduled grouped GEMMs (K9/K10) against a per-expert fp32 loop; experts with zero rows, tiles that
    straddle two experts, NaN in the unowned tail of the slab."""
    E, cap, K, N = len(counts), 1024, 256, 384
    x, w, cnt, total = _grouped_case(E, cap, counts, K, N, 17)
    xr = x[:total].float().requires_grad_(True)
    wr = w.float().requires_grad_(True)
    outs, o = [], 0
    for e, c in enumerate(counts):
        outs.append(xr[o:o + c] @ wr[e])
        o += c
    ref = torch.cat(outs)
    xg = x.clone().requires_grad_(True)
    wg = w.clone().requires_grad_(True)
    y = ops.grouped_linear(xg, wg, cnt)
    assert y.shape == (cap, N)
    torch.testing.assert_close(y[:total].float(), ref, rtol=2e-2, atol=2e-2 * (K ** 0.5) * 0.25)
    g = (torch.randn(cap, N, device=DEV) * 0.5).bfloat16()
    g[total:] = float("nan")
    ref.backward(g[:total].float())
    y.backward(g)
    torch.testing.assert_close(xg.grad[:total].float(), xr.grad, rtol=2e-2, atol=5e-2)
    assert torch.isfinite(wg.grad.float()).all(), "unowned rows leaked into the weight gradient"
    torch.testing.assert_close(wg.grad.float(), wr.grad, rtol=2e-2, atol=2e-2 * (max(counts) ** 0.5) * 0.25 + 1e-3)


def test_grouped_experts_ffn_matches_expert_list():
    import paddle_b200 as paddle
    from paddle_b200.incubate.distributed.models.moe import GroupedExpertsFFN
    from paddle_b200.models.moe import ExpertFFN, MoEConfig
    cfg = MoEConfig.tiny(hidden_size=128, intermediate_size=256)
    paddle.seed(5)
    paddle.set_default_dtype("bfloat16")
    try:
        experts = [ExpertFFN(cfg).to(DEV) for _ in range(3)]
        grouped = GroupedExpertsFFN.from_experts(experts).to(DEV)
    finally:
        paddle.set_default_dtype("float32")
    counts = [200, 0, 150]
    cap = 512
    x = (torch.randn(cap, 128, device=DEV) * 0.5).bfloat16()
    cnt = torch.tensor(counts, dtype=torch.int64, device=DEV)
    torch.cuda.set_sync_debug_mode("error")          # the grouped path must not synchronise with the host
    try:
        y = grouped(x, cnt)
        y[:350].float().sum().backward()
    finally:
        torch.cuda.set_sync_debug_mode("default")
    o = 0
    for e, c in enumerate(counts):
        if c:
            ref = experts[e](x[o:o + c])
            torch.testing.assert_close(y[o:o + c].float(), ref.float(), rtol=3e-2, atol=3e-2)
        o += c
    assert torch.isfinite(grouped.gate_up_weight.grad.float()).all()
    assert float(grouped.gate_up_weight.grad[1].abs().max()) == 0.0      # expert 1 saw no rows
