"""paddle.nn transformer layers against the reference's contract (python/paddle/nn/layer/transformer.py): output
tuples, masks, Cache / StaticCache incremental decoding equal to full causal decoding, stacks built from the first
layer's configuration, attr lists."""
import numpy as np
import pytest
import torch

import paddle
import paddle.nn as nn


def _mha_reference(m, q_in, k_in, v_in, mask):
    """Plain restatement: per-head scaled dot-product attention on the layer's own projections."""
    b, sq, e = q_in.shape
    h, d = m.num_heads, m.head_dim
    q = (q_in @ m.q_proj.weight + m.q_proj.bias).reshape(b, sq, h, d).permute(0, 2, 1, 3)
    k = (k_in @ m.k_proj.weight + m.k_proj.bias).reshape(b, -1, h, d).permute(0, 2, 1, 3)
    v = (v_in @ m.v_proj.weight + m.v_proj.bias).reshape(b, -1, h, d).permute(0, 2, 1, 3)
    s = q @ k.transpose(-1, -2) / d ** 0.5
    if mask is not None:
        s = s + mask
    w = torch.softmax(s, -1)
    o = (w @ v).permute(0, 2, 1, 3).reshape(b, sq, e)
    return o @ m.out_proj.weight + m.out_proj.bias, w


def test_multi_head_attention_outputs_masks_and_caches():
    paddle.seed(0)
    m = nn.MultiHeadAttention(16, 4, kdim=12, vdim=10, need_weights=True)
    m.eval()
    assert sorted(m.state_dict()) == ["k_proj.bias", "k_proj.weight", "out_proj.bias", "out_proj.weight", "q_proj.bias",
                                      "q_proj.weight", "v_proj.bias", "v_proj.weight"]
    assert m.k_proj.weight.shape == (12, 16) and m.v_proj.weight.shape == (10, 16)
    q, k, v = torch.randn(2, 5, 16), torch.randn(2, 7, 12), torch.randn(2, 7, 10)
    fmask = torch.randn(2, 1, 5, 7)
    with torch.no_grad():
        out, w = m(q, k, v, fmask)
        ref_out, ref_w = _mha_reference(m, q, k, v, fmask)
    assert w.shape == (2, 4, 5, 7)
    torch.testing.assert_close(out, ref_out, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(w, ref_w, rtol=1e-4, atol=1e-5)
    # bool and int masks keep the True / non-zero positions; the same result as the additive float mask
    keep = torch.rand(2, 1, 5, 7) > 0.3
    keep[..., 0] = True
    with torch.no_grad():
        o_bool, w_bool = m(q, k, v, keep)
        o_int, _ = m(q, k, v, keep.to(torch.int32))
        o_float, _ = m(q, k, v, torch.where(keep, 0.0, -1e9))
    torch.testing.assert_close(o_bool, o_float)
    torch.testing.assert_close(o_int, o_float)
    assert float(w_bool[~keep.expand_as(w_bool)].abs().max()) < 1e-6
    # the flash path (need_weights=False) computes the same thing and returns the tensor alone
    m.need_weights = False
    with torch.no_grad():
        o = m(q, k, v, keep)
    assert isinstance(o, torch.Tensor)
    torch.testing.assert_close(o, o_float, rtol=1e-4, atol=1e-5)
    # self attention with an incremental Cache == full attention under a causal mask
    sa = nn.MultiHeadAttention(16, 4)
    sa.eval()
    x = torch.randn(2, 6, 16)
    cache = sa.gen_cache(x)
    assert isinstance(cache, nn.MultiHeadAttention.Cache) and cache.k.shape == (2, 4, 0, 4)
    steps = []
    with torch.no_grad():
        for t in range(6):
            o, cache = sa(x[:, t:t + 1], None, None, None, cache)
            steps.append(o)
        full = sa(x, attn_mask=torch.tril(torch.ones(6, 6, dtype=torch.bool)))
    assert cache.k.shape == (2, 4, 6, 4)
    torch.testing.assert_close(torch.cat(steps, 1), full, rtol=1e-4, atol=1e-5)
    # StaticCache: memory projected once and returned untouched
    mem = torch.randn(2, 9, 16)
    static = sa.gen_cache(mem, mem, type=nn.MultiHeadAttention.StaticCache)
    assert isinstance(static, nn.MultiHeadAttention.StaticCache) and static.k.shape == (2, 4, 9, 4)
    with torch.no_grad():
        o, back = sa(x, None, None, None, static)
        want = sa(x, mem, mem)
    assert back is static
    torch.testing.assert_close(o, want, rtol=1e-4, atol=1e-5)
    given = sa.gen_cache(static.k, static.v)
    assert isinstance(given, nn.MultiHeadAttention.Cache) and given.k is static.k
    # need_weights + cache -> 3-tuple
    sa.need_weights = True
    o3 = sa(x[:, :1], None, None, None, sa.gen_cache(x))
    assert len(o3) == 3 and o3[1].shape == (2, 4, 1, 1)
    with pytest.raises(AssertionError):
        nn.MultiHeadAttention(10, 3)


def test_encoder_decoder_stacks_and_incremental_decoding():
    paddle.seed(1)
    layer = nn.TransformerEncoderLayer(16, 4, 32, dropout=0.0, activation="gelu", normalize_before=True, layer_norm_eps=1e-6)
    enc = nn.TransformerEncoder(layer, 3, nn.LayerNorm(16))
    assert enc.layers[0] is layer and len(enc.layers) == 3
    # further layers are built from the first one's arguments: same config, new parameters
    assert enc.layers[2].normalize_before and enc.layers[2].norm1._epsilon == 1e-6
    assert not torch.equal(enc.layers[0].linear1.weight, enc.layers[1].linear1.weight)
    names = sorted(enc.state_dict())
    assert "layers.0.self_attn.q_proj.weight" in names and "layers.2.linear2.bias" in names and "norm.weight" in names
    assert len(names) == 3 * 16 + 2
    src = torch.randn(2, 7, 16)
    pad = torch.ones(2, 1, 1, 7, dtype=torch.bool)
    pad[1, ..., 5:] = False
    enc.eval()
    with torch.no_grad():
        out = enc(src, pad)
        # padded keys do not influence the valid positions
        src2 = src.clone()
        src2[1, 5:] = 100.0
        out2 = enc(src2, pad)
    torch.testing.assert_close(out[1, :5], out2[1, :5], rtol=1e-4, atol=1e-5)
    # encoder cache: (output, new caches), one Cache per layer
    caches = enc.gen_cache(src)
    assert len(caches) == 3 and isinstance(caches[0], nn.MultiHeadAttention.Cache)
    with torch.no_grad():
        o, new = enc(src[:, :1], None, caches)
    assert o.shape == (2, 1, 16) and len(new) == 3 and new[0].k.shape == (2, 4, 1, 4)
    # post-norm layer, relu, separate dropouts
    post = nn.TransformerEncoderLayer(16, 4, 32, dropout=0.5, attn_dropout=0.0, act_dropout=0.25)
    assert post.dropout1.p == 0.5 and post.dropout.p == 0.25 and post.self_attn.dropout == 0.0 and not post.normalize_before
    # full transformer: decoding step by step with caches == teacher-forced decoding with the subsequent mask
    tr = nn.Transformer(16, 4, 2, 2, 32, dropout=0.0)
    tr.eval()
    assert isinstance(tr.encoder.norm, nn.LayerNorm) and isinstance(tr.decoder.norm, nn.LayerNorm)
    tgt = torch.randn(2, 5, 16)
    mask = tr.generate_square_subsequent_mask(5)
    assert mask.shape == (5, 5) and float(mask[0, 0]) == 0.0 and mask[0, 1] == float("-inf") and float(mask[4, 0]) == 0.0
    with torch.no_grad():
        full = tr(src, tgt, tgt_mask=mask)
        mem = tr.encoder(src)
        caches = tr.decoder.gen_cache(mem)
        assert len(caches) == 2 and isinstance(caches[0][0], nn.MultiHeadAttention.Cache) \
            and isinstance(caches[0][1], nn.MultiHeadAttention.StaticCache)
        zipped = tr.decoder.gen_cache(mem, do_zip=True)
        assert len(zipped) == 2 and len(zipped[0]) == 2 and isinstance(zipped[1][0], nn.MultiHeadAttention.StaticCache)
        steps = []
        for t in range(5):
            o, caches = tr.decoder(tgt[:, t:t + 1], mem, None, None, caches)
            steps.append(o)
    torch.testing.assert_close(torch.cat(steps, 1), full, rtol=1e-4, atol=1e-5)
    assert caches[1][0].k.shape == (2, 4, 5, 4)
    # memory mask reaches the cross attention
    mm = torch.ones(2, 1, 1, 7, dtype=torch.bool)
    mm[..., 3:] = False
    with torch.no_grad():
        a = tr(src, tgt, memory_mask=mm)
        src3 = src.clone()
        b = tr.decoder(tgt, tr.encoder(src3), memory_mask=mm)
    torch.testing.assert_close(a, b)
    # training step runs and reaches every parameter
    tr.train()
    loss = tr(src, tgt, tgt_mask=mask).square().mean()
    loss.backward()
    assert all(p.grad is not None for p in tr.parameters())


def test_attr_lists_and_custom_stacks():
    from paddle import ParamAttr
    from paddle.nn.initializer import Constant
    # one entry per sub-block: [self-attn, ffn] for the encoder layer
    layer = nn.TransformerEncoderLayer(8, 2, 16, weight_attr=[ParamAttr(initializer=Constant(0.5)), ParamAttr(initializer=Constant(0.25))],
                                       bias_attr=[False, None])
    assert float(layer.self_attn.q_proj.weight.mean()) == 0.5 and float(layer.linear1.weight.mean()) == 0.25
    assert layer.self_attn.q_proj.bias is None and layer.linear1.bias is not None
    dec = nn.TransformerDecoderLayer(8, 2, 16, bias_attr=[None, False, None])
    assert dec.self_attn.out_proj.bias is not None and dec.cross_attn.out_proj.bias is None and dec.linear2.bias is not None
    with pytest.raises(AssertionError):
        nn.TransformerDecoderLayer(8, 2, 16, bias_attr=[None, False])
    # Transformer spreads a 2-list as (attention, ffn) and a 3-list as (self, cross, ffn)
    tr = nn.Transformer(8, 2, 1, 1, 16, bias_attr=[False, None])
    assert tr.encoder.layers[0].self_attn.q_proj.bias is None and tr.encoder.layers[0].linear1.bias is not None
    assert tr.decoder.layers[0].cross_attn.q_proj.bias is None and tr.decoder.layers[0].linear1.bias is not None
    tr = nn.Transformer(8, 2, 1, 1, 16, bias_attr=[None, False, None])
    assert tr.encoder.layers[0].self_attn.q_proj.bias is not None and tr.decoder.layers[0].cross_attn.q_proj.bias is None
    # a named attr is suffixed per block so names stay unique
    layer = nn.TransformerEncoderLayer(8, 2, 16, weight_attr=ParamAttr(name="w"))
    assert layer.self_attn.q_proj.weight.name.startswith("w_0") and layer.linear1.weight.name.startswith("w_1")
    # custom encoder / decoder are used as given
    enc = nn.TransformerEncoder(nn.TransformerEncoderLayer(8, 2, 16), 1)
    tr = nn.Transformer(8, 2, custom_encoder=enc, num_decoder_layers=1, dim_feedforward=16)
    assert tr.encoder is enc and tr.encoder.norm is None
    assert tr(torch.randn(1, 3, 8), torch.randn(1, 2, 8)).shape == (1, 2, 8)
