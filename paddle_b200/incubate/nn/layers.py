"""Fused layers (reference: python/paddle/incubate/nn/layer/{fused_linear,fused_dropout_add}.py)."""
from paddle_b200.nn.layer.layers import Layer

from . import functional as IF


class FusedLinear(Layer):
    def __init__(self, in_features, out_features, weight_attr=None, bias_attr=None, transpose_weight=False, name=None):
        super().__init__()
        shape = [out_features, in_features] if transpose_weight else [in_features, out_features]
        self.weight = self.create_parameter(shape, attr=weight_attr)
        self.bias = self.create_parameter([out_features], attr=bias_attr, is_bias=True)
        self.transpose_weight = transpose_weight

    def forward(self, x):
        return IF.fused_linear(x, self.weight, self.bias, self.transpose_weight)


class FusedDropoutAdd(Layer):
    def __init__(self, p=0.5, mode="upscale_in_train", name=None):
        super().__init__()
        self.p, self.mode = p, mode

    def forward(self, x, y):
        return IF.fused_dropout_add(x, y, self.p, self.training, self.mode)


# ---- fused transformer layers (reference: python/paddle/incubate/nn/layer/fused_transformer.py) -------------------
# The reference backs these with monolithic fused_attention / fused_feedforward ops; here they are compositions of this
# framework's kernels (flash attention, fused residual + norm, GEMM epilogues) behind the same constructor signatures.
class FusedBiasDropoutResidualLayerNorm(Layer):
    def __init__(self, embed_dim, dropout_rate=0.5, weight_attr=None, bias_attr=None, epsilon=1e-5, name=None):
        super().__init__()
        from paddle_b200.nn import initializer as I
        self.embed_dim, self.dropout_rate, self._epsilon = embed_dim, dropout_rate, epsilon
        self.linear_bias = self.create_parameter([embed_dim], attr=bias_attr, is_bias=True)
        self.ln_scale = self.create_parameter([embed_dim], attr=weight_attr, default_initializer=I.Constant(1.0))
        self.ln_bias = self.create_parameter([embed_dim], attr=bias_attr, is_bias=True)

    def forward(self, x, residual):
        from .functional import fused_bias_dropout_residual_layer_norm
        return fused_bias_dropout_residual_layer_norm(x, residual, self.linear_bias, self.ln_scale, self.ln_bias,
                                                      self.dropout_rate, self._epsilon, self.training)


class FusedMultiHeadAttention(Layer):
    def __init__(self, embed_dim, num_heads, dropout_rate=0.5, attn_dropout_rate=0.5, kdim=None, vdim=None,
                 normalize_before=False, need_weights=False, qkv_weight_attr=None, qkv_bias_attr=None,
                 linear_weight_attr=None, linear_bias_attr=None, pre_ln_scale_attr=None, pre_ln_bias_attr=None,
                 ln_scale_attr=None, ln_bias_attr=None, epsilon=1e-5, nranks=1, ring_id=-1, transpose_qkv_wb=False, name=None):
        super().__init__()
        import paddle_b200.nn as nn
        assert embed_dim % num_heads == 0
        self.embed_dim, self.num_heads, self.head_dim = embed_dim, num_heads, embed_dim // num_heads
        self.normalize_before, self.dropout_rate, self.attn_dropout_rate = normalize_before, dropout_rate, attn_dropout_rate
        self.qkv = nn.Linear(embed_dim, 3 * embed_dim, weight_attr=qkv_weight_attr, bias_attr=qkv_bias_attr)
        self.out = nn.Linear(embed_dim, embed_dim, weight_attr=linear_weight_attr, bias_attr=linear_bias_attr)
        self.pre_ln = nn.LayerNorm(embed_dim, epsilon) if normalize_before else None
        self.ln = None if normalize_before else nn.LayerNorm(embed_dim, epsilon)

    def forward(self, query, key=None, value=None, attn_mask=None, cache=None):
        import torch

        from paddle_b200 import ops
        x = self.pre_ln(query) if self.pre_ln is not None else query
        b, s, _ = x.shape
        q, k, v = self.qkv(x).reshape(b, s, 3, self.num_heads, self.head_dim).unbind(2)
        o = ops.flash_attention(q, k, v, dropout=self.attn_dropout_rate, causal=False, attn_mask=attn_mask, training=self.training)
        o = self.out(o.reshape(b, s, self.embed_dim))
        o = query + torch.nn.functional.dropout(o, self.dropout_rate, self.training)
        return o if self.ln is None else self.ln(o)


class FusedFeedForward(Layer):
    def __init__(self, d_model, dim_feedforward, dropout_rate=0.1, epsilon=1e-05, activation="relu", act_dropout_rate=None,
                 normalize_before=False, linear1_weight_attr=None, linear1_bias_attr=None, linear2_weight_attr=None,
                 linear2_bias_attr=None, ln1_scale_attr=None, ln1_bias_attr=None, ln2_scale_attr=None, ln2_bias_attr=None,
                 nranks=1, ring_id=-1, name=None):
        super().__init__()
        import paddle_b200.nn as nn
        self.normalize_before, self.dropout_rate = normalize_before, dropout_rate
        self.act_dropout_rate = dropout_rate if act_dropout_rate is None else act_dropout_rate
        self.linear1 = nn.Linear(d_model, dim_feedforward, weight_attr=linear1_weight_attr, bias_attr=linear1_bias_attr)
        self.linear2 = nn.Linear(dim_feedforward, d_model, weight_attr=linear2_weight_attr, bias_attr=linear2_bias_attr)
        self.norm = nn.LayerNorm(d_model, epsilon)
        self.activation = activation

    def forward(self, src, cache=None):
        import torch

        from paddle_b200.nn import functional as F
        x = self.norm(src) if self.normalize_before else src
        h = torch.nn.functional.dropout(getattr(F, self.activation)(self.linear1(x)), self.act_dropout_rate, self.training)
        y = src + torch.nn.functional.dropout(self.linear2(h), self.dropout_rate, self.training)
        return y if self.normalize_before else self.norm(y)


class FusedTransformerEncoderLayer(Layer):
    def __init__(self, d_model, nhead, dim_feedforward, dropout_rate=0.1, activation="relu", attn_dropout_rate=None,
                 act_dropout_rate=None, normalize_before=False, weight_attr=None, bias_attr=None):
        super().__init__()
        ad = dropout_rate if attn_dropout_rate is None else attn_dropout_rate
        self.fused_attn = FusedMultiHeadAttention(d_model, nhead, dropout_rate, ad, normalize_before=normalize_before)
        self.ffn = FusedFeedForward(d_model, dim_feedforward, dropout_rate, activation=activation,
                                    act_dropout_rate=act_dropout_rate, normalize_before=normalize_before)

    def forward(self, src, src_mask=None, cache=None):
        return self.ffn(self.fused_attn(src, attn_mask=src_mask))


class FusedEcMoe(Layer):
    def __init__(self, hidden_size, inter_size, num_experts, act_type, weight_attr=None, bias_attr=None):
        super().__init__()
        self.act_type = act_type
        self.bmm_weight0 = self.create_parameter([num_experts, hidden_size, inter_size], attr=weight_attr)
        self.bmm_bias0 = self.create_parameter([num_experts, 1, inter_size], attr=bias_attr, is_bias=True)
        self.bmm_weight1 = self.create_parameter([num_experts, inter_size, hidden_size], attr=weight_attr)
        self.bmm_bias1 = self.create_parameter([num_experts, 1, hidden_size], attr=bias_attr, is_bias=True)

    def forward(self, x, gate):
        from .functional import fused_ec_moe
        return fused_ec_moe(x, gate, self.bmm_weight0, self.bmm_bias0, self.bmm_weight1, self.bmm_bias1, self.act_type)


class FusedMultiTransformer(Layer):
    def __init__(self, *a, **k):
        super().__init__()
        raise NotImplementedError("FusedMultiTransformer is the reference's decode-time inference stack; this framework "
                                  "targets training (SURVEY 2.6.b, out of scope)")
