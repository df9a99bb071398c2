"""paddle.nn transformer layers (reference: python/paddle/nn/layer/transformer.py): MultiHeadAttention with the
Cache / StaticCache decoding contract, encoder / decoder layers and stacks, the full Transformer.

Contract points kept from the reference:
  * inputs are ``[batch, seq, embed]``; ``attn_mask`` broadcasts to ``[batch, heads, q_len, k_len]``; a bool / int mask
    keeps positions where it is True / non-zero, a float mask is added to the scores;
  * ``forward`` returns ``out`` alone, or a tuple growing with ``need_weights`` (the softmax weights) and ``cache``;
  * ``Cache`` (k, v of past positions, ``[batch, heads, len, head_dim]``, grown by concatenation) is for self
    attention while decoding, ``StaticCache`` (projected encoder memory, computed once) for cross attention;
  * stacks build layers 1..n-1 from the first layer's constructor arguments - new parameters, freshly initialised;
  * ``weight_attr`` / ``bias_attr`` may be a list with one entry per sub-block (self-attn[, cross-attn], ffn).

Attention runs through ``F.scaled_dot_product_attention`` (the flash kernels on GPU) unless the softmax weights are
asked for, which needs the explicit matmul-softmax-matmul.
"""
from __future__ import annotations

import collections
import copy

import torch

from ...framework.param_attr import ParamAttr
from .. import functional as F
from .common import Dropout, LayerNorm, Linear
from .layers import Layer, LayerList

__all__ = ["MultiHeadAttention", "TransformerEncoderLayer", "TransformerEncoder", "TransformerDecoderLayer",
           "TransformerDecoder", "Transformer"]


def _convert_param_attr_to_list(param_attr, n):
    """One attr per sub-block.  A list is taken entry by entry (short lists repeat their last entry is NOT done: the
    reference asserts the length); a single attr is copied, and a copy with a name gets a per-block suffix so two
    parameters never share a name."""
    if isinstance(param_attr, (list, tuple)):
        assert len(param_attr) == n, f"length of param_attr should be {n} when it is a list/tuple"
        return [ParamAttr._to_attr(a) if a is not False and a is not None else a for a in param_attr]
    if param_attr is None or param_attr is False or isinstance(param_attr, bool):
        return [param_attr if param_attr is not True else None] * n
    attrs = []
    base = ParamAttr._to_attr(param_attr)
    for i in range(n):
        a = copy.deepcopy(base)
        if getattr(a, "name", None):
            a.name = f"{a.name}_{i}"
        attrs.append(a)
    return attrs


def _convert_attention_mask(attn_mask, dtype):
    """bool / int mask -> additive float mask (0 where kept, -1e9 where masked); float masks pass through."""
    if attn_mask is None:
        return None
    if attn_mask.dtype == torch.bool or not attn_mask.dtype.is_floating_point:
        return (attn_mask.to(dtype) - 1.0) * 1e9
    return attn_mask.to(dtype)


class MultiHeadAttention(Layer):
    Cache = collections.namedtuple("Cache", ["k", "v"])
    StaticCache = collections.namedtuple("StaticCache", ["k", "v"])

    def __init__(self, embed_dim, num_heads, dropout=0.0, kdim=None, vdim=None, need_weights=False, weight_attr=None,
                 bias_attr=None):
        super().__init__()
        assert embed_dim > 0, f"Expected embed_dim to be greater than 0, but received {embed_dim}"
        assert num_heads > 0, f"Expected num_heads to be greater than 0, but received {num_heads}"
        self.embed_dim = embed_dim
        self.kdim = kdim if kdim is not None else embed_dim
        self.vdim = vdim if vdim is not None else embed_dim
        self.num_heads, self.dropout, self.need_weights = num_heads, dropout, need_weights
        self.head_dim = embed_dim // num_heads
        assert self.head_dim * num_heads == embed_dim, "embed_dim must be divisible by num_heads"
        self.q_proj = Linear(embed_dim, embed_dim, weight_attr, bias_attr=bias_attr)
        self.k_proj = Linear(self.kdim, embed_dim, weight_attr, bias_attr=bias_attr)
        self.v_proj = Linear(self.vdim, embed_dim, weight_attr, bias_attr=bias_attr)
        self.out_proj = Linear(embed_dim, embed_dim, weight_attr, bias_attr=bias_attr)

    # [b, s, e] -> [b, heads, s, head_dim]
    def _split(self, x):
        b, s, _ = x.shape
        return x.reshape(b, s, self.num_heads, self.head_dim).transpose(1, 2)

    def compute_kv(self, key, value):
        return self._split(self.k_proj(key)), self._split(self.v_proj(value))

    def gen_cache(self, key, value=None, type=Cache):
        """``StaticCache``: project ``key`` / ``value`` (the encoder memory) once.  ``Cache`` with ``value=None``: an
        empty past for ``key``'s batch.  ``Cache`` with both: wrap given k / v."""
        if type == MultiHeadAttention.StaticCache:
            k, v = self.compute_kv(key, key if value is None else value)
            return self.StaticCache(k, v)
        if value is None:
            shape = (key.shape[0], self.num_heads, 0, self.head_dim)
            return self.Cache(key.new_zeros(shape), key.new_zeros(shape))
        return self.Cache(key, value)

    def _prepare_qkv(self, query, key, value, cache):
        q = self._split(self.q_proj(query))
        if isinstance(cache, self.StaticCache):
            k, v = cache.k, cache.v
        else:
            k, v = self.compute_kv(key, value)
        if isinstance(cache, self.Cache):
            k = torch.cat([cache.k, k], dim=2)
            v = torch.cat([cache.v, v], dim=2)
            cache = self.Cache(k, v)
        return q, k, v, cache

    def forward(self, query, key=None, value=None, attn_mask=None, cache=None):
        key = query if key is None else key
        value = query if value is None else value
        q, k, v, cache = self._prepare_qkv(query, key, value, cache)
        mask = _convert_attention_mask(attn_mask, q.dtype)
        if self.need_weights:
            scores = torch.matmul(q * (self.head_dim ** -0.5), k.transpose(-1, -2))
            if mask is not None:
                scores = scores + mask
            weights = torch.softmax(scores, dim=-1)
            if self.dropout:
                weights = F.dropout(weights, self.dropout, training=self.training, mode="upscale_in_train")
            out = torch.matmul(weights, v).transpose(1, 2)
        else:
            weights = None
            out = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), mask,
                                                 self.dropout, False, self.training)
        b, s = out.shape[0], out.shape[1]
        out = self.out_proj(out.reshape(b, s, self.embed_dim))
        outs = [out]
        if self.need_weights:
            outs.append(weights)
        if cache is not None:
            outs.append(cache)
        return out if len(outs) == 1 else tuple(outs)


def _activation(name):
    fn = getattr(F, name, None)
    if fn is None:
        raise ValueError(f"unknown activation {name!r}")
    return fn


class TransformerEncoderLayer(Layer):
    def __init__(self, d_model, nhead, dim_feedforward, dropout=0.1, activation="relu", attn_dropout=None,
                 act_dropout=None, normalize_before=False, weight_attr=None, bias_attr=None, layer_norm_eps=1e-5):
        config = {k: v for k, v in locals().items() if k not in ("self", "__class__")}
        super().__init__()
        self._config = config
        assert d_model > 0 and nhead > 0 and dim_feedforward > 0
        attn_dropout = dropout if attn_dropout is None else attn_dropout
        act_dropout = dropout if act_dropout is None else act_dropout
        self.normalize_before = normalize_before
        w = _convert_param_attr_to_list(weight_attr, 2)
        b = _convert_param_attr_to_list(bias_attr, 2)
        self.self_attn = MultiHeadAttention(d_model, nhead, dropout=attn_dropout, weight_attr=w[0], bias_attr=b[0])
        self.linear1 = Linear(d_model, dim_feedforward, w[1], bias_attr=b[1])
        self.dropout = Dropout(act_dropout, mode="upscale_in_train")
        self.linear2 = Linear(dim_feedforward, d_model, w[1], bias_attr=b[1])
        self.norm1 = LayerNorm(d_model, layer_norm_eps)
        self.norm2 = LayerNorm(d_model, layer_norm_eps)
        self.dropout1 = Dropout(dropout, mode="upscale_in_train")
        self.dropout2 = Dropout(dropout, mode="upscale_in_train")
        self.activation = _activation(activation)

    def forward(self, src, src_mask=None, cache=None):
        residual = src
        if self.normalize_before:
            src = self.norm1(src)
        if cache is None:
            src = self.self_attn(src, src, src, src_mask)
        else:
            src, incremental_cache = self.self_attn(src, src, src, src_mask, cache)
        src = residual + self.dropout1(src)
        if not self.normalize_before:
            src = self.norm1(src)
        residual = src
        if self.normalize_before:
            src = self.norm2(src)
        src = self.linear2(self.dropout(self.activation(self.linear1(src))))
        src = residual + self.dropout2(src)
        if not self.normalize_before:
            src = self.norm2(src)
        return src if cache is None else (src, incremental_cache)

    def gen_cache(self, src):
        return self.self_attn.gen_cache(src, type=self.self_attn.Cache)


def _stack(first, num_layers):
    """[first] + (num_layers - 1) new layers built from ``first``'s constructor arguments."""
    return LayerList([first if i == 0 else type(first)(**first._config) for i in range(num_layers)])


class TransformerEncoder(Layer):
    def __init__(self, encoder_layer, num_layers, norm=None):
        super().__init__()
        self.layers = _stack(encoder_layer, num_layers)
        self.num_layers, self.norm = num_layers, norm

    def forward(self, src, src_mask=None, cache=None):
        output, new_caches = src, []
        for i, mod in enumerate(self.layers):
            if cache is None:
                output = mod(output, src_mask=src_mask)
            else:
                output, new_cache = mod(output, src_mask=src_mask, cache=cache[i])
                new_caches.append(new_cache)
        if self.norm is not None:
            output = self.norm(output)
        return output if cache is None else (output, new_caches)

    def gen_cache(self, src):
        return [layer.gen_cache(src) for layer in self.layers]


class TransformerDecoderLayer(Layer):
    def __init__(self, d_model, nhead, dim_feedforward, dropout=0.1, activation="relu", attn_dropout=None,
                 act_dropout=None, normalize_before=False, weight_attr=None, bias_attr=None, layer_norm_eps=1e-5):
        config = {k: v for k, v in locals().items() if k not in ("self", "__class__")}
        super().__init__()
        self._config = config
        assert d_model > 0 and nhead > 0 and dim_feedforward > 0
        attn_dropout = dropout if attn_dropout is None else attn_dropout
        act_dropout = dropout if act_dropout is None else act_dropout
        self.normalize_before = normalize_before
        w = _convert_param_attr_to_list(weight_attr, 3)
        b = _convert_param_attr_to_list(bias_attr, 3)
        self.self_attn = MultiHeadAttention(d_model, nhead, dropout=attn_dropout, weight_attr=w[0], bias_attr=b[0])
        self.cross_attn = MultiHeadAttention(d_model, nhead, dropout=attn_dropout, weight_attr=w[1], bias_attr=b[1])
        self.linear1 = Linear(d_model, dim_feedforward, w[2], bias_attr=b[2])
        self.dropout = Dropout(act_dropout, mode="upscale_in_train")
        self.linear2 = Linear(dim_feedforward, d_model, w[2], bias_attr=b[2])
        self.norm1 = LayerNorm(d_model, layer_norm_eps)
        self.norm2 = LayerNorm(d_model, layer_norm_eps)
        self.norm3 = LayerNorm(d_model, layer_norm_eps)
        self.dropout1 = Dropout(dropout, mode="upscale_in_train")
        self.dropout2 = Dropout(dropout, mode="upscale_in_train")
        self.dropout3 = Dropout(dropout, mode="upscale_in_train")
        self.activation = _activation(activation)

    def forward(self, tgt, memory, tgt_mask=None, memory_mask=None, cache=None):
        """``cache`` = (incremental ``Cache`` of the self attention, ``StaticCache`` of the cross attention)."""
        residual = tgt
        if self.normalize_before:
            tgt = self.norm1(tgt)
        if cache is None:
            tgt = self.self_attn(tgt, tgt, tgt, tgt_mask, None)
        else:
            tgt, incremental_cache = self.self_attn(tgt, tgt, tgt, tgt_mask, cache[0])
        tgt = residual + self.dropout1(tgt)
        if not self.normalize_before:
            tgt = self.norm1(tgt)
        residual = tgt
        if self.normalize_before:
            tgt = self.norm2(tgt)
        if cache is None:
            tgt = self.cross_attn(tgt, memory, memory, memory_mask, None)
        else:
            tgt, static_cache = self.cross_attn(tgt, memory, memory, memory_mask, cache[1])
        tgt = residual + self.dropout2(tgt)
        if not self.normalize_before:
            tgt = self.norm2(tgt)
        residual = tgt
        if self.normalize_before:
            tgt = self.norm3(tgt)
        tgt = self.linear2(self.dropout(self.activation(self.linear1(tgt))))
        tgt = residual + self.dropout3(tgt)
        if not self.normalize_before:
            tgt = self.norm3(tgt)
        return tgt if cache is None else (tgt, (incremental_cache, static_cache))

    def gen_cache(self, memory):
        incremental = self.self_attn.gen_cache(memory, type=self.self_attn.Cache)
        static = self.cross_attn.gen_cache(memory, memory, type=self.cross_attn.StaticCache)
        return incremental, static


class TransformerDecoder(Layer):
    def __init__(self, decoder_layer, num_layers, norm=None):
        super().__init__()
        self.layers = _stack(decoder_layer, num_layers)
        self.num_layers, self.norm = num_layers, norm

    def forward(self, tgt, memory, tgt_mask=None, memory_mask=None, cache=None):
        output, new_caches = tgt, []
        for i, mod in enumerate(self.layers):
            if cache is None:
                output = mod(output, memory, tgt_mask=tgt_mask, memory_mask=memory_mask, cache=None)
            else:
                output, new_cache = mod(output, memory, tgt_mask=tgt_mask, memory_mask=memory_mask, cache=cache[i])
                new_caches.append(new_cache)
        if self.norm is not None:
            output = self.norm(output)
        return output if cache is None else (output, new_caches)

    def gen_cache(self, memory, do_zip=False):
        """One (incremental, static) pair per layer; ``do_zip`` regroups them as (all incrementals, all statics)."""
        cache = [layer.gen_cache(memory) for layer in self.layers]
        return list(zip(*cache)) if do_zip else cache


class Transformer(Layer):
    """Encoder-decoder of "Attention is all you need".  Both stacks end in a LayerNorm.  ``weight_attr`` / ``bias_attr``
    lists: 1 entry = every block; 2 = (attention blocks, ffn); 3 = (self-attn, cross-attn, ffn) - the encoder takes
    the first and the last."""

    def __init__(self, d_model=512, nhead=8, num_encoder_layers=6, num_decoder_layers=6, dim_feedforward=2048,
                 dropout=0.1, activation="relu", attn_dropout=None, act_dropout=None, normalize_before=False,
                 weight_attr=None, bias_attr=None, custom_encoder=None, custom_decoder=None):
        super().__init__()
        assert d_model > 0 and nhead > 0 and dim_feedforward > 0

        def split(attr):
            if isinstance(attr, (list, tuple)):
                if len(attr) == 1:
                    return [attr[0]] * 2, [attr[0]] * 3
                if len(attr) == 2:
                    return list(attr), [attr[0], attr[0], attr[-1]]
                if len(attr) == 3:
                    return [attr[0], attr[-1]], list(attr)
                raise AssertionError("length of bias_attr / weight_attr should be 1 or 2 or 3 when it is a list/tuple")
            return attr, attr
        enc_w, dec_w = split(weight_attr)
        enc_b, dec_b = split(bias_attr)
        if custom_encoder is not None:
            self.encoder = custom_encoder
        else:
            layer = TransformerEncoderLayer(d_model, nhead, dim_feedforward, dropout, activation, attn_dropout,
                                            act_dropout, normalize_before, enc_w, enc_b)
            self.encoder = TransformerEncoder(layer, num_encoder_layers, LayerNorm(d_model))
        if custom_decoder is not None:
            self.decoder = custom_decoder
        else:
            layer = TransformerDecoderLayer(d_model, nhead, dim_feedforward, dropout, activation, attn_dropout,
                                            act_dropout, normalize_before, dec_w, dec_b)
            self.decoder = TransformerDecoder(layer, num_decoder_layers, LayerNorm(d_model))
        self.d_model, self.nhead = d_model, nhead

    def forward(self, src, tgt, src_mask=None, tgt_mask=None, memory_mask=None):
        memory = self.encoder(src, src_mask=src_mask)
        return self.decoder(tgt, memory, tgt_mask=tgt_mask, memory_mask=memory_mask)

    @staticmethod
    def generate_square_subsequent_mask(length):
        """[length, length] float mask: 0 on and below the diagonal, -inf above (a position sees itself and the past)."""
        return torch.triu(torch.full((int(length), int(length)), float("-inf"), dtype=torch.get_default_dtype()), 1)
