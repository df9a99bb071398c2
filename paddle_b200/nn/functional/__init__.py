"""paddle.nn.functional (reference: python/paddle/nn/functional/*.py).

Generic ops map to torch; the hot set routes to the hand-written sm_100a kernels in
``paddle_b200.ops`` when the input lives on a GPU: flash attention, fused softmax-cross-entropy, rms/layer norm,
swiglu, embedding gather stays on torch.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as TF

from ...framework import dtype as _dt
from . import flash_attention  # noqa: F401  (module, callable)
from .flash_attention import (flash_attn_unpadded, scaled_dot_product_attention,  # noqa: F401
                              flash_attn_qkvpacked, flash_attn_varlen_qkvpacked, flash_attention_with_sparse_mask,
                              sdp_kernel)


# ------------------------------------------------------------------ linear / embedding
def linear(x, weight, bias=None, name=None):
    """``x @ weight + bias`` with Paddle's ``[in, out]`` weight layout (nn/functional/common.py:linear)."""
    from ... import ops
    return ops.linear(x, weight, bias)


def embedding(x, weight, padding_idx=None, sparse=False, name=None, max_norm=None, norm_type=2.0, scale_grad_by_freq=False):
    if padding_idx is not None and padding_idx < 0:
        padding_idx = weight.shape[0] + padding_idx
    ids = x.long() if x.dtype not in (torch.int32, torch.int64) else x
    out = TF.embedding(ids, weight, padding_idx=padding_idx, sparse=bool(sparse))
    if padding_idx is not None:
        # the reference returns zeros for padding ids whatever the table row holds (and sends it no gradient)
        out = out * (ids != padding_idx).unsqueeze(-1).to(out.dtype)
    return out


def one_hot(x, num_classes, name=None):
    return TF.one_hot(x.long(), num_classes).to(_dt.default_torch_dtype())


def bilinear(x1, x2, weight, bias=None, name=None):
    return TF.bilinear(x1, x2, weight, bias)


# ------------------------------------------------------------------ activations
def relu(x, name=None):
    return TF.relu(x)


def relu_(x, name=None):
    return TF.relu_(x)


def relu6(x, name=None):
    return TF.relu6(x)


def leaky_relu(x, negative_slope=0.01, name=None):
    return TF.leaky_relu(x, negative_slope)


def elu(x, alpha=1.0, name=None):
    return TF.elu(x, alpha)


def selu(x, scale=1.0507009873554804934193349852946, alpha=1.6732632423543772848170429916717, name=None):
    return scale * torch.where(x > 0, x, alpha * (torch.exp(x) - 1))


def celu(x, alpha=1.0, name=None):
    return TF.celu(x, alpha)


def gelu(x, approximate=False, name=None):
    return TF.gelu(x, approximate="tanh" if approximate else "none")


def silu(x, name=None):
    return TF.silu(x)


swish = silu


def mish(x, name=None):
    return TF.mish(x)


def sigmoid(x, name=None):
    return torch.sigmoid(x)


def hardsigmoid(x, slope=0.1666667, offset=0.5, name=None):
    return torch.clamp(x * slope + offset, 0.0, 1.0)


def hardswish(x, name=None):
    return TF.hardswish(x)


def hardtanh(x, min=-1.0, max=1.0, name=None):
    return TF.hardtanh(x, min, max)


def tanh(x, name=None):
    return torch.tanh(x)


def softplus(x, beta=1, threshold=20, name=None):
    return TF.softplus(x, beta, threshold)


def softsign(x, name=None):
    return TF.softsign(x)


def softshrink(x, threshold=0.5, name=None):
    return TF.softshrink(x, threshold)


def hardshrink(x, threshold=0.5, name=None):
    return TF.hardshrink(x, threshold)


def tanhshrink(x, name=None):
    return TF.tanhshrink(x)


def thresholded_relu(x, threshold=1.0, value=0.0, name=None):
    return torch.where(x > threshold, x, torch.full_like(x, value))


def prelu(x, weight, data_format="NCHW", name=None):
    return TF.prelu(x, weight)


def log_sigmoid(x, name=None):
    return TF.logsigmoid(x)


def maxout(x, groups, axis=1, name=None):
    shp = list(x.shape)
    c = shp[axis]
    new = shp[:axis] + [c // groups, groups] + shp[axis + 1:]
    return x.reshape(new).amax(axis + 1)


def glu(x, axis=-1, name=None):
    return TF.glu(x, axis)


def swiglu(x, y=None, name=None):
    from ... import ops
    return ops.swiglu(x, y)


def softmax(x, axis=-1, dtype=None, name=None):
    return TF.softmax(x, dim=axis, dtype=_dt.to_torch_dtype(dtype))


def log_softmax(x, axis=-1, dtype=None, name=None):
    return TF.log_softmax(x, dim=axis, dtype=_dt.to_torch_dtype(dtype))


def gumbel_softmax(x, temperature=1.0, hard=False, axis=-1, name=None):
    return TF.gumbel_softmax(x, tau=temperature, hard=hard, dim=axis)


# ------------------------------------------------------------------ dropout / norm
def dropout(x, p=0.5, axis=None, training=True, mode="upscale_in_train", name=None):
    if not training or p == 0.0:
        return x if mode == "upscale_in_train" or not training and mode == "upscale_in_train" else (
            x * (1.0 - p) if not training else x)
    if axis is not None:
        axes = [axis] if isinstance(axis, int) else list(axis)
        mshape = [s if i in axes else 1 for i, s in enumerate(x.shape)]
        mask = (torch.rand(mshape, device=x.device) >= p).to(x.dtype)
        out = x * mask
        return out / (1.0 - p) if mode == "upscale_in_train" else out
    if mode == "upscale_in_train":
        return TF.dropout(x, p, training=True)
    return TF.dropout(x, p, training=True) * (1.0 - p)


def dropout2d(x, p=0.5, training=True, data_format="NCHW", name=None):
    return TF.dropout2d(x, p, training)


def dropout3d(x, p=0.5, training=True, data_format="NCDHW", name=None):
    return TF.dropout3d(x, p, training)


def alpha_dropout(x, p=0.5, training=True, name=None):
    return TF.alpha_dropout(x, p, training)


def layer_norm(x, normalized_shape, weight=None, bias=None, epsilon=1e-05, name=None):
    if isinstance(normalized_shape, int):
        normalized_shape = [normalized_shape]
    from ... import ops
    return ops.layer_norm(x, list(normalized_shape), weight, bias, epsilon)


def rms_norm(x, weight, epsilon=1e-6):
    from ... import ops
    return ops.rms_norm(x, weight, epsilon)


def normalize(x, p=2, axis=1, epsilon=1e-12, name=None):
    return TF.normalize(x, p=p, dim=axis, eps=epsilon)


# ------------------------------------------------------------------ conv / pool
def grid_sample(x, grid, mode="bilinear", padding_mode="zeros", align_corners=True, name=None):
    return TF.grid_sample(x, grid, mode, padding_mode, align_corners)


# ------------------------------------------------------------------ losses
def _reduce(loss, reduction):
    if reduction == "mean":
        return loss.mean()
    if reduction == "sum":
        return loss.sum()
    return loss


def cross_entropy(input, label, weight=None, ignore_index=-100, reduction="mean", soft_label=False, axis=-1,
                  use_softmax=True, label_smoothing=0.0, name=None):
    """Reference: python/paddle/nn/functional/loss.py:cross_entropy.  Hard labels may carry a trailing 1-dim;
    ``mean`` divides by the number of non-ignored rows."""
    from ... import ops
    if axis not in (-1, input.dim() - 1):
        input = input.movedim(axis, -1)
        if soft_label:
            label = label.movedim(axis, -1)
    if soft_label:
        logp = TF.log_softmax(input.float(), dim=-1) if use_softmax else torch.log(input.float())
        loss = -(label.float() * logp).sum(-1)
        if weight is not None:
            loss = loss * (label.float() * weight.float()).sum(-1)
        return _reduce(loss, reduction).to(input.dtype) if reduction != "none" else loss.unsqueeze(-1).to(input.dtype)
    lab = label
    if lab.dim() == input.dim() and lab.shape[-1] == 1:
        lab = lab.squeeze(-1)
    lab = lab.long()
    if not use_softmax:
        logp = torch.log(input.float())
        loss = TF.nll_loss(logp.reshape(-1, logp.shape[-1]), lab.reshape(-1), weight=weight, ignore_index=ignore_index,
                           reduction="none").reshape(lab.shape)
    elif weight is None and label_smoothing == 0.0:
        loss = ops.softmax_cross_entropy(input, lab, ignore_index)
    else:
        loss = TF.cross_entropy(input.reshape(-1, input.shape[-1]).float(), lab.reshape(-1), weight=weight,
                                ignore_index=ignore_index, reduction="none",
                                label_smoothing=label_smoothing).reshape(lab.shape)
    if reduction == "none":
        return loss.unsqueeze(-1) if label.dim() == input.dim() else loss
    if reduction == "sum":
        return loss.sum()
    valid = (lab != ignore_index)
    if weight is not None:
        denom = (weight.float()[lab.clamp(min=0)] * valid).sum()
    else:
        denom = valid.sum()
    return loss.sum() / denom.clamp(min=1).to(loss.dtype)


def softmax_with_cross_entropy(logits, label, soft_label=False, ignore_index=-100, numeric_stable_mode=True,
                               return_softmax=False, axis=-1):
    loss = cross_entropy(logits, label, soft_label=soft_label, ignore_index=ignore_index, reduction="none", axis=axis)
    if loss.dim() < logits.dim():
        loss = loss.unsqueeze(axis)
    if return_softmax:
        return loss, TF.softmax(logits, dim=axis)
    return loss


def nll_loss(input, label, weight=None, ignore_index=-100, reduction="mean", name=None):
    return TF.nll_loss(input, label.long(), weight, ignore_index=ignore_index, reduction=reduction)


def mse_loss(input, label, reduction="mean", name=None):
    return TF.mse_loss(input, label, reduction=reduction)


def l1_loss(input, label, reduction="mean", name=None):
    return TF.l1_loss(input, label, reduction=reduction)


def smooth_l1_loss(input, label, reduction="mean", delta=1.0, name=None):
    return TF.huber_loss(input, label, reduction=reduction, delta=delta)


def binary_cross_entropy(input, label, weight=None, reduction="mean", name=None):
    return TF.binary_cross_entropy(input, label, weight, reduction=reduction)


def binary_cross_entropy_with_logits(logit, label, weight=None, reduction="mean", pos_weight=None, name=None):
    return TF.binary_cross_entropy_with_logits(logit, label, weight, reduction=reduction, pos_weight=pos_weight)


def kl_div(input, label, reduction="mean", log_target=False, name=None):
    return TF.kl_div(input, label, reduction=reduction, log_target=log_target)


def cosine_similarity(x1, x2, axis=1, eps=1e-8):
    return TF.cosine_similarity(x1, x2, dim=axis, eps=eps)


def margin_ranking_loss(input, other, label, margin=0.0, reduction="mean", name=None):
    return TF.margin_ranking_loss(input, other, label, margin=margin, reduction=reduction)


def sigmoid_focal_loss(logit, label, normalizer=None, alpha=0.25, gamma=2.0, reduction="sum", name=None):
    p = torch.sigmoid(logit)
    ce = TF.binary_cross_entropy_with_logits(logit, label, reduction="none")
    p_t = p * label + (1 - p) * (1 - label)
    loss = ce * ((1 - p_t) ** gamma) * (alpha * label + (1 - alpha) * (1 - label))
    if normalizer is not None:
        loss = loss / normalizer
    return _reduce(loss, reduction)


def square_error_cost(input, label):
    return (input - label) ** 2


def label_smooth(label, prior_dist=None, epsilon=0.1, name=None):
    k = label.shape[-1]
    if prior_dist is None:
        return (1 - epsilon) * label + epsilon / k
    return (1 - epsilon) * label + epsilon * prior_dist


def ctc_loss(log_probs, labels, input_lengths, label_lengths, blank=0, reduction="mean", norm_by_times=False):
    return TF.ctc_loss(log_probs, labels, input_lengths, label_lengths, blank, reduction)


def sequence_mask(x, maxlen=None, dtype="int64", name=None):
    maxlen = int(maxlen if maxlen is not None else x.max().item())
    return (torch.arange(maxlen, device=x.device) < x.unsqueeze(-1)).to(_dt.to_torch_dtype(dtype))


from .extra import *  # noqa: E402,F401,F403
from .conv_pool import *  # noqa: E402,F401,F403
from .extra import _inplace as _mk_inplace  # noqa: E402

elu_, hardtanh_, leaky_relu_, softmax_, tanh_, thresholded_relu_ = (_mk_inplace(f) for f in (elu, hardtanh, leaky_relu, softmax, tanh, thresholded_relu))
