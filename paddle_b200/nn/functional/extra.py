"""Long tail of paddle.nn.functional (reference: python/paddle/nn/functional/{pooling,loss,activation,common,vision,
distance,extension}.py) on torch.nn.functional, Paddle argument names kept."""
from __future__ import annotations

import torch
import torch.nn.functional as TF


def pairwise_distance(x, y, p=2.0, epsilon=1e-6, keepdim=False, name=None):
    return TF.pairwise_distance(x, y, p, epsilon, keepdim)


def _inplace(fn):
    def f(x, *a, **k):
        with torch.no_grad():
            x.copy_(fn(x, *a, **k))
        return x
    f.__name__ = fn.__name__ + "_"
    return f


def fractional_max_pool2d(x, output_size, kernel_size=None, random_u=None, return_mask=False, name=None):
    return TF.fractional_max_pool2d(x, kernel_size or 2, output_size=output_size, return_indices=return_mask)


def fractional_max_pool3d(x, output_size, kernel_size=None, random_u=None, return_mask=False, name=None):
    return TF.fractional_max_pool3d(x, kernel_size or 2, output_size=output_size, return_indices=return_mask)


def _red(loss, reduction):
    return loss.mean() if reduction == "mean" else loss.sum() if reduction == "sum" else loss


def dice_loss(input, label, epsilon=1e-5, name=None):
    lab = TF.one_hot(label.squeeze(-1).long(), input.shape[-1]).to(input.dtype)
    dims = tuple(range(1, input.dim()))
    inter = (input * lab).sum(dims)
    return (1 - 2 * inter / (input.sum(dims) + lab.sum(dims) + epsilon)).mean()


def log_loss(input, label, epsilon=1e-4, name=None):
    return -label * torch.log(input + epsilon) - (1 - label) * torch.log(1 - input + epsilon)


def hsigmoid_loss(input, label, num_classes, weight, bias=None, path_table=None, path_code=None, is_sparse=False, name=None):
    from ..layer.extra import HSigmoidLoss
    m = HSigmoidLoss.__new__(HSigmoidLoss)
    torch.nn.Module.__init__(m)
    m.num_classes, m.weight, m.bias = num_classes, weight, bias
    return HSigmoidLoss.forward(m, input, label)


def multi_label_soft_margin_loss(input, label, weight=None, reduction="mean", name=None):
    return TF.multilabel_soft_margin_loss(input, label, weight, reduction=reduction)


def poisson_nll_loss(input, label, log_input=True, full=False, epsilon=1e-8, reduction="mean", name=None):
    return TF.poisson_nll_loss(input, label, log_input, full, eps=epsilon, reduction=reduction)


def gaussian_nll_loss(input, label, variance, full=False, epsilon=1e-6, reduction="mean", name=None):
    return TF.gaussian_nll_loss(input, label, variance, full, epsilon, reduction)


def npair_loss(anchor, positive, labels, l2_reg=0.002):
    lab = labels.reshape(-1, 1).float()
    same = (lab == lab.t()).float()
    same = same / same.sum(1, keepdim=True)
    l2 = (anchor.pow(2).sum(1).mean() + positive.pow(2).sum(1).mean()) * 0.25 * l2_reg
    logits = anchor @ positive.t()
    return (-(same * torch.log_softmax(logits, 1)).sum(1)).mean() + l2


def hinge_embedding_loss(input, label, margin=1.0, reduction="mean", name=None):
    return TF.hinge_embedding_loss(input, label, margin, reduction=reduction)


def cosine_embedding_loss(input1, input2, label, margin=0, reduction="mean", name=None):
    return TF.cosine_embedding_loss(input1, input2, label, margin, reduction=reduction)


def triplet_margin_loss(input, positive, negative, margin=1.0, p=2, epsilon=1e-6, swap=False, reduction="mean", name=None):
    return TF.triplet_margin_loss(input, positive, negative, margin, p, epsilon, swap, reduction=reduction)


def triplet_margin_with_distance_loss(input, positive, negative, distance_function=None, margin=1.0, swap=False,
                                      reduction="mean", name=None):
    return TF.triplet_margin_with_distance_loss(input, positive, negative, distance_function=distance_function,
                                                margin=margin, swap=swap, reduction=reduction)


def multi_margin_loss(input, label, p=1, margin=1.0, weight=None, reduction="mean", name=None):
    return TF.multi_margin_loss(input, label.long(), p, margin, weight, reduction=reduction)


def soft_margin_loss(input, label, reduction="mean", name=None):
    return TF.soft_margin_loss(input, label.to(input.dtype), reduction=reduction)


def rnnt_loss(input, label, input_lengths, label_lengths, blank=0, fastemit_lambda=0.001, reduction="mean", name=None):
    raise NotImplementedError("rnnt_loss needs the warp-transducer kernels, which this framework does not ship")


def margin_cross_entropy(logits, label, margin1=1.0, margin2=0.5, margin3=0.0, scale=64.0, group=None,
                         return_softmax=False, reduction="mean"):
    """ArcFace-style margin softmax on cosine logits (single-rank form of the model-parallel op)."""
    lab = label.reshape(-1).long()
    theta = torch.acos(logits.float().clamp(-1 + 1e-7, 1 - 1e-7))
    tgt = torch.cos(margin1 * theta + margin2) - margin3
    onehot = TF.one_hot(lab, logits.shape[-1]).bool()
    out = torch.where(onehot, tgt, logits.float()) * scale
    loss = _red(TF.cross_entropy(out, lab, reduction="none").unsqueeze(-1), reduction)
    return (loss, torch.softmax(out, -1)) if return_softmax else loss


def class_center_sample(label, num_classes, num_samples, group=None):
    pos = torch.unique(label)
    if pos.numel() < num_samples:
        rest = torch.tensor(sorted(set(range(num_classes)) - set(pos.tolist())), device=label.device, dtype=label.dtype)
        extra = rest[torch.randperm(rest.numel(), device=label.device)[: num_samples - pos.numel()]]
        sampled = torch.sort(torch.cat([pos, extra])).values
    else:
        sampled = pos
    remap = torch.full((num_classes,), -1, dtype=label.dtype, device=label.device)
    remap[sampled] = torch.arange(sampled.numel(), dtype=label.dtype, device=label.device)
    return remap[label], sampled


def affine_grid(theta, out_shape, align_corners=True, name=None):
    return TF.affine_grid(theta, list(out_shape), align_corners=align_corners)


def gather_tree(ids, parents):
    """Back-trace beam-search ids [T, B, beam] through the parent pointers."""
    T = ids.shape[0]
    out = torch.empty_like(ids)
    out[T - 1] = ids[T - 1]
    beam = parents[T - 1]
    for t in range(T - 2, -1, -1):
        out[t] = ids[t].gather(-1, beam)
        beam = parents[t].gather(-1, beam)
    return out


def temporal_shift(x, seg_num, shift_ratio=0.25, name=None, data_format="NCHW"):
    nt, c, h, w = x.shape
    x5 = x.reshape(nt // seg_num, seg_num, c, h, w)
    c1, c2 = int(c * shift_ratio), int(c * 2 * shift_ratio)
    out = torch.zeros_like(x5)
    out[:, :-1, :c1] = x5[:, 1:, :c1]
    out[:, 1:, c1:c2] = x5[:, :-1, c1:c2]
    out[:, :, c2:] = x5[:, :, c2:]
    return out.reshape(nt, c, h, w)


def sparse_attention(query, key, value, sparse_csr_offset, sparse_csr_columns, key_padding_mask=None, attn_mask=None, name=None):
    """Block-sparse attention given a CSR sparsity pattern per (batch, head): evaluated densely with the pattern as mask."""
    b, h, s, d = query.shape
    mask = torch.zeros(b, h, s, s, dtype=torch.bool, device=query.device)
    for bi in range(b):
        for hi in range(h):
            off, cols = sparse_csr_offset[bi, hi].long(), sparse_csr_columns[bi, hi].long()
            for r in range(s):
                mask[bi, hi, r, cols[off[r]:off[r + 1]]] = True
    sc = (query @ key.transpose(-1, -2)) / d ** 0.5
    sc = sc.masked_fill(~mask, float("-inf"))
    return torch.softmax(sc, -1) @ value


def rrelu(x, lower=1.0 / 8.0, upper=1.0 / 3.0, training=True, name=None):
    return TF.rrelu(x, lower, upper, training)


def adaptive_log_softmax_with_loss(input, label, head_weight, tail_weights, cutoffs, head_bias=None, name=None):
    raise NotImplementedError("adaptive_log_softmax_with_loss: use paddle.nn.AdaptiveLogSoftmaxWithLoss")
