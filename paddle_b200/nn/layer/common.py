"""Common layers (reference: python/paddle/nn/layer/{common,norm,conv,activation,loss,pooling,transformer}.py).
Weight layouts follow the reference: Linear ``[in, out]``, Conv ``[out, in/groups, kh, kw]``, Embedding ``[V, h]``."""
from __future__ import annotations

import math

import torch

from ...framework.param_attr import ParamAttr
from .. import functional as F
from .. import initializer as I
from .layers import Layer, LayerList


class Identity(Layer):
    def forward(self, x):
        return x


class Linear(Layer):
    def __init__(self, in_features, out_features, weight_attr=None, bias_attr=None, name=None):
        super().__init__()
        self._in_features, self._out_features = in_features, out_features
        self.weight = self.create_parameter([in_features, out_features], attr=weight_attr, dtype=self._dtype)
        self.bias = self.create_parameter([out_features], attr=bias_attr, dtype=self._dtype, is_bias=True)
        self.name = name

    def forward(self, input):
        return F.linear(input, self.weight, self.bias)

    def extra_repr(self):
        return f"in_features={self._in_features}, out_features={self._out_features}, dtype={self._dtype}"


class Embedding(Layer):
    def __init__(self, num_embeddings, embedding_dim, padding_idx=None, sparse=False, weight_attr=None, name=None):
        super().__init__()
        self._num_embeddings, self._embedding_dim = num_embeddings, embedding_dim
        self._padding_idx = None if padding_idx is None else (padding_idx if padding_idx >= 0 else num_embeddings + padding_idx)
        self._sparse = bool(sparse)
        # no layer-specific default: create_parameter's own (Xavier uniform over [num_embeddings, embedding_dim]), as in
        # the reference (python/paddle/nn/layer/common.py Embedding -> LayerHelperBase.create_parameter)
        self.weight = self.create_parameter([num_embeddings, embedding_dim], attr=weight_attr, dtype=self._dtype)
        self.weight._is_sparse_grad = self._sparse      # DataParallel's reducer routes it through the sparse path
        if self._padding_idx is not None:
            with torch.no_grad():
                self.weight[self._padding_idx].zero_()

    def forward(self, x):
        return F.embedding(x, self.weight, padding_idx=self._padding_idx, sparse=self._sparse)


class Dropout(Layer):
    def __init__(self, p=0.5, axis=None, mode="upscale_in_train", name=None):
        super().__init__()
        self.p, self.axis, self.mode = p, axis, mode

    def forward(self, input):
        return F.dropout(input, self.p, self.axis, self.training, self.mode)


class Dropout2D(Layer):
    def __init__(self, p=0.5, data_format="NCHW", name=None):
        super().__init__()
        self.p = p

    def forward(self, x):
        return F.dropout2d(x, self.p, self.training)


class Flatten(Layer):
    def __init__(self, start_axis=1, stop_axis=-1):
        super().__init__()
        self.start_axis, self.stop_axis = start_axis, stop_axis

    def forward(self, x):
        return torch.flatten(x, self.start_axis, self.stop_axis)


# ------------------------------------------------------------------ norms
class LayerNorm(Layer):
    def __init__(self, normalized_shape, epsilon=1e-05, weight_attr=None, bias_attr=None, name=None):
        super().__init__()
        if isinstance(normalized_shape, int):
            normalized_shape = [normalized_shape]
        self._normalized_shape = list(normalized_shape)
        self._epsilon = epsilon
        self.weight = None if weight_attr is False else self.create_parameter(
            self._normalized_shape, attr=weight_attr, default_initializer=I.Constant(1.0))
        self.bias = None if bias_attr is False else self.create_parameter(
            self._normalized_shape, attr=bias_attr, is_bias=True)

    def forward(self, input):
        return F.layer_norm(input, self._normalized_shape, self.weight, self.bias, self._epsilon)


class RMSNorm(Layer):
    """y = x * rsqrt(mean(x^2) + eps) * weight  (K18; python/paddle/incubate/nn/functional/fused_rms_norm.py)."""

    def __init__(self, hidden_size, epsilon=1e-6, weight_attr=None, name=None):
        super().__init__()
        self._epsilon = epsilon
        self.weight = self.create_parameter([hidden_size], attr=weight_attr, default_initializer=I.Constant(1.0))

    def forward(self, x):
        return F.rms_norm(x, self.weight, self._epsilon)


class _BatchNormBase(Layer):
    def __init__(self, num_features, momentum=0.9, epsilon=1e-05, weight_attr=None, bias_attr=None, data_format="NCHW",
                 use_global_stats=None, name=None):
        super().__init__()
        self._num_features, self._momentum, self._epsilon = num_features, momentum, epsilon
        self._data_format, self._use_global_stats = data_format, use_global_stats
        self.weight = None if weight_attr is False else self.create_parameter(
            [num_features], attr=weight_attr, default_initializer=I.Constant(1.0))
        self.bias = None if bias_attr is False else self.create_parameter([num_features], attr=bias_attr, is_bias=True)
        self.register_buffer("_mean", torch.zeros(num_features, device=self.weight.device if self.weight is not None else None))
        self.register_buffer("_variance", torch.ones(num_features, device=self.weight.device if self.weight is not None else None))

    def forward(self, input):
        return F.batch_norm(input, self._mean, self._variance, self.weight, self.bias, self.training, self._momentum,
                            self._epsilon, self._data_format, self._use_global_stats)


class BatchNorm1D(_BatchNormBase):
    pass


class BatchNorm2D(_BatchNormBase):
    pass


class BatchNorm3D(_BatchNormBase):
    pass


class BatchNorm(_BatchNormBase):
    def __init__(self, num_channels, act=None, is_test=False, momentum=0.9, epsilon=1e-05, param_attr=None,
                 bias_attr=None, dtype="float32", data_layout="NCHW", in_place=False, moving_mean_name=None,
                 moving_variance_name=None, do_model_average_for_mean_and_var=True, use_global_stats=False,
                 trainable_statistics=False):
        super().__init__(num_channels, momentum, epsilon, param_attr, bias_attr, data_layout, use_global_stats or None)
        self._act = act

    def forward(self, x):
        y = super().forward(x)
        return getattr(F, self._act)(y) if self._act else y


class SyncBatchNorm(_BatchNormBase):
    """Cross-rank batch statistics (reference: python/paddle/nn/layer/norm.py SyncBatchNorm,
    paddle/fluid/operators/sync_batch_norm_utils.h: all-reduce of [2C+1] sums).  K14."""

    def forward(self, x):
        from ... import distributed as dist
        if not self.training or not dist.is_initialized() or dist.get_world_size() == 1:
            return super().forward(x)
        return _SyncBNFn.apply(x, self.weight, self.bias, self._mean, self._variance, self._momentum, self._epsilon)

    @classmethod
    def convert_sync_batchnorm(cls, layer):
        out = layer
        if isinstance(layer, _BatchNormBase) and not isinstance(layer, SyncBatchNorm):
            out = SyncBatchNorm(layer._num_features, layer._momentum, layer._epsilon, data_format=layer._data_format)
            with torch.no_grad():
                if layer.weight is not None:
                    out.weight.copy_(layer.weight)
                    out.bias.copy_(layer.bias)
                out._mean.copy_(layer._mean)
                out._variance.copy_(layer._variance)
        for name, child in list(layer.named_children()):
            out.add_sublayer(name, cls.convert_sync_batchnorm(child))
        return out


def _stats_allreduce(t):
    """K14: the [2C+1] statistics vector is summed with the one-kernel NVLS all-reduce (K1) on a symmetric scratch
    when the world group has an arena; NCCL otherwise."""
    import torch.distributed as tdist
    if t.is_cuda:
        from ...distributed import collective as C
        from ...parallel import symm
        g = C._get_default_group()
        if symm.available(g):
            try:
                n = (t.numel() + 3) // 4 * 4
                buf = symm._staging(symm.get_arena(g), "syncbn", (n,), torch.float32)
                buf[:t.numel()].copy_(t)
                if n != t.numel():
                    buf[t.numel():].zero_()
                symm.allreduce_(buf)
                return buf[:t.numel()].clone()
            except Exception:  # noqa: BLE001  (arena could not be created: fall through to NCCL on every rank alike)
                pass
    tdist.all_reduce(t)
    return t


class _SyncBNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, rmean, rvar, momentum, eps):
        import torch.distributed as tdist
        dims = [0] + list(range(2, x.dim()))
        n_local = x.numel() // x.shape[1]
        xf = x.float()
        stats = torch.cat([xf.sum(dims), (xf * xf).sum(dims), torch.tensor([float(n_local)], device=x.device)])
        stats = _stats_allreduce(stats)
        C = x.shape[1]
        n = stats[-1]
        mean = stats[:C] / n
        var = stats[C:2 * C] / n - mean * mean
        rstd = torch.rsqrt(var + eps)
        with torch.no_grad():
            rmean.mul_(momentum).add_(mean, alpha=1 - momentum)
            rvar.mul_(momentum).add_(var, alpha=1 - momentum)     # biased, as sync_batch_norm_utils.h:95-106
        shp = [1, C] + [1] * (x.dim() - 2)
        xhat = (xf - mean.reshape(shp)) * rstd.reshape(shp)
        y = xhat * w.float().reshape(shp) + b.float().reshape(shp)
        ctx.save_for_backward(xhat, w, rstd, n)
        return y.to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        import torch.distributed as tdist
        xhat, w, rstd, n = ctx.saved_tensors
        dims = [0] + list(range(2, dy.dim()))
        C = dy.shape[1]
        shp = [1, C] + [1] * (dy.dim() - 2)
        dyf = dy.float()
        db = dyf.sum(dims)
        dw = (dyf * xhat).sum(dims)
        red = torch.cat([db, dw])
        red = _stats_allreduce(red)
        mean_dy = (red[:C] / n).reshape(shp)
        mean_dyx = (red[C:] / n).reshape(shp)
        dx = (dyf - mean_dy - xhat * mean_dyx) * (w.float() * rstd).reshape(shp)
        return dx.to(dy.dtype), dw.to(w.dtype), db.to(w.dtype), None, None, None, None


# ------------------------------------------------------------------ conv / pool
# ------------------------------------------------------------------ activations
def _act_layer(name, fn, **defaults):
    class _A(Layer):
        def __init__(self, *args, **kwargs):
            super().__init__()
            kwargs.pop("name", None)
            self._args, self._kwargs = args, {**defaults, **kwargs}

        def forward(self, x):
            return fn(x, *self._args, **self._kwargs)

    _A.__name__ = _A.__qualname__ = name
    return _A


ReLU = _act_layer("ReLU", F.relu)
ReLU6 = _act_layer("ReLU6", F.relu6)
LeakyReLU = _act_layer("LeakyReLU", F.leaky_relu)
ELU = _act_layer("ELU", F.elu)
SELU = _act_layer("SELU", F.selu)
CELU = _act_layer("CELU", F.celu)
GELU = _act_layer("GELU", F.gelu)
Silu = _act_layer("Silu", F.silu)
Swish = _act_layer("Swish", F.silu)
Mish = _act_layer("Mish", F.mish)
Sigmoid = _act_layer("Sigmoid", F.sigmoid)
Hardsigmoid = _act_layer("Hardsigmoid", F.hardsigmoid)
Hardswish = _act_layer("Hardswish", F.hardswish)
Hardtanh = _act_layer("Hardtanh", F.hardtanh)
Tanh = _act_layer("Tanh", F.tanh)
Softplus = _act_layer("Softplus", F.softplus)
Softsign = _act_layer("Softsign", F.softsign)
Softshrink = _act_layer("Softshrink", F.softshrink)
Hardshrink = _act_layer("Hardshrink", F.hardshrink)
Tanhshrink = _act_layer("Tanhshrink", F.tanhshrink)
LogSigmoid = _act_layer("LogSigmoid", F.log_sigmoid)
GLU = _act_layer("GLU", F.glu)


class Softmax(Layer):
    def __init__(self, axis=-1, name=None):
        super().__init__()
        self._axis = axis

    def forward(self, x):
        return F.softmax(x, self._axis)


class LogSoftmax(Layer):
    def __init__(self, axis=-1, name=None):
        super().__init__()
        self._axis = axis

    def forward(self, x):
        return F.log_softmax(x, self._axis)


class PReLU(Layer):
    def __init__(self, num_parameters=1, init=0.25, weight_attr=None, data_format="NCHW", name=None):
        super().__init__()
        self._weight = self.create_parameter([num_parameters], attr=weight_attr, default_initializer=I.Constant(init))

    def forward(self, x):
        return F.prelu(x, self._weight)


# ------------------------------------------------------------------ losses
class CrossEntropyLoss(Layer):
    def __init__(self, weight=None, ignore_index=-100, reduction="mean", soft_label=False, axis=-1, use_softmax=True,
                 label_smoothing=0.0, name=None):
        super().__init__()
        self.weight, self.ignore_index, self.reduction = weight, ignore_index, reduction
        self.soft_label, self.axis, self.use_softmax, self.label_smoothing = soft_label, axis, use_softmax, label_smoothing

    def forward(self, input, label):
        return F.cross_entropy(input, label, self.weight, self.ignore_index, self.reduction, self.soft_label, self.axis,
                               self.use_softmax, self.label_smoothing)


class MSELoss(Layer):
    def __init__(self, reduction="mean"):
        super().__init__()
        self.reduction = reduction

    def forward(self, input, label):
        return F.mse_loss(input, label, self.reduction)


class L1Loss(Layer):
    def __init__(self, reduction="mean", name=None):
        super().__init__()
        self.reduction = reduction

    def forward(self, input, label):
        return F.l1_loss(input, label, self.reduction)


class SmoothL1Loss(Layer):
    def __init__(self, reduction="mean", delta=1.0, name=None):
        super().__init__()
        self.reduction, self.delta = reduction, delta

    def forward(self, input, label):
        return F.smooth_l1_loss(input, label, self.reduction, self.delta)


class NLLLoss(Layer):
    def __init__(self, weight=None, ignore_index=-100, reduction="mean", name=None):
        super().__init__()
        self.weight, self.ignore_index, self.reduction = weight, ignore_index, reduction

    def forward(self, input, label):
        return F.nll_loss(input, label, self.weight, self.ignore_index, self.reduction)


class BCELoss(Layer):
    def __init__(self, weight=None, reduction="mean", name=None):
        super().__init__()
        self.weight, self.reduction = weight, reduction

    def forward(self, input, label):
        return F.binary_cross_entropy(input, label, self.weight, self.reduction)


class BCEWithLogitsLoss(Layer):
    def __init__(self, weight=None, reduction="mean", pos_weight=None, name=None):
        super().__init__()
        self.weight, self.reduction, self.pos_weight = weight, reduction, pos_weight

    def forward(self, logit, label):
        return F.binary_cross_entropy_with_logits(logit, label, self.weight, self.reduction, self.pos_weight)


class KLDivLoss(Layer):
    def __init__(self, reduction="mean", log_target=False):
        super().__init__()
        self.reduction, self.log_target = reduction, log_target

    def forward(self, input, label):
        return F.kl_div(input, label, self.reduction, self.log_target)


class CosineSimilarity(Layer):
    def __init__(self, axis=1, eps=1e-8):
        super().__init__()
        self.axis, self.eps = axis, eps

    def forward(self, x1, x2):
        return F.cosine_similarity(x1, x2, self.axis, self.eps)


# ------------------------------------------------------------------ attention / transformer
