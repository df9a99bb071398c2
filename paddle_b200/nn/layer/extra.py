"""Less common ``paddle.nn`` layers (reference: python/paddle/nn/layer/{rnn,transformer,conv,pooling,common,loss,
activation,vision,distance}.py).  None of them is on a hot path of the training product, so they are thin layers over
torch's implementations that keep Paddle's constructor arguments, parameter naming scheme and return conventions."""
from __future__ import annotations

import torch
import torch.nn.functional as TF

from .. import functional as F
from .common import Dropout, LayerNorm, Linear, _act_layer
from .layers import Layer, LayerList

# ---------------------------------------------------------------------------------------------- activations
SiLU = _act_layer("SiLU", F.silu)
ThresholdedReLU = _act_layer("ThresholdedReLU", F.thresholded_relu)
Maxout = _act_layer("Maxout", F.maxout)


class Softmax2D(Layer):
    def forward(self, x):
        assert x.dim() in (3, 4), "Softmax2D requires a 3D or 4D tensor as input"
        return torch.softmax(x, dim=-3)


def _adopt(layer, inner, attr):
    """Keep a torch module that does the computing OUT of the sublayer registry and make its parameters / buffers this
    layer's own, under their own names - so that state dicts carry the reference's keys ("weight", "weight_ih", ...)
    instead of a wrapper prefix.  Both objects share the very same tensors."""
    object.__setattr__(layer, attr, inner)
    for pname, prm in list(inner._parameters.items()):
        if prm is None:
            continue
        own = layer.create_parameter(list(prm.shape), dtype=prm.dtype)
        with torch.no_grad():
            own.copy_(prm)
        layer.add_parameter(pname, own)
        inner._parameters[pname] = own
    for bname, buf in list(inner._buffers.items()):
        if buf is not None:
            layer.register_buffer(bname, buf, persistable=bname not in inner._non_persistent_buffers_set)
    return inner


class _TrainFollows:
    """Mixin: the adopted torch module follows this layer's train / eval mode."""
    _inner_attr = "_m"

    def train(self, mode=True):
        getattr(self, self._inner_attr).train(mode)
        return super().train(mode)

    def eval(self):
        getattr(self, self._inner_attr).eval()
        return super().eval()

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        inner = getattr(self, self._inner_attr)
        for n in list(inner._parameters):               # a cast / move may have replaced tensors: share them again
            if n in self._parameters:
                inner._parameters[n] = self._parameters[n]
        for n in list(inner._buffers):
            if n in self._buffers:
                inner._buffers[n] = self._buffers[n]
        if hasattr(inner, "_flat_weights_names"):
            inner._flat_weights = [inner._parameters.get(n) for n in inner._flat_weights_names]
        self.__dict__.pop("_flat_for", None)
        return out


# ---------------------------------------------------------------------------------------------- recurrent layers
class _RNNBase(_TrainFollows, Layer):
    _mode = "LSTM"
    _inner_attr = "_rnn"

    def __init__(self, input_size, hidden_size, num_layers=1, direction="forward", time_major=False, dropout=0.0,
                 weight_ih_attr=None, weight_hh_attr=None, bias_ih_attr=None, bias_hh_attr=None, name=None, activation="tanh"):
        super().__init__()
        bidir = direction in ("bidirect", "bidirectional")
        kw = dict(input_size=input_size, hidden_size=hidden_size, num_layers=num_layers, batch_first=not time_major,
                  dropout=dropout, bidirectional=bidir, bias=bias_ih_attr is not False)
        if self._mode == "RNN":
            kw["nonlinearity"] = activation
        rnn = getattr(torch.nn, self._mode)(**kw)
        # The parameters belong to THIS layer under the reference's names (weight_ih_l0, weight_hh_l0_reverse, ...:
        # python/paddle/nn/layer/rnn.py RNNBase), which are also what a .pdparams file uses; the torch module that
        # computes with them is held outside the sublayer registry so that no "_rnn." prefix appears in state dicts.
        object.__setattr__(self, "_rnn", rnn)
        for pname, prm in list(rnn._parameters.items()):
            own = self.create_parameter(list(prm.shape), dtype=prm.dtype)
            with torch.no_grad():
                own.copy_(prm)
            self.add_parameter(pname, own)
            rnn._parameters[pname] = own
        rnn._flat_weights = [rnn._parameters.get(n) for n in rnn._flat_weights_names]
        self.input_size, self.hidden_size, self.num_layers = input_size, hidden_size, num_layers
        self.num_directions = 2 if bidir else 1
        self.time_major = time_major

    def forward(self, inputs, initial_states=None, sequence_length=None):
        if inputs.is_cuda and self.__dict__.get("_flat_for") != inputs.device:
            self._rnn.flatten_parameters()          # one contiguous weight block for cuDNN, once per device
            self.__dict__["_flat_for"] = inputs.device
        if sequence_length is not None:
            lens = torch.as_tensor(sequence_length).cpu()
            packed = torch.nn.utils.rnn.pack_padded_sequence(inputs, lens, batch_first=not self.time_major, enforce_sorted=False)
            out, st = self._rnn(packed, initial_states)
            out, _ = torch.nn.utils.rnn.pad_packed_sequence(out, batch_first=not self.time_major,
                                                            total_length=inputs.shape[0 if self.time_major else 1])
            return out, st
        return self._rnn(inputs, initial_states)



class LSTM(_RNNBase):
    _mode = "LSTM"


class GRU(_RNNBase):
    _mode = "GRU"


class SimpleRNN(_RNNBase):
    _mode = "RNN"


class _CellBase(_TrainFollows, Layer):
    _cls = torch.nn.LSTMCell
    _inner_attr = "_cell"

    def __init__(self, input_size, hidden_size, weight_ih_attr=None, weight_hh_attr=None, bias_ih_attr=None,
                 bias_hh_attr=None, name=None, **kw):
        super().__init__()
        _adopt(self, self._cls(input_size, hidden_size, bias=bias_ih_attr is not False, **kw), "_cell")
        self.input_size, self.hidden_size = input_size, hidden_size

    @property
    def state_shape(self):
        return (self.hidden_size,)


class LSTMCell(_CellBase):
    _cls = torch.nn.LSTMCell

    def forward(self, inputs, states=None):
        h, c = self._cell(inputs, states)
        return h, (h, c)

    @property
    def state_shape(self):
        return ((self.hidden_size,), (self.hidden_size,))


class GRUCell(_CellBase):
    _cls = torch.nn.GRUCell

    def forward(self, inputs, states=None):
        h = self._cell(inputs, states)
        return h, h


class SimpleRNNCell(_CellBase):
    _cls = torch.nn.RNNCell

    def __init__(self, input_size, hidden_size, activation="tanh", **kw):
        super().__init__(input_size, hidden_size, nonlinearity=activation, **{k: v for k, v in kw.items() if k.endswith("_attr")})

    def forward(self, inputs, states=None):
        h = self._cell(inputs, states)
        return h, h


class RNN(Layer):
    """Runs a cell over the time dimension (reference: nn.RNN(cell, is_reverse=False, time_major=False))."""

    def __init__(self, cell, is_reverse=False, time_major=False):
        super().__init__()
        self.cell, self.is_reverse, self.time_major = cell, is_reverse, time_major

    def forward(self, inputs, initial_states=None, sequence_length=None, **kwargs):
        x = inputs if self.time_major else inputs.transpose(0, 1)
        steps = range(x.shape[0] - 1, -1, -1) if self.is_reverse else range(x.shape[0])
        states, outs = initial_states, [None] * x.shape[0]
        for t in steps:
            o, states = self.cell(x[t], states)
            outs[t] = o
        out = torch.stack(outs, 0)
        return (out if self.time_major else out.transpose(0, 1)), states


class BiRNN(Layer):
    def __init__(self, cell_fw, cell_bw, time_major=False):
        super().__init__()
        self.fw, self.bw = RNN(cell_fw, False, time_major), RNN(cell_bw, True, time_major)

    def forward(self, inputs, initial_states=None, sequence_length=None, **kwargs):
        s_fw, s_bw = initial_states if initial_states is not None else (None, None)
        o1, f1 = self.fw(inputs, s_fw)
        o2, f2 = self.bw(inputs, s_bw)
        return torch.cat([o1, o2], -1), (f1, f2)


# ---------------------------------------------------------------------------------------------- transformer decoder
# ---------------------------------------------------------------------------------------------- thin torch-backed layers
def _wrap(name, factory, doc=""):
    """Layer whose forward is a torch module built from Paddle-style constructor arguments by ``factory``."""
    class _W(_TrainFollows, Layer):
        def __init__(self, *args, **kwargs):
            super().__init__()
            kwargs.pop("name", None)
            _adopt(self, factory(*args, **kwargs), "_m")

        def forward(self, *x):
            return self._m(*x)
    _W.__name__ = _W.__qualname__ = name
    _W.__doc__ = doc
    return _W


def _bias(bias_attr):
    return bias_attr is not False


Unflatten = _wrap("Unflatten", lambda axis, shape: torch.nn.Unflatten(axis, tuple(shape)))
AlphaDropout = _wrap("AlphaDropout", lambda p=0.5: torch.nn.AlphaDropout(p))
Dropout3D = _wrap("Dropout3D", lambda p=0.5, data_format="NCDHW": torch.nn.Dropout3d(p))
Bilinear = _wrap("Bilinear", lambda in1_features, in2_features, out_features, weight_attr=None, bias_attr=None:
                 torch.nn.Bilinear(in1_features, in2_features, out_features, _bias(bias_attr)))
PairwiseDistance = _wrap("PairwiseDistance", lambda p=2.0, epsilon=1e-6, keepdim=False: torch.nn.PairwiseDistance(p, epsilon, keepdim))
CosineEmbeddingLoss = _wrap("CosineEmbeddingLoss", lambda margin=0, reduction="mean": torch.nn.CosineEmbeddingLoss(margin, reduction=reduction))
MarginRankingLoss = _wrap("MarginRankingLoss", lambda margin=0.0, reduction="mean": torch.nn.MarginRankingLoss(margin, reduction=reduction))
HingeEmbeddingLoss = _wrap("HingeEmbeddingLoss", lambda margin=1.0, reduction="mean": torch.nn.HingeEmbeddingLoss(margin, reduction=reduction))
TripletMarginLoss = _wrap("TripletMarginLoss", lambda margin=1.0, p=2.0, epsilon=1e-6, swap=False, reduction="mean":
                          torch.nn.TripletMarginLoss(margin, p, epsilon, swap, reduction=reduction))
MultiLabelSoftMarginLoss = _wrap("MultiLabelSoftMarginLoss", lambda weight=None, reduction="mean":
                                 torch.nn.MultiLabelSoftMarginLoss(weight, reduction=reduction))
SoftMarginLoss = _wrap("SoftMarginLoss", lambda reduction="mean": torch.nn.SoftMarginLoss(reduction=reduction))


class CTCLoss(Layer):
    def __init__(self, blank=0, reduction="mean"):
        super().__init__()
        self.blank, self.reduction = blank, reduction

    def forward(self, log_probs, labels, input_lengths, label_lengths, norm_by_times=False):
        return F.ctc_loss(log_probs, labels, input_lengths, label_lengths, self.blank, self.reduction)


class SpectralNorm(Layer):
    """Power-iteration spectral normalisation of a weight tensor (reference: nn.SpectralNorm(weight_shape, dim, power_iters))."""

    def __init__(self, weight_shape, dim=0, power_iters=1, epsilon=1e-12, dtype="float32"):
        super().__init__()
        self.dim, self.power_iters, self.eps = dim, power_iters, epsilon
        h = weight_shape[dim]
        w = 1
        for i, s in enumerate(weight_shape):
            if i != dim:
                w *= s
        self.register_buffer("weight_u", torch.nn.functional.normalize(torch.randn(h), dim=0))
        self.register_buffer("weight_v", torch.nn.functional.normalize(torch.randn(w), dim=0))

    def forward(self, weight):
        mat = weight.transpose(0, self.dim).reshape(weight.shape[self.dim], -1) if self.dim != 0 else weight.reshape(weight.shape[0], -1)
        u, v = self.weight_u, self.weight_v
        with torch.no_grad():
            for _ in range(self.power_iters):
                v = torch.nn.functional.normalize(mat.t() @ u, dim=0, eps=self.eps)
                u = torch.nn.functional.normalize(mat @ v, dim=0, eps=self.eps)
            self.weight_u.copy_(u)
            self.weight_v.copy_(v)
        sigma = u @ (mat @ v)
        return weight / sigma


class HSigmoidLoss(Layer):
    """Hierarchical sigmoid over the default complete binary tree (reference: nn.HSigmoidLoss, non-custom-tree mode)."""

    def __init__(self, feature_size, num_classes, weight_attr=None, bias_attr=None, is_custom=False, is_sparse=False, name=None):
        super().__init__()
        if is_custom:
            raise NotImplementedError("HSigmoidLoss: custom trees are not supported")
        self.num_classes = num_classes
        self.weight = self.create_parameter([num_classes - 1, feature_size], attr=weight_attr)
        self.bias = self.create_parameter([num_classes - 1, 1], attr=bias_attr, is_bias=True) if bias_attr is not False else None

    def forward(self, input, label, path_table=None, path_code=None):
        # node ids / codes of the complete binary tree: leaf = label + num_classes, walk up to the root (1)
        lab = label.reshape(-1).long() + self.num_classes
        depth = int(self.num_classes - 1).bit_length()
        loss = input.new_zeros(input.shape[0], dtype=torch.float32)
        for _ in range(depth):
            parent, code = lab // 2, (lab % 2).float()
            valid = parent >= 1
            idx = (parent - 1).clamp(0, self.num_classes - 2)
            z = (input.float() * self.weight[idx].float()).sum(-1)
            if self.bias is not None:
                z = z + self.bias[idx, 0].float()
            loss = loss + torch.where(valid, TF.binary_cross_entropy_with_logits(z, code, reduction="none"), torch.zeros_like(z))
            lab = torch.where(valid, parent, lab)
            lab = torch.where(lab <= 1, torch.zeros_like(lab), lab)
        return loss.reshape(-1, 1)


# ---------------------------------------------------------------------------------------------- remaining layers
RNNCellBase = _CellBase
PoissonNLLLoss = _wrap("PoissonNLLLoss", lambda log_input=True, full=False, epsilon=1e-8, reduction="mean":
                       torch.nn.PoissonNLLLoss(log_input, full, eps=epsilon, reduction=reduction))
GaussianNLLLoss = _wrap("GaussianNLLLoss", lambda full=False, epsilon=1e-6, reduction="mean":
                        torch.nn.GaussianNLLLoss(full=full, eps=epsilon, reduction=reduction))
MultiMarginLoss = _wrap("MultiMarginLoss", lambda p=1, margin=1.0, weight=None, reduction="mean":
                        torch.nn.MultiMarginLoss(p, margin, weight, reduction=reduction))
TripletMarginWithDistanceLoss = _wrap("TripletMarginWithDistanceLoss", lambda distance_function=None, margin=1.0, swap=False,
                                      reduction="mean": torch.nn.TripletMarginWithDistanceLoss(
                                          distance_function=distance_function, margin=margin, swap=swap, reduction=reduction))
RReLU = _wrap("RReLU", lambda lower=1.0 / 8.0, upper=1.0 / 3.0: torch.nn.RReLU(lower, upper))
FractionalMaxPool2D = _wrap("FractionalMaxPool2D", lambda output_size, kernel_size=None, random_u=None, return_mask=False:
                            torch.nn.FractionalMaxPool2d(kernel_size or 2, output_size=output_size, return_indices=return_mask))
FractionalMaxPool3D = _wrap("FractionalMaxPool3D", lambda output_size, kernel_size=None, random_u=None, return_mask=False:
                            torch.nn.FractionalMaxPool3d(kernel_size or 2, output_size=output_size, return_indices=return_mask))
AdaptiveLogSoftmaxWithLoss = _wrap("AdaptiveLogSoftmaxWithLoss", lambda in_features, n_classes, cutoffs, div_value=4.0,
                                   head_bias=False: torch.nn.AdaptiveLogSoftmaxWithLoss(in_features, n_classes, cutoffs,
                                                                                        div_value, head_bias))


class RNNTLoss(Layer):
    def __init__(self, blank=0, fastemit_lambda=0.001, reduction="mean", name=None):
        super().__init__()

    def forward(self, *a, **k):
        return F.rnnt_loss(*a, **k)


class BeamSearchDecoder:
    """Beam-search decoding over an RNN cell (reference: nn.BeamSearchDecoder + nn.dynamic_decode), greedy-free minimal
    form: keeps ``beam_size`` hypotheses per batch entry, log-prob scored, finished beams frozen on ``end_token``."""

    def __init__(self, cell, start_token, end_token, beam_size, embedding_fn=None, output_fn=None):
        self.cell, self.start, self.end, self.k = cell, start_token, end_token, beam_size
        self.embedding_fn, self.output_fn = embedding_fn, output_fn


def dynamic_decode(decoder, inits=None, max_step_num=None, output_time_major=False, impute_finished=False,
                   is_test=False, return_length=False, **kwargs):
    cell, k, end = decoder.cell, decoder.k, decoder.end
    state = inits
    h0 = state[0] if isinstance(state, (tuple, list)) else state
    b = h0.shape[0]
    dev = h0.device

    def tile(s):
        return s.unsqueeze(1).expand(b, k, *s.shape[1:]).reshape(b * k, *s.shape[1:])
    state = tuple(tile(s) for s in state) if isinstance(state, (tuple, list)) else tile(state)
    tokens = torch.full((b * k,), decoder.start, dtype=torch.long, device=dev)
    scores = torch.zeros(b, k, device=dev)
    scores[:, 1:] = -1e9
    finished = torch.zeros(b, k, dtype=torch.bool, device=dev)
    ids, parents = [], []
    for _ in range(max_step_num or 32):
        inp = decoder.embedding_fn(tokens) if decoder.embedding_fn else tokens
        out, state = cell(inp, state)
        logits = decoder.output_fn(out) if decoder.output_fn else out
        logp = torch.log_softmax(logits.float(), -1).reshape(b, k, -1)
        V = logp.shape[-1]
        frozen = torch.full_like(logp, -1e9)
        frozen[..., end] = 0
        logp = torch.where(finished[..., None], frozen, logp)
        total = (scores[..., None] + logp).reshape(b, k * V)
        scores, flat = total.topk(k, -1)
        parent, tok = flat // V, flat % V
        ids.append(tok)
        parents.append(parent)
        gidx = (parent + torch.arange(b, device=dev)[:, None] * k).reshape(-1)
        state = tuple(s[gidx] for s in state) if isinstance(state, tuple) else state[gidx]
        finished = finished.gather(1, parent) | (tok == end)
        tokens = tok.reshape(-1)
        if bool(finished.all()):
            break
    ids_t, par_t = torch.stack(ids), torch.stack(parents)
    seq = F.gather_tree(ids_t, par_t)
    seq = seq if output_time_major else seq.permute(1, 0, 2)
    lengths = (seq != end).sum(0 if output_time_major else 1) if return_length else None
    return (seq, state, lengths) if return_length else (seq, state)
