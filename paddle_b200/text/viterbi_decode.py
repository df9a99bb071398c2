"""paddle.text.viterbi_decode / ViterbiDecoder (reference: python/paddle/text/viterbi_decode.py; kernel
paddle/phi/kernels/cpu/viterbi_decode_kernel.cc:150-310).

``transition_params[i, j]`` scores the step tag i -> tag j.  With ``include_bos_eos_tag`` the LAST row is the start
tag (added to the first emission) and the SECOND TO LAST column the stop tag (added where each sequence ends).  Returns
(best score per sequence, best path ``[batch, max(lengths)]`` with zeros after each sequence's end).  The recursion is
vectorised over the batch; sequences that have ended carry their scores forward unchanged.
"""
from __future__ import annotations

import torch

from ..nn.layer.layers import Layer

__all__ = ["viterbi_decode", "ViterbiDecoder"]


def viterbi_decode(potentials, transition_params, lengths, include_bos_eos_tag=True, name=None):
    if potentials.dim() != 3:
        raise ValueError(f"potentials should be [batch, seq_len, num_tags], got shape {tuple(potentials.shape)}")
    b, t_max, n = potentials.shape
    if tuple(transition_params.shape) != (n, n):
        raise ValueError(f"transition_params should be [{n}, {n}], got {tuple(transition_params.shape)}")
    lengths = lengths.to(torch.int64).reshape(b)
    steps = int(lengths.max()) if b else 0
    if steps == 0:
        return potentials.new_zeros(b), torch.zeros((b, 0), dtype=torch.int64, device=potentials.device)
    trans = transition_params
    alpha = potentials[:, 0]
    if include_bos_eos_tag:
        start, stop = trans[-1], trans[:, -2]
        alpha = alpha + start
        alpha = alpha + stop * (lengths == 1).to(alpha.dtype).unsqueeze(1)
    history = []
    for i in range(1, steps):
        scores = alpha.unsqueeze(2) + trans.unsqueeze(0)                  # [b, from, to]
        best, arg = scores.max(1)
        history.append(arg)
        nxt = best + potentials[:, i]
        live = (lengths > i).to(alpha.dtype).unsqueeze(1)
        alpha = nxt * live + alpha * (1 - live)
        if include_bos_eos_tag:
            alpha = alpha + stop * (lengths == i + 1).to(alpha.dtype).unsqueeze(1)
    score, last = alpha.max(1)
    path = torch.zeros((b, steps), dtype=torch.int64, device=potentials.device)
    # walk back: position p of sequence s is decided once p < length[s]; its last position takes the arg-max tag
    cur = last
    for p in range(steps - 1, -1, -1):
        is_last = lengths == p + 1
        inside = lengths > p
        if p < steps - 1:
            prev = history[p].gather(1, cur.unsqueeze(1)).squeeze(1)      # best tag at p given the tag at p + 1
            cur = torch.where(lengths > p + 1, prev, cur)
        cur = torch.where(is_last, last, cur)
        path[:, p] = torch.where(inside, cur, torch.zeros_like(cur))
    return score, path


class ViterbiDecoder(Layer):
    def __init__(self, transitions, include_bos_eos_tag=True, name=None):
        super().__init__()
        self.transitions = transitions
        self.include_bos_eos_tag = include_bos_eos_tag
        self.name = name

    def forward(self, potentials, lengths):
        return viterbi_decode(potentials, self.transitions, lengths, self.include_bos_eos_tag, self.name)
