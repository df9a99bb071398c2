"""Multi-tensor optimizer kernels (K12 fused AdamW, K15 squared-L2-norm, AMP unscale/check).

One launch per step for any number of tensors: a descriptor table (``MTDesc``, 64 B per tensor, layout mirrored
from csrc/kernels/optim.cu) plus a chunk map are staged in pinned memory and copied with one async H2D each.
Reference: paddle/phi/kernels/gpu/adamw_kernel.cu, fused_adam_kernel.cu, squared_l2_norm_kernel.cu, amp_kernel.cu."""
from __future__ import annotations

import numpy as np
import torch

from ._util import K, check, dtype_code, ptr, stream_ptr

MTDESC = np.dtype([("p", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("master", "<u8"), ("numel", "<i8"),
                   ("lr_ratio", "<f4"), ("decay", "<f4"), ("p_dtype", "<i4"), ("g_dtype", "<i4")])
assert MTDESC.itemsize == 64


class TensorTable:
    """Device-resident descriptor table + chunk map for a fixed list of tensors."""

    def __init__(self, device):
        self.device = device
        self.n_chunks = 0
        self._host = None
        self._dev_table = None
        self._dev_chunks = None
        self._key = None
        self.chunk = K().pdb200_mt_chunk_elems()

    def update(self, rows):
        """rows: list of dict(p,g,m,v,master,numel,lr_ratio,decay,p_dtype,g_dtype) with raw pointers."""
        n = len(rows)
        arr = np.zeros(n, dtype=MTDESC)
        for i, r in enumerate(rows):
            arr[i] = (r.get("p") or 0, r.get("g") or 0, r.get("m") or 0, r.get("v") or 0, r.get("master") or 0,
                      r["numel"], r.get("lr_ratio", 1.0), r.get("decay", 1.0), r.get("p_dtype", 0), r.get("g_dtype", 0))
        key = arr.tobytes()
        if key == self._key:
            return
        self._key = key
        numels = [int(r["numel"]) for r in rows]
        cm = []
        for ti, ne in enumerate(numels):
            nchunk = (ne + self.chunk - 1) // self.chunk
            if nchunk:
                c = np.empty((nchunk, 2), dtype=np.int32)
                c[:, 0] = ti
                c[:, 1] = np.arange(nchunk, dtype=np.int32)
                cm.append(c)
        cm = np.concatenate(cm, 0) if cm else np.zeros((0, 2), np.int32)
        self.n_chunks = int(cm.shape[0])
        tb = torch.from_numpy(arr.view(np.uint8).copy())
        cb = torch.from_numpy(np.ascontiguousarray(cm).view(np.uint8).reshape(-1).copy())
        if self.device.type == "cuda":
            tb, cb = tb.pin_memory(), cb.pin_memory()
        self._host = (tb, cb)  # keep pinned staging alive until the next update
        self._dev_table = tb.to(self.device, non_blocking=True)
        self._dev_chunks = cb.to(self.device, non_blocking=True) if cb.numel() else torch.zeros(8, dtype=torch.uint8, device=self.device)

    @property
    def table_ptr(self):
        return self._dev_table.data_ptr()

    @property
    def chunks_ptr(self):
        return self._dev_chunks.data_ptr()


def adamw_step(table: TensorTable, lr, beta1, beta2, eps, weight_decay, beta1_pow=0.0, beta2_pow=0.0, beta_pows=None,
               grad_scale=None, found_inf=None):
    """beta_pows: optional device fp32[2] = (beta1^t, beta2^t) for this step; overrides the host scalars."""
    check(K().pdb200_adamw_multi(table.table_ptr, table.chunks_ptr, table.n_chunks, float(lr), float(beta1),
                                 float(beta2), float(eps), float(weight_decay), float(beta1_pow), float(beta2_pow),
                                 ptr(beta_pows), ptr(grad_scale), ptr(found_inf), stream_ptr(table.device)), "adamw_multi")


def momentum_step(table: TensorTable, lr, mu, l2_decay, nesterov, grad_scale=None, found_inf=None):
    check(K().pdb200_momentum_multi(table.table_ptr, table.chunks_ptr, table.n_chunks, float(lr), float(mu),
                                    float(l2_decay), int(nesterov), ptr(grad_scale), ptr(found_inf),
                                    stream_ptr(table.device)), "momentum_multi")


def l2norm_sq(table: TensorTable, out=None, accumulate=False):
    """Sum of squares over every table entry's ``g`` tensor -> fp32[1] on device (no host sync)."""
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=table.device)
    scratch = torch.empty(max(table.n_chunks, 1), dtype=torch.float32, device=table.device)
    check(K().pdb200_l2norm_multi(table.table_ptr, table.chunks_ptr, table.n_chunks, ptr(scratch), scratch.numel(),
                                  ptr(out), int(accumulate), stream_ptr(table.device)), "l2norm_multi")
    return out


def unscale_and_check(table: TensorTable, inv_scale, found_inf):
    check(K().pdb200_unscale_check_multi(table.table_ptr, table.chunks_ptr, table.n_chunks, ptr(inv_scale),
                                         ptr(found_inf), stream_ptr(table.device)), "unscale_check_multi")


def grads_table(tensors, device=None):
    """Table whose ``g`` slots point at ``tensors`` (for l2norm / unscale)."""
    tensors = [t for t in tensors if t is not None and t.numel() > 0]
    device = device or (tensors[0].device if tensors else torch.device("cuda"))
    tab = TensorTable(device)
    tab.update([dict(g=t.data_ptr(), numel=t.numel(), g_dtype=dtype_code(t)) for t in tensors])
    return tab


def squared_l2_norm(tensors):
    """sum_i ||t_i||^2 as fp32[1]; native multi-tensor kernel on GPU, torch on CPU."""
    tensors = [t for t in tensors if t is not None and t.numel() > 0]
    if not tensors:
        return torch.zeros(1)
    if tensors[0].is_cuda:
        from ._util import use_native
        if use_native(tensors[0]):
            return l2norm_sq(grads_table(tensors))
    return torch.stack([t.float().pow(2).sum() for t in tensors]).sum().reshape(1)
