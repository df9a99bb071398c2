"""Pipeline p2p (reference: python/paddle/distributed/fleet/meta_parallel/pp_utils/p2p_communication.py:52-882).

Stays on NCCL p2p by design (north star): a first-use shape/dtype handshake (``SendRecvMeta``), then batched
isend/irecv (ncclGroupStart/End) between neighbouring stages, issued on the communication stream so the
transfers overlap compute.  ``enable_partial_send_recv`` (send 1/mp of the activation, all-gather inside the mp
group) is honoured when the tensor divides evenly.
"""
from __future__ import annotations

import torch

from paddle_b200.distributed import collective as C
from paddle_b200.distributed import communication as comm
from paddle_b200.framework import dtype as _dt

_hcg = None
_enable_partial_send_recv = True
_use_batch_p2p = True

_DTYPE_IDS = [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64, torch.int32, torch.int8,
              torch.uint8, torch.bool]


def initialize_p2p_groups(hcg, enable_partial_send_recv=True, use_batch_p2p_comm=True):
    global _hcg, _enable_partial_send_recv, _use_batch_p2p
    _hcg = hcg
    _enable_partial_send_recv = enable_partial_send_recv
    _use_batch_p2p = use_batch_p2p_comm


class SendRecvMeta:
    """Shape / dtype / stop_gradient of what travels between two stages (exchanged once)."""

    def __init__(self):
        self.init_or_erase_meta()

    def init_or_erase_meta(self):
        self.send_shape_message = None
        self.send_dtype_message = None
        self.recv_shape_message = None
        self.recv_dtype_message = None
        self.recv_stop_gradient = None
        self.has_send_meta = False
        self.has_recv_meta = False

    def _dev(self):
        return torch.device("cuda", torch.cuda.current_device()) if C._state.backend == "nccl" else torch.device("cpu")

    def _send_one(self, t, peer, group):
        meta = [t.dim(), _DTYPE_IDS.index(t.dtype), int(not t.requires_grad)] + list(t.shape)
        buf = torch.tensor([len(meta)] + meta + [0] * (15 - len(meta)), dtype=torch.int64, device=self._dev())
        comm.send(buf, peer, group)

    def _recv_one(self, peer, group):
        buf = torch.zeros(16, dtype=torch.int64, device=self._dev())
        comm.recv(buf, peer, group)
        v = buf.tolist()
        n = v[0]
        meta = v[1:1 + n]
        return tuple(meta[3:3 + meta[0]]), _DTYPE_IDS[meta[1]], bool(meta[2])

    def send_meta(self, tensor, group, peer):
        tensors = tensor if isinstance(tensor, tuple) else (tensor,)
        cnt = torch.tensor([len(tensors), int(isinstance(tensor, tuple))], dtype=torch.int64, device=self._dev())
        comm.send(cnt, peer, group)
        for t in tensors:
            self._send_one(t, peer, group)
        self.has_send_meta = True

    def recv_meta(self, group, peer):
        cnt = torch.zeros(2, dtype=torch.int64, device=self._dev())
        comm.recv(cnt, peer, group)
        n, is_tuple = cnt.tolist()
        metas = [self._recv_one(peer, group) for _ in range(n)]
        if is_tuple:
            self.recv_shape_message = tuple(m[0] for m in metas)
            self.recv_dtype_message = tuple(m[1] for m in metas)
            self.recv_stop_gradient = tuple(m[2] for m in metas)
        else:
            self.recv_shape_message, self.recv_dtype_message, self.recv_stop_gradient = metas[0]
        self.has_recv_meta = True

    def set_send_message(self, tensor):
        if isinstance(tensor, tuple):
            self.send_shape_message = tuple(tuple(t.shape) for t in tensor)
            self.send_dtype_message = tuple(t.dtype for t in tensor)
        else:
            self.send_shape_message = tuple(tensor.shape)
            self.send_dtype_message = tensor.dtype


def _partial_ok(t):
    mp = _hcg.get_model_parallel_world_size()
    return _enable_partial_send_recv and mp > 1 and t.numel() % mp == 0


class P2pHelper:
    def __init__(self, use_cache=True):
        self._send_recv_meta = SendRecvMeta()
        self._use_cache = use_cache

    def clear_meta_cache(self):
        self._send_recv_meta.init_or_erase_meta()

    @property
    def _pp_group(self):
        return _hcg.get_pipe_parallel_group()

    def _send_meta(self, output_tensor):
        if not self._send_recv_meta.has_send_meta:
            self._send_recv_meta.set_send_message(output_tensor)
            self._send_recv_meta.send_meta(output_tensor, self._pp_group, _hcg._get_p2p_next_rank())
            self._send_recv_meta.has_send_meta = self._use_cache

    def _recv_meta(self):
        if not self._send_recv_meta.has_recv_meta:
            self._send_recv_meta.recv_meta(self._pp_group, _hcg._get_p2p_prev_rank())
            self._send_recv_meta.has_recv_meta = self._use_cache

    # ---- core exchange -------------------------------------------------------------------------------
    def _alloc_recv(self, shapes, dtypes, stop_grads):
        dev = self._send_recv_meta._dev()
        if shapes is None:
            return None
        if isinstance(shapes, tuple) and shapes and isinstance(shapes[0], tuple):
            outs = []
            for s, d, sg in zip(shapes, dtypes, stop_grads if stop_grads is not None else [False] * len(shapes)):
                t = torch.empty(s, dtype=d, device=dev)
                t.requires_grad_(not sg and t.is_floating_point())
                outs.append(t)
            return tuple(outs)
        t = torch.empty(shapes, dtype=dtypes, device=dev)
        t.requires_grad_(not (stop_grads or False) and t.is_floating_point())
        return t

    def _p2p(self, tensor_send_next, tensor_send_prev, recv_prev, recv_next, sync_recv=True):
        m = self._send_recv_meta
        ops = []
        group = self._pp_group
        prev_rank = _hcg._get_p2p_prev_rank() if hasattr(_hcg, "prev_rank") else None
        next_rank = _hcg._get_p2p_next_rank() if hasattr(_hcg, "next_rank") else None
        tensor_recv_prev = self._alloc_recv(m.recv_shape_message, m.recv_dtype_message, m.recv_stop_gradient) if recv_prev else None
        # the gradient that comes back has the shape of what this stage sent forward
        tensor_recv_next = self._alloc_recv(m.send_shape_message, m.send_dtype_message, None) if recv_next else None
        if tensor_recv_next is not None:
            for t in (tensor_recv_next if isinstance(tensor_recv_next, tuple) else (tensor_recv_next,)):
                t.requires_grad_(False)

        def each(x):
            return x if isinstance(x, tuple) else (x,)

        with torch.no_grad():
            # Order matters when prev and next are the SAME peer (pp == 2 with virtual stages): messages between one
            # pair match in issue order, and the peer's "send to prev" is this rank's "recv from next".  Issuing
            # [send_prev, recv_next, send_next, recv_prev] on every rank makes the pairs line up for any pp.
            if tensor_send_prev is not None:
                for t in each(tensor_send_prev):
                    ops.append(comm.P2POp(comm.isend, t.contiguous(), prev_rank, group))
            if tensor_recv_next is not None:
                for t in each(tensor_recv_next):
                    ops.append(comm.P2POp(comm.irecv, t, next_rank, group))
            if tensor_send_next is not None:
                for t in each(tensor_send_next):
                    ops.append(comm.P2POp(comm.isend, t.detach().contiguous(), next_rank, group))
            if tensor_recv_prev is not None:
                for t in each(tensor_recv_prev):
                    ops.append(comm.P2POp(comm.irecv, t, prev_rank, group))
            tasks = []
            if ops:
                if _use_batch_p2p and C._state.backend == "nccl":
                    tasks = comm.batch_isend_irecv(ops)
                else:
                    # isend/irecv are non-blocking on gloo and NCCL: issue in list order, wait afterwards
                    tasks = [o.op(o.tensor, o.peer, o.group) for o in ops]
            if sync_recv:
                for t in tasks:
                    t.wait()
                tasks = []
        return tensor_recv_prev, tensor_recv_next, tasks

    # ---- public verbs (reference :637-882) --------------------------------------------------------------
    def recv_forward(self, pp_first_stage, sync_recv=True, batch_p2p_comm=True):
        if pp_first_stage:
            return None
        self._recv_meta()
        out, _, _ = self._p2p(None, None, True, False, sync_recv)
        return out

    def recv_backward(self, pp_last_stage, sync_recv=True, batch_p2p_comm=True):
        if pp_last_stage:
            return None
        _, out, _ = self._p2p(None, None, False, True, sync_recv)
        return out

    def send_forward(self, output_tensor, pp_last_stage, batch_p2p_comm=True):
        if not pp_last_stage:
            self._send_meta(output_tensor)
            self._p2p(output_tensor, None, False, False)

    def send_backward(self, input_tensor_grad, pp_first_stage, batch_p2p_comm=True):
        if not pp_first_stage:
            self._p2p(None, input_tensor_grad, False, False)

    def send_forward_recv_backward(self, output_tensor, pp_last_stage, batch_p2p_comm=True):
        if pp_last_stage:
            return None
        self._send_meta(output_tensor)
        _, out, _ = self._p2p(output_tensor, None, False, True)
        return out

    def send_backward_recv_forward(self, input_tensor_grad, pp_first_stage, batch_p2p_comm=True):
        if pp_first_stage:
            return None
        out, _, _ = self._p2p(None, input_tensor_grad, True, False)
        return out

    def send_forward_backward_recv_forward_backward(self, output_tensor, input_tensor_grad, recv_prev, recv_next,
                                                    batch_p2p_comm=True):
        if output_tensor is not None:
            self._send_meta(output_tensor)
        if recv_prev:
            self._recv_meta()
        a, b, _ = self._p2p(output_tensor, input_tensor_grad, recv_prev, recv_next)
        return a, b

    def send_forward_recv_forward(self, output_tensor, recv_prev, batch_p2p_comm=True, overlap_p2p_comm=False):
        if output_tensor is not None:
            self._send_meta(output_tensor)
        if recv_prev:
            self._recv_meta()
        a, _, tasks = self._p2p(output_tensor, None, recv_prev, False, sync_recv=not overlap_p2p_comm)
        return (a, tasks) if overlap_p2p_comm else a

    def send_backward_recv_backward(self, input_tensor_grad, recv_next, batch_p2p_comm=True, overlap_p2p_comm=False):
        _, b, tasks = self._p2p(None, input_tensor_grad, False, recv_next, sync_recv=not overlap_p2p_comm)
        return (b, tasks) if overlap_p2p_comm else b
