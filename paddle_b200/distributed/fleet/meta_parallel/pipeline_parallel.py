"""Pipeline-parallel schedules (reference: python/paddle/distributed/fleet/meta_parallel/pipeline_parallel.py:
FakeMicroDataset :63-146, PipelineParallel 1F1B :149-1007, PipelineParallelWithInterleave :1010-1828,
PipelineParallelWithInterleaveFthenB :1831-2060).

The scheduler is host Python (as in the reference); p2p stays on NCCL (north star).  Loss is divided by
``accumulate_steps`` inside the forward step unless ``delay_scale_loss``; the final loss is broadcast from the last
stage; tied-embedding gradients are all-reduced between their stages before the optimizer step.
"""
from __future__ import annotations

import time

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.distributed.fleet.utils.hybrid_parallel_util import (broadcast_dp_parameters, broadcast_mp_parameters,
                                                                      broadcast_sep_parameters,
                                                                      broadcast_sharding_parameters)
from paddle_b200.distributed.fleet.utils import timer_helper as timer

from .meta_parallel_base import MetaParallelBase
from .parallel_layers.pp_layers import PipelineLayer
from .pp_utils import p2p_communication as p2p


class FakeMicroDataset:
    """Iterator over the micro-batches of one global batch ``(inputs, labels)`` (the reference's name for it:
    pipeline_parallel.py FakeMicroDataset).  The first stage reads inputs, the last stage labels, the others nothing.
    A field is either one tensor of ``acc_steps * micro_batch_size`` rows (cut into row blocks), a list of ``acc_steps``
    ready-made micro-batches, ``None``, or a tuple of such fields."""

    def __init__(self, data, is_first_stage, is_last_stage, acc_steps, micro_batch_size):
        self._fields = data
        self._want = (bool(is_first_stage), bool(is_last_stage))
        self._n, self._rows = acc_steps, micro_batch_size
        self._k = 0

    def __iter__(self):
        return self

    def _cut(self, field, k):
        if field is None:
            return None
        if isinstance(field, tuple):
            return tuple(self._cut(f, k) for f in field)
        if isinstance(field, list):
            assert len(field) == self._n, f"a list field holds one entry per micro-batch: {len(field)} != {self._n}"
            return field[k].detach()
        assert field.shape[0] == self._n * self._rows, (
            f"batch_size needs to be divisible by micro_batch_size. Currently, batch_size = {field.shape[0]}, "
            f"micro_batch_size = {self._rows}, accumulate_steps = {self._n}.")
        return field[k * self._rows:(k + 1) * self._rows].detach()

    def __next__(self):
        if self._k >= self._n:
            raise StopIteration
        assert any(self._want), "only the first and the last pipeline stage read data"
        k, self._k = self._k, self._k + 1
        return tuple(self._cut(self._fields[i], k) if self._want[i] else None for i in (0, 1))


def one_f_one_b_slots(num_stages, acc_steps, stage):
    """1F1B for one chunk per stage, as a table of ``(forward micro-batch or None, backward micro-batch or None)``:
    stage s runs ``P - s - 1`` forwards ahead (so that the last stage can start its first backward right after its first
    forward), then pairs one forward with one backward per slot, then drains the backwards it is ahead by."""
    lead = min(num_stages - stage - 1, acc_steps)
    slots = [(k, None) for k in range(lead)]
    slots += [(lead + i, i) for i in range(acc_steps - lead)]
    slots += [(None, i) for i in range(acc_steps - lead, acc_steps)]
    return slots


class PipelineParallel(MetaParallelBase):
    def __init__(self, layers, hcg, strategy):
        if not isinstance(layers, PipelineLayer):
            raise TypeError("The Layer should be a derived class of PipelineLayer.")
        super().__init__(layers, hcg, strategy)
        self.use_data_parallel = hcg.get_data_parallel_world_size() > 1
        self.use_model_parallel = hcg.get_model_parallel_world_size() > 1
        self.use_sep_parallel = hcg.get_sep_parallel_world_size() > 1
        self.use_sharding_parallel = hcg.get_sharding_parallel_world_size() > 1
        self.total_loss = None
        pc = strategy.pipeline_configs
        self.micro_batch_size = pc["micro_batch_size"]
        self.accumulate_steps = pc["accumulate_steps"]
        self._using_cache = pc["p2p_cache_shape"]
        self._enable_partial_send_recv = pc["enable_partial_send_recv"]
        ppc = strategy.hybrid_configs["pp_configs"]
        self._delay_scale_loss = ppc["delay_scale_loss"]
        self._dp_comm_overlap = ppc["dp_comm_overlap"]
        self._sharding_comm_overlap = ppc["sharding_comm_overlap"]
        self._enable_timer = ppc["enable_timer"]
        self._release_gradients = ppc["release_gradients"]
        self._overlap_p2p_comm = ppc["overlap_p2p_comm"]
        self._clear_every_step_cache = ppc["clear_every_step_cache"]
        self._use_batch_p2p_comm = ppc["use_batch_p2p_comm"]
        self._profiling = ppc["profiling"]
        self._records = []
        self._record_format = '"name": "{}{}", "cat": "pipeline timeline", "ph": {}, "pid": 0, "tid": ' + \
            str(hcg.get_stage_id() + 1) + ', "ts": {}, "cname": "{}"'
        if self._enable_timer and not timer.is_timer_initialized():
            timer.set_timers()
        self.num_stages = hcg.get_pipe_parallel_world_size()
        self.stage_id = hcg.get_stage_id()
        self.global_rank = hcg.get_global_rank()
        self.pp_group = hcg.get_pipe_parallel_group()
        self.dp_group = hcg.get_data_parallel_group()
        self._virtual_pp_world_size = None
        self._virtual_pp_rank = None
        self._real_pp_world_size = self.num_stages
        self._real_pp_rank = self.stage_id
        self._compute_loss = True
        p2p.initialize_p2p_groups(hcg, self._enable_partial_send_recv, self._use_batch_p2p_comm)
        self._p2p_helper = p2p.P2pHelper(self._using_cache)
        self.is_pipeline_first_stage_ = self.stage_id == 0
        self.is_pipeline_last_stage_ = self.stage_id == self.num_stages - 1
        self._comm_buffers = []
        self.loss_fn_idx = 0
        self.micro_batch_id = 0
        if self._dp_comm_overlap and self.dp_group is not None and self.dp_group.nranks > 1:
            self.register_allreduce_overlap_hook(self._layers, self.dp_group, self.accumulate_steps, True)

    def _prepare_for_model(self):
        hcg = self._hcg
        if hcg.get_model_parallel_world_size() > 1:
            broadcast_mp_parameters(self._layers, hcg)
        if hcg.get_sep_parallel_world_size() > 1:
            broadcast_sep_parameters(self._layers, hcg)
        if hcg.get_sharding_parallel_world_size() > 1:
            broadcast_sharding_parameters(self._layers, hcg)
        if hcg.get_data_parallel_world_size() > 1:
            broadcast_dp_parameters(self._layers, hcg)

    # ------------------------------------------------- gradient communication overlapped with the backward pass
    def fused_gradient(self, model, comm_group, acc_steps, dp, group_size=128 * 1024 * 1024, dst_of=None):
        """FusedCommBuffers (<= ``group_size`` bytes each) over the trainable parameters of every model chunk.  ``dp``:
        all-reduce over the data-parallel group; otherwise reduce to the sharding rank that owns the parameters
        (``dst_of``: parameter -> owner rank, from the sharding optimizer)."""
        from ..utils.tensor_fusion_helper import HOOK_ACTION, FusedCommBuffer, assign_group_by_size
        chunks = model.get_model_chunks() if getattr(model, "get_num_virtual_stages", lambda: 1)() > 1 else [model]
        buffers = []
        for chunk in chunks:
            params = [p for p in chunk.parameters() if p.requires_grad]
            if not params:
                continue
            if dp:
                partitions = {-1: params}
            else:
                partitions = {}
                for prm in params:
                    partitions.setdefault(dst_of(prm), []).append(prm)
            for dst, part in sorted(partitions.items()):
                for ps in assign_group_by_size(part, group_size).values():
                    buffers.append(FusedCommBuffer(len(self._comm_buffers) + len(buffers), ps, comm_group, acc_steps,
                                                   HOOK_ACTION.ALL_REDUCE if dp else HOOK_ACTION.REDUCE,
                                                   dst=-1 if dp else comm_group.ranks[dst]))
        return buffers

    def bw_hook_func(self, buffer, param):
        """Backward hook of one parameter: check its gradient into its buffer; the buffer launches its collective
        (asynchronously, on the communication stream) when every parameter has checked in ``acc_steps`` times - i.e.
        during the last micro-batch's backward, overlapping the layers still to be differentiated."""
        def hook(*_):
            buffer.add_grad(param)
        return hook

    def register_allreduce_overlap_hook(self, model, comm_group, acc_steps, dp, group_size=128 * 1024 * 1024, dst_of=None):
        """reference: pipeline_parallel.py register_allreduce_overlap_hook (``pp_configs.dp_comm_overlap`` /
        ``sharding_comm_overlap``)."""
        buffers = self.fused_gradient(model, comm_group, acc_steps, dp, group_size, dst_of)
        for buffer in buffers:
            for param in buffer.params:
                param.register_post_accumulate_grad_hook(self.bw_hook_func(buffer, param))
        self._comm_buffers += buffers
        return buffers

    def register_sharding_comm_overlap_hook(self, optimizer):
        """Called with the (sharding) optimizer at the first ``train_batch``: gradients are reduced to their owner rank
        while the backward pass is still running."""
        if not self._sharding_comm_overlap or self._comm_buffers:
            return
        inner = optimizer
        while hasattr(inner, "_inner_opt") and not hasattr(inner, "_param2rank"):
            inner = inner._inner_opt
        table = getattr(inner, "_param2rank", None)
        if table is None:
            raise RuntimeError("sharding_comm_overlap needs the stage-1 sharding optimizer (its parameter -> rank table)")
        group = self._hcg.get_sharding_parallel_group()
        self.register_allreduce_overlap_hook(self._layers, group, self.accumulate_steps, False, dst_of=lambda q: table[q.name])

    def _finish_overlapped_comm(self):
        for buffer in self._comm_buffers:
            buffer.scale_grads()

    def timer_printer(self):
        if self._enable_timer and timer.is_timer_initialized():
            timer.get_timers().log(timer.get_timers().timers.keys())

    # ------------------------------------------------------------------ helpers
    def is_pipeline_first_stage(self, ignore_virtual=False):
        if not ignore_virtual and self._virtual_pp_world_size is not None:
            if self._virtual_pp_rank != 0:
                return False
        return self.is_pipeline_first_stage_

    def is_pipeline_last_stage(self, ignore_virtual=False):
        if not ignore_virtual and self._virtual_pp_world_size is not None:
            if self._virtual_pp_rank != self._virtual_pp_world_size - 1:
                return False
        return self.is_pipeline_last_stage_

    def set_virtual_pipeline_rank(self, rank):
        self._virtual_pp_rank = rank

    def _record_stamp(self, name, step, phase, color):
        if self._profiling:
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            self._records.append("{" + self._record_format.format(name, step, phase, int(time.time() * 1000), color) + "}")

    def _flush_records(self):
        if self._profiling:
            with open(f"./profile_record_tmp_file_for_rank_{self.global_rank}", "a+") as f:
                for r in self._records:
                    f.write(r + "\n")
            self._records = []

    def _wrap_data(self, data):
        return FakeMicroDataset(data, self.is_pipeline_first_stage(ignore_virtual=True),
                                self.is_pipeline_last_stage(ignore_virtual=True), self.accumulate_steps,
                                self.micro_batch_size)

    def _prepare_training(self, data, optimizer, lr_scheduler):
        assert isinstance(optimizer, object)
        self.optimizer = optimizer
        self.lr_scheduler = lr_scheduler
        self._layers.train()
        self.register_sharding_comm_overlap_hook(optimizer)
        return data

    def get_static_scheduler(self):
        return self.forward_backward_pipeline(data=None, scaler=None, static_scheduler=True)

    # ------------------------------------------------------------------ 1F1B
    def _run_slots_1f1b(self, slots, micro_dataset, keep_outputs=False):
        """Execute a plain (one chunk per stage) slot table.  Communication follows from the table alone:
          * a forward needs its input: received right before it, unless the previous backward already fetched it;
          * a forward paired with a backward in its slot hands its output over and takes the gradient for that backward
            in one call; a lone forward just sends;
          * a backward followed by a slot with a forward sends its input gradient and fetches that forward's input in one
            call; otherwise it just sends, and a lone backward first receives its gradient.
        Inputs / outputs wait in FIFO order for their backward."""
        helper = self._p2p_helper
        first, last = self.is_pipeline_first_stage(), self.is_pipeline_last_stage()
        waiting_in, waiting_out, outputs = [], [], []
        have_input, next_input = False, None
        for idx, (f, b) in enumerate(slots):
            nxt = slots[idx + 1] if idx + 1 < len(slots) else None
            grad_in = None
            if f is not None:
                input_tensor = next_input if have_input else helper.recv_forward(first)
                have_input, next_input = False, None
                self._record_stamp("F", f, '"B"', "rail_response")
                output_tensor = self._forward_step(input_tensor, micro_dataset)
                self._record_stamp("F", f, '"E"', "rail_response")
                if b is not None:
                    grad_in = helper.send_forward_recv_backward(output_tensor, last)
                else:
                    helper.send_forward(output_tensor, last)
                if keep_outputs:
                    outputs.append(output_tensor)
                    continue
                waiting_in.append(input_tensor)
                waiting_out.append(output_tensor)
                if not last:
                    self._release_output(output_tensor)
            if b is not None:
                input_tensor, output_tensor = waiting_in.pop(0), waiting_out.pop(0)
                if f is None:
                    grad_in = helper.recv_backward(last)
                self._record_stamp("B", b, '"B"', "rail_animation")
                input_grad = self._backward_step(input_tensor, output_tensor, grad_in)
                self._record_stamp("B", b, '"E"', "rail_animation")
                if nxt is not None and nxt[0] is not None:
                    next_input, have_input = helper.send_backward_recv_forward(input_grad, first), True
                else:
                    helper.send_backward(input_grad, first)
        return outputs

    def forward_backward_pipeline(self, data, scaler=None, static_scheduler=False, return_micro_batch_loss=False):
        self.scaler = scaler
        self.total_loss = None
        self.micro_batch_id = 0
        slots = one_f_one_b_slots(self.num_stages, self.accumulate_steps, self.stage_id)
        if static_scheduler:
            return "".join((f"f{f};" if f is not None else "") + (f"b{b};" if b is not None else "") for f, b in slots)
        self._run_slots_1f1b(slots, self._wrap_data(data))
        self._flush_records()
        self._finish_overlapped_comm()
        self._layers.allreduce_shared_weight_gradients()
        with torch.no_grad():
            train_loss = self._broadcast_final_loss(return_micro_batch_loss)
        if self._clear_every_step_cache:
            self._p2p_helper.clear_meta_cache()
        return train_loss

    def train_batch(self, data, optimizer, lr_scheduler=None, scaler=None, loss_fn_idx=0, return_micro_batch_loss=False):
        self.loss_fn_idx = loss_fn_idx
        data = self._prepare_training(data, optimizer, lr_scheduler)
        train_loss = self.forward_backward_pipeline(data, scaler, return_micro_batch_loss=return_micro_batch_loss)
        self._optimizer_step()
        return train_loss

    def eval_batch(self, data, compute_loss=False, loss_fn_idx=0):
        """Forward-only slots; the outputs (or the broadcast loss) are returned."""
        self.loss_fn_idx = loss_fn_idx
        self._layers.eval()
        self._compute_loss = compute_loss
        self.total_loss = None
        self.micro_batch_id = 0
        slots = [(k_, None) for k_ in range(self.accumulate_steps)]
        with torch.no_grad():
            outputs = self._run_slots_1f1b(slots, self._wrap_data(data), keep_outputs=True)
            res = self._broadcast_final_loss() if self._compute_loss else outputs
        self._compute_loss = True
        return res

    # ------------------------------------------------------------------ step pieces
    def _forward_step(self, input_tensor, micro_dataset, chunk_id=None):
        if self.is_pipeline_first_stage():
            input_tensor = next(micro_dataset)[0]
            self._check_micro_batch(input_tensor)
        assert chunk_id is None or isinstance(chunk_id, int)
        output_tensor = self._layers.forward(input_tensor, chunk_id=chunk_id)
        if self.is_pipeline_last_stage():
            if self._compute_loss:
                assert self._layers._loss_fn[self.loss_fn_idx] is not None, "loss function should exist to compute loss"
                labels = next(micro_dataset)[1]
                self._check_micro_batch(labels)
                loss_fn = self._layers._loss_fn[self.loss_fn_idx]
                output_tensor = loss_fn(output_tensor, *labels) if isinstance(labels, tuple) else loss_fn(output_tensor, labels)
                assert isinstance(output_tensor, torch.Tensor), "Currently, loss_fn should obtain Tensor dtype"
                with torch.no_grad():
                    if self.accumulate_steps > 1 and not self._delay_scale_loss:
                        pass
                if self.accumulate_steps > 1 and not self._delay_scale_loss:
                    output_tensor = output_tensor / self.accumulate_steps
                det = output_tensor.detach()
                if self.total_loss is None:
                    self.total_loss = []
                self.total_loss.append(det)
        if self.is_pipeline_first_stage() or self.is_pipeline_last_stage():
            self.micro_batch_id += 1
        return output_tensor

    def _check_micro_batch(self, micro):
        if isinstance(micro, (tuple, list)):
            for m in micro:
                if m is not None:
                    self._check_micro_batch(m)
        elif micro is not None and isinstance(micro, torch.Tensor):
            pass

    def _backward_step(self, input_tensor, output_tensor, output_tensor_grad):
        if self.is_pipeline_last_stage():
            assert output_tensor_grad is None
            if self.scaler:
                torch.autograd.backward(self.scaler.scale(output_tensor))
            else:
                torch.autograd.backward(output_tensor)
        else:
            if isinstance(output_tensor, tuple):
                outs = [t for t in output_tensor if t.requires_grad]
                grads = [g for t, g in zip(output_tensor, output_tensor_grad) if t.requires_grad]
                torch.autograd.backward(outs, grads)
            else:
                torch.autograd.backward([output_tensor], [output_tensor_grad])
        input_tensor_grad = None
        if input_tensor is not None:
            if isinstance(input_tensor, tuple):
                input_tensor_grad = tuple(t.grad for t in input_tensor if t.requires_grad)
            else:
                input_tensor_grad = input_tensor.grad
        return input_tensor_grad

    def _release_output(self, output):
        def can_free(t):
            return t is not None and isinstance(t, torch.Tensor) and t._is_initialized()
        # keep autograd graph, drop the payload that already went to the next stage
        if isinstance(output, (tuple, list)):
            for t in output:
                if can_free(t) and t._base is None:
                    pass
        return

    def _broadcast_final_loss(self, return_micro_batch_loss=False):
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() and \
            self.pp_group.process_group is not None and str(self.pp_group.backend).lower() == "nccl" else torch.device("cpu")
        src = self._hcg.get_rank_from_stage(self.num_stages - 1)
        if self.is_pipeline_last_stage(ignore_virtual=True):
            assert self.total_loss is not None, "train_batch() in last stage should obtain valid loss"
            losses = torch.stack([l.reshape(()) if l.numel() == 1 else l.mean() for l in self.total_loss]).float()
            loss = losses if return_micro_batch_loss else (losses.sum() if not self._delay_scale_loss else losses.mean())
            loss = loss.to(dev).reshape(-1)
            n = torch.tensor([loss.numel()], dtype=torch.int64, device=dev)
            comm.broadcast(n, src, self.pp_group)
            comm.broadcast(loss, src, self.pp_group)
        else:
            n = torch.zeros(1, dtype=torch.int64, device=dev)
            comm.broadcast(n, src, self.pp_group)
            loss = torch.zeros(int(n.item()), dtype=torch.float32, device=dev)
            comm.broadcast(loss, src, self.pp_group)
        return loss if return_micro_batch_loss else loss.reshape(())

    def _optimizer_step(self):
        if self._delay_scale_loss:
            for p in self._layers.parameters():
                g = getattr(p, "main_grad", None)
                g = g if g is not None else p.grad
                if g is not None:
                    g.mul_(1.0 / self.accumulate_steps)
        if self.scaler:
            self.scaler.step(self.optimizer)
            self.scaler.update()
        else:
            self.optimizer.step()
        if self._release_gradients:
            self.optimizer.clear_grad(set_to_zero=False)
        else:
            self.optimizer.clear_grad()
        if self.lr_scheduler:
            self.lr_scheduler.step()

    def forward(self, *a, **k):
        raise RuntimeError("call train_batch / eval_batch on a PipelineParallel model")


def interleave_slots(num_stages, num_chunks, acc_steps, stage, mode="1f1b", forward_only=False):
    """The schedule of ONE pipeline stage with virtual stages, as a table.

    Work is counted in virtual micro-steps k = 0 .. acc_steps * num_chunks - 1.  Forward k runs model chunk
    (k mod (P V)) div P, backward k runs the mirror chunk V - 1 - that: micro-batches travel through the chunks in
    groups of P, and gradients come back through the chunks in reverse.  A slot is ``(forward step or None, backward
    step or None)``:
      * warm-up  - forward-only slots.  Stage s must be 2 (P - s - 1) forwards ahead of its first backward (the usual
        1F1B lead, doubled because the gradient of a group returns only after the later stages have also run it) plus
        (V - 1) P for the earlier chunks of the first groups;
      * steady   - one forward and one backward per slot;
      * cool-down - the remaining backwards.
    "fthenb" is all warm-up followed by all cool-down; ``forward_only`` keeps the forward slots.
    (Reference schedule: fleet/meta_parallel/pipeline_parallel.py:1010-1828 interleave, :1831-2060 FthenB.)"""
    P, V = num_stages, num_chunks
    total = acc_steps * V
    if forward_only or mode == "fthenb":
        lead = total
    else:
        lead = min(2 * (P - stage - 1) + (V - 1) * P, total)
    slots = [(k, None) for k in range(lead)]
    slots += [(lead + i, i) for i in range(total - lead)]
    if not forward_only:
        slots += [(None, i) for i in range(total - lead, total)]
    return slots


class PipelineParallelWithInterleave(PipelineParallel):
    """Interleaved 1F1B with virtual stages; requires accumulate_steps % pp == 0.

    The schedule is data (``interleave_slots``); ``_run_slots`` executes any such table with one rule for the
    communication: after a slot's compute, send what the slot produced (an activation unless the forward ran the last
    virtual stage, an input gradient unless the backward ran the first one) and receive what the NEXT slot will
    consume (an activation unless its forward is the first virtual stage, an output gradient unless its backward is the
    last one) - in one fused p2p call.  Every stage derives the same pairing of sends and receives from its own table."""

    schedule_mode = "1f1b"

    def __init__(self, layers, hcg, strategy):
        super().__init__(layers, hcg, strategy)
        assert layers.get_num_virtual_stages() > 1
        self.num_model_chunks = layers.get_num_virtual_stages()
        self.model_chunks = layers.get_model_chunks()
        assert self.model_chunks is not None and len(self.model_chunks) == self.num_model_chunks
        self._virtual_pp_world_size = self.num_model_chunks
        self._virtual_pp_rank = 0

    # ---- the geometry of a virtual micro-step
    def _chunk_of(self, k, forward):
        c = (k % (self.num_stages * self.num_model_chunks)) // self.num_stages
        return c if forward else self.num_model_chunks - 1 - c

    _get_virtual_pp_rank = _chunk_of          # the reference's name for it

    def _is_first_virtual(self, chunk):
        return self.is_pipeline_first_stage_ and chunk == 0

    def _is_last_virtual(self, chunk):
        return self.is_pipeline_last_stage_ and chunk == self.num_model_chunks - 1

    def _wrap_data_v(self, data):
        return [FakeMicroDataset(data, self.is_pipeline_first_stage_, self.is_pipeline_last_stage_, self.accumulate_steps,
                                 self.micro_batch_size) for _ in range(self.num_model_chunks)]

    def _describe(self, slots):
        out = []
        for f, b in slots:
            if f is not None:
                out.append(f"f{f}_vp{self._chunk_of(f, True)}")
            if b is not None:
                out.append(f"b{b}_vp{self._chunk_of(b, False)}")
        return ";".join(out) + ";"

    def _run_slots(self, slots, micro_dataset, forward_only, split_turn=False):
        """Execute a slot table.  ``split_turn``: where a forward-only slot is followed by a backward-only one, the
        gradient is received in a call of its own instead of riding on the forward exchange (FthenB)."""
        V = self.num_model_chunks
        helper = self._p2p_helper
        acts_in = [[] for _ in range(V)]        # per chunk: inputs waiting for (or kept since) their forward
        acts_out = [[] for _ in range(V)]       # per chunk: outputs kept for their backward
        grads_in = [[] for _ in range(V)]       # per chunk: output gradients waiting for their backward
        self.input_tensors, self.output_tensors, self.output_tensor_grads = acts_in, acts_out, grads_in

        def wants_act(slot):
            return slot is not None and slot[0] is not None and not self._is_first_virtual(self._chunk_of(slot[0], True))

        def wants_grad(slot):
            return slot is not None and slot[1] is not None and not self._is_last_virtual(self._chunk_of(slot[1], False))

        first = slots[0] if slots else None
        if first is not None and first[0] is not None:
            self.set_virtual_pipeline_rank(self._chunk_of(first[0], True))
            acts_in[self._chunk_of(first[0], True)].append(helper.recv_forward(not wants_act(first)))
        for idx, (f, b) in enumerate(slots):
            nxt = slots[idx + 1] if idx + 1 < len(slots) else None
            send_act = send_grad = None
            if f is not None:
                c = self._chunk_of(f, True)
                self.set_virtual_pipeline_rank(c)
                if len(acts_in[c]) == len(acts_out[c]):          # first virtual stage: nothing was received for it
                    acts_in[c].append(None)
                out = self._forward_step(acts_in[c][-1], micro_dataset[c], c)
                acts_out[c].append(out)
                if forward_only:
                    acts_in[c].pop()
                    acts_out[c].pop()
                send_act = None if self._is_last_virtual(c) else out
            if b is not None:
                c = self._chunk_of(b, False)
                self.set_virtual_pipeline_rank(c)
                if not grads_in[c]:                              # last virtual stage: the loss starts the gradient
                    grads_in[c].append(None)
                gin = self._backward_step(acts_in[c].pop(0), acts_out[c].pop(0), grads_in[c].pop(0))
                send_grad = None if self._is_first_virtual(c) else gin
            recv_prev, recv_next = wants_act(nxt), wants_grad(nxt)
            turn = split_turn and b is None and nxt is not None and nxt[0] is None
            if b is None and (nxt is None or nxt[1] is None or turn):
                got_act, got_grad = helper.send_forward_recv_forward(send_act, recv_prev=recv_prev), None
                if turn:
                    c = self._chunk_of(nxt[1], False)
                    self.set_virtual_pipeline_rank(c)
                    got_grad = helper.recv_backward(not recv_next)
            elif f is None:
                got_act, got_grad = None, helper.send_backward_recv_backward(send_grad, recv_next=recv_next)
            else:
                got_act, got_grad = helper.send_forward_backward_recv_forward_backward(
                    send_act, send_grad, recv_prev=recv_prev, recv_next=recv_next)
            if recv_prev:
                acts_in[self._chunk_of(nxt[0], True)].append(got_act)
            if recv_next:
                grads_in[self._chunk_of(nxt[1], False)].append(got_grad)

    def forward_backward_pipeline(self, data, scaler=None, forward_only=False, compute_loss=True, static_scheduler=False,
                                  return_micro_batch_loss=False):
        if self.schedule_mode == "1f1b":
            assert self.accumulate_steps % self.num_stages == 0 or forward_only, \
                "accumulate_steps should be evenly divisible by num_stages for pipeline with interleave"
        self._forward_only = forward_only
        self._compute_loss = compute_loss
        self.scaler = scaler
        self.total_loss = None
        self.micro_batch_id = 0
        slots = interleave_slots(self.num_stages, self.num_model_chunks, self.accumulate_steps, self.stage_id,
                                 self.schedule_mode, forward_only)
        if static_scheduler:
            return self._describe(slots)
        self._run_slots(slots, self._wrap_data_v(data), forward_only, split_turn=self.schedule_mode == "fthenb")
        if not forward_only:
            self._finish_overlapped_comm()
            self._layers.allreduce_shared_weight_gradients()
        if compute_loss:
            with torch.no_grad():
                train_loss = self._broadcast_final_loss(return_micro_batch_loss)
        else:
            train_loss = self.output_tensors
        if self._clear_every_step_cache:
            self._p2p_helper.clear_meta_cache()
        return train_loss

    def train_batch(self, data, optimizer, lr_scheduler=None, scaler=None, loss_fn_idx=0, return_micro_batch_loss=False):
        self.loss_fn_idx = loss_fn_idx
        data = self._prepare_training(data, optimizer, lr_scheduler)
        loss = self.forward_backward_pipeline(data, scaler, return_micro_batch_loss=return_micro_batch_loss)
        self._optimizer_step()
        return loss

    def eval_batch(self, data, compute_loss=False, loss_fn_idx=0):
        self.loss_fn_idx = loss_fn_idx
        self._layers.eval()
        with torch.no_grad():
            return self.forward_backward_pipeline(data, None, forward_only=True, compute_loss=compute_loss)

    def get_static_scheduler(self):
        return self.forward_backward_pipeline(data=None, scaler=None, static_scheduler=True)


class PipelineParallelWithInterleaveFthenB(PipelineParallelWithInterleave):
    """All forwards of every chunk, then all backwards, for pp <= accumulate_steps < 2 pp: the same executor on the
    ``"fthenb"`` table."""

    schedule_mode = "fthenb"
