"""PipelineLayer / LayerDesc / SharedLayerDesc / SegmentLayers
(reference: python/paddle/distributed/fleet/meta_parallel/parallel_layers/pp_layers.py:56-934)."""
from __future__ import annotations

import glob
import math
import os
import re
from functools import partial

import torch

from paddle_b200.distributed import communication as comm
from paddle_b200.framework import io as _io
from paddle_b200.nn.layer.layers import Layer, LayerDict


class LayerDesc:
    def __init__(self, layer_func, *inputs, **kwargs):
        self.layer_func = layer_func
        self.inputs = inputs
        self.kwargs = kwargs
        if not issubclass(layer_func, Layer) and not issubclass(layer_func, torch.nn.Module):
            raise TypeError("The input(layer_func) should be a derived class of Layer.")

    def build_layer(self):
        return self.layer_func(*self.inputs, **self.kwargs)

    def __repr__(self):
        return f"{self.layer_func.__name__}({', '.join(map(str, self.inputs))})"


class SharedLayerDesc(LayerDesc):
    def __init__(self, key, layer_func, forward_func=None, shared_weight_attr="weight", *inputs, **kwargs):
        super().__init__(layer_func, *inputs, **kwargs)
        self.layer_name = key
        self.forward_func = forward_func
        self.shared_weight_attr = shared_weight_attr


class SegmentLayers:
    def __init__(self, layers_desc, num_parts, method="uniform", num_virtual_pipeline_stage=None):
        self._layers_desc = layers_desc
        self.method = method
        self.num_parts = num_parts
        self.num_items = len(layers_desc)
        self.num_virtual_pipeline_stage = num_virtual_pipeline_stage
        if num_virtual_pipeline_stage is not None:
            self.total_parts = num_parts * num_virtual_pipeline_stage
        assert self.num_items >= self.num_parts, "layer number should be greater than number of segments"

    def do_segment(self):
        if isinstance(self.method, list):
            seg = list(self.method)
            assert seg[0] == 0 and len(seg) == self.num_parts + 1, "seg_method[0] should be 0 and length num_parts+1"
            return seg
        if self.method == "uniform":
            return self.uniform(self.num_items, self.num_parts)
        if self.method.startswith("layer:"):
            name = self.method.split(":")[1]
            weights = [0] * len(self._layers_desc)
            for idx in self._gen_layer_weight(name):
                weights[idx] = 1
            actual = self.num_parts if self.num_virtual_pipeline_stage is None else self.total_parts
            assert sum(weights) % actual == 0, f"number of layers ({sum(weights)}) should be divided by part number({actual})"
            part = sum(weights) // actual
            result = [0] * (actual + 1)
            mem, cur = 0, 1
            for idx, w in enumerate(weights):
                mem += w
                if mem == part:
                    result[cur] = idx + 1
                    cur += 1
                    mem = 0
            result[actual] = len(weights)
            return result
        raise ValueError(f"method {self.method} is not supported")

    def _gen_layer_weight(self, layername):
        regex = re.compile(layername, re.IGNORECASE)
        idxs = []
        for idx, layer in enumerate(self._layers_desc):
            name = None
            if isinstance(layer, (Layer, torch.nn.Module)):
                name = layer.__class__.__name__
            elif isinstance(layer, LayerDesc):
                name = layer.layer_func.__name__
            else:
                try:
                    name = layer.__name__
                except AttributeError:
                    continue
            if regex.search(name):
                idxs.append(idx)
        assert idxs, "weight_idxs' length should be greater than 0"
        return idxs

    @staticmethod
    def uniform(num_items, num_parts):
        result = [0] * (num_parts + 1)
        part = math.floor(num_items / num_parts)
        extra = num_items % num_parts
        for i in range(1, num_parts):
            offset = 1 if i > (num_parts - extra) else 0
            result[i] = int(min(result[i - 1] + part + offset, num_items))
        result[num_parts] = num_items
        return result


class PipelineLayerChunk(Layer):
    def __init__(self):
        super().__init__()
        self.run_function = []

    def append(self, sublayer):
        if isinstance(sublayer, (Layer, torch.nn.Module)):
            self.add_sublayer(str(len(self.run_function)), sublayer)
        self.run_function.append(sublayer)

    def get_run_function(self):
        return self.run_function

    def forward(self, *args, **kwargs):
        raise PermissionError("The forward function of PipelineLayerChunk cannot be called directly. "
                              "Please call forward function of PipelineLayer.")


class PipelineLayer(Layer):
    def __init__(self, layers, num_stages=None, topology=None, loss_fn=None, seg_method="uniform",
                 recompute_interval=0, recompute_ctx=None, num_virtual_pipeline_stages=None, use_cudagraph=False):
        super().__init__()
        if num_stages is None and topology is None:
            raise ValueError("should provide num_stages or topology")
        if num_virtual_pipeline_stages:
            assert isinstance(num_virtual_pipeline_stages, int), "virtual_pipeline_stage should be None or an int"
            if num_virtual_pipeline_stages > 1:
                assert isinstance(seg_method, str) and seg_method.startswith("layer:"), \
                    "seg_method should be a str for interleave scheduler"
        self._num_virtual_pipeline_stages = 1 if num_virtual_pipeline_stages is None else num_virtual_pipeline_stages
        self.device_id = 0
        self.layers = layers
        self._loss_fn = loss_fn if isinstance(loss_fn, list) else [loss_fn]
        self._topo = topology
        self._recompute_interval = recompute_interval
        self.recompute_ctx = recompute_ctx
        if recompute_interval > 0:
            assert recompute_ctx is not None, "recompute_ctx must be not None for recompute."
        from paddle_b200.distributed import collective as C
        from paddle_b200.distributed import fleet
        self.world_size = C.get_world_size()
        self.global_rank = C.get_rank()
        hcg = fleet.get_hybrid_communicate_group()
        if self._topo is None and hcg is not None:
            self._topo = hcg.topology()
        if self._topo is not None:
            self._stage_id = self._topo.get_coord(self.global_rank).pipe
            self._num_stages = self._topo.get_dim_size("pipe")
            if num_stages:
                assert self._num_stages == num_stages, f"num_stages should be equal to be {self._num_stages}"
        else:
            self._num_stages = num_stages
            self._stage_id = self.global_rank % num_stages if self.world_size > 1 else 0
        self._total_stages_with_virtual_stages = self._num_stages * self._num_virtual_pipeline_stages
        self._layers_desc = list(self.layers)
        self._num_layers = len(self._layers_desc)
        self.shared_layers = LayerDict()
        self.shared_weight_attrs = {}
        self.shared_comm = {}
        if self._num_virtual_pipeline_stages > 1:
            self._start_poss, self._end_poss = [], []
            self._segment_network_for_interleave(seg_method)
            self._model_chunks = []
            self._build_layer_with_interleave()
        else:
            self._start_pos, self._end_pos = 0, self._num_layers
            self._segment_network(seg_method)
            self.run_function = []
            self._build_layer()
        self.shared_comm = self._construct_shared_comm()
        self._synchronize_shared_weights()

    # ------------------------------------------------------------------ queries
    def get_stage_from_index(self, layer_idx):
        assert 0 <= layer_idx < self._num_layers, "layer_idx is out of bound"
        for vpp in range(self._num_virtual_pipeline_stages):
            start = vpp * self._num_stages
            for stage in range(self._num_stages):
                if self.segment_parts[start + stage] <= layer_idx < self.segment_parts[start + stage + 1]:
                    return stage

    def get_num_virtual_stages(self):
        return self._num_virtual_pipeline_stages

    def get_model_chunks(self):
        return None if self._num_virtual_pipeline_stages == 1 else self._model_chunks

    def get_num_stages(self):
        return self._num_stages

    # ------------------------------------------------------------------ shared (tied) weights
    def _construct_shared_comm(self):
        shared_comm = {}
        if self._topo is None or self._topo.get_dim("pipe") == 1:
            return shared_comm
        from paddle_b200.distributed import collective as C
        layers_desc = self._layers_desc
        shared_names = {s.layer_name for s in layers_desc if isinstance(s, SharedLayerDesc)}
        for key in shared_names:
            shared_layers = [i for i, l in enumerate(layers_desc) if isinstance(l, SharedLayerDesc) and l.layer_name == key]
            shared_stages = sorted({self.get_stage_from_index(i) for i in shared_layers})
            names = self._topo.get_hybrid_group_names()
            others = [n for n in names if n != "pipe"]
            import itertools
            for combo in itertools.product(*[range(self._topo.get_dim(n)) for n in others]):
                kw = dict(zip(others, combo))
                ranks = sorted(self._topo.get_rank(pipe=s, **kw) for s in shared_stages)
                group = C.new_group(ranks=ranks)
                if self.global_rank in ranks:
                    assert key in self.shared_layers
                    shared_comm[key] = {"ranks": ranks, "group": group,
                                        "weight_attr": self.shared_weight_attrs[key], "layer": self.shared_layers[key]}
        return shared_comm

    def _synchronize_shared_weights(self):
        for key, c in self.shared_comm.items():
            with torch.no_grad():
                comm.broadcast(getattr(c["layer"], c["weight_attr"]), min(c["ranks"]), c["group"])
            for p in c["layer"].parameters():
                if self.global_rank != min(c["ranks"]):
                    p.is_firstly_shared = False

    def allreduce_shared_weight_gradients(self):
        for key, c in self.shared_comm.items():
            p = getattr(self.shared_layers[key], c["weight_attr"])
            g = getattr(p, "main_grad", None)
            g = g if g is not None else p.grad
            if g is None:
                g = torch.zeros_like(p, dtype=torch.float32 if hasattr(p, "main_grad") else p.dtype)
                if hasattr(p, "main_grad"):
                    p.main_grad = g
                else:
                    p.grad = g
            with torch.no_grad():
                comm.all_reduce(g, group=c["group"])

    # ------------------------------------------------------------------ segmentation / build
    def _segment_network_for_interleave(self, seg_method):
        seg = SegmentLayers(self._layers_desc, num_parts=self._num_stages, method=seg_method,
                            num_virtual_pipeline_stage=self._num_virtual_pipeline_stages)
        self.segment_parts = seg.do_segment()
        for i in range(self._stage_id, self._total_stages_with_virtual_stages, self._num_stages):
            self._start_poss.append(self.segment_parts[i])
            self._end_poss.append(self.segment_parts[i + 1])
        assert len(self._start_poss) == len(self._end_poss)

    def _segment_network(self, seg_method):
        seg = SegmentLayers(self._layers_desc, num_parts=self._num_stages, method=seg_method)
        self.segment_parts = seg.do_segment()
        self._start_pos = self.segment_parts[self._stage_id]
        self._end_pos = self.segment_parts[self._stage_id + 1]

    def _build_layer_with_interleave(self):
        for i in range(len(self._start_poss)):
            chunk = PipelineLayerChunk()
            for fn in self._build_layer_impl(self._start_poss[i], self._end_poss[i]):
                chunk.append(fn)
            self._model_chunks.append(chunk)
            self.add_sublayer(f"chunk_{i}", chunk)

    def _build_layer(self):
        self.run_function = self._build_layer_impl(self._start_pos, self._end_pos)

    def _build_layer_impl(self, start, end):
        run_function = []
        for index, layer in enumerate(self._layers_desc[start:end]):
            layer_index = start + index
            if isinstance(layer, (Layer, torch.nn.Module)):
                run_function.append(layer)
                if self._num_virtual_pipeline_stages == 1:
                    self.add_sublayer(str(layer_index), layer)
            elif isinstance(layer, SharedLayerDesc):
                if layer.layer_name not in self.shared_layers:
                    self.shared_layers[layer.layer_name] = layer.build_layer()
                    self.shared_weight_attrs[layer.layer_name] = layer.shared_weight_attr
                    for p in self.shared_layers[layer.layer_name].parameters():
                        p.is_firstly_shared = True
                if layer.forward_func is None:
                    run_function.append(self.shared_layers[layer.layer_name])
                else:
                    run_function.append(partial(layer.forward_func, self.shared_layers[layer.layer_name]))
            elif isinstance(layer, LayerDesc):
                model = layer.build_layer()
                run_function.append(model)
                if self._num_virtual_pipeline_stages == 1:
                    self.add_sublayer(str(layer_index), model)
            else:
                run_function.append(layer)
        return run_function

    # ------------------------------------------------------------------ forward
    def forward_function(self, start, end):
        def execute_func(*x):
            if len(x) == 1:
                x = x[0]
            for layer in self.run_function[start:end]:
                x = layer(*x) if isinstance(x, tuple) else layer(x)
            return x
        return execute_func

    def forward(self, input, chunk_id=None):
        if chunk_id is not None:
            assert isinstance(chunk_id, int) and self._num_virtual_pipeline_stages > 1
            assert chunk_id < len(self._model_chunks)
            run_function = self._model_chunks[chunk_id].get_run_function()
        else:
            run_function = self.run_function
        if self._recompute_interval == 0:
            for layer in run_function:
                input = layer(*input) if isinstance(input, tuple) else layer(input)
            return input
        from paddle_b200.distributed.fleet.recompute import recompute
        n = len(run_function)
        for start in range(0, n, self._recompute_interval):
            end = min(start + self._recompute_interval, n)
            funcs = run_function[start:end]

            def seg(*x, _funcs=funcs):
                if len(x) == 1:
                    x = x[0]
                for l in _funcs:
                    x = l(*x) if isinstance(x, tuple) else l(x)
                return x
            need = any(isinstance(f, torch.nn.Module) and any(p.requires_grad for p in f.parameters()) for f in funcs)
            if need:
                args = input if isinstance(input, tuple) else (input,)
                input = recompute(seg, *args, preserve_rng_state=self.recompute_ctx.get("preserve_rng_state", True))
            else:
                input = seg(*input) if isinstance(input, tuple) else seg(input)
        return input

    # ------------------------------------------------------------------ per-layer checkpoints (:841-934)
    def _ckpt_name(self, idx, vpp=None):
        from paddle_b200.distributed import fleet
        hcg = fleet.get_hybrid_communicate_group()
        mp_rank = hcg.get_model_parallel_rank() if hcg is not None else 0
        s = f"layer_{idx:02d}"
        if vpp is not None:
            s += f"-virtual_pp_stage_{vpp:02d}"
        return s + f"-tensor_{mp_rank:02d}-model_states.pdparams"

    def save_state_dict(self, path):
        from paddle_b200.distributed import fleet
        hcg = fleet.get_hybrid_communicate_group()
        if hcg is not None and hcg.get_data_parallel_rank() != 0:
            return
        os.makedirs(path, exist_ok=True)

        def save(run_function, start, vpp=None):
            for i, layer in enumerate(run_function):
                if isinstance(layer, torch.nn.Module) and hasattr(layer, "state_dict"):
                    _io.save(layer.state_dict(), os.path.join(path, self._ckpt_name(start + i, vpp)))
        if self._num_virtual_pipeline_stages > 1:
            for v, chunk in enumerate(self._model_chunks):
                save(chunk.get_run_function(), self._start_poss[v], v)
        else:
            save(self.run_function, self._start_pos)

    def set_state_dir(self, path):
        assert os.path.exists(path), f"{path} not found, please check the path"

        def load(run_function, start, vpp=None):
            for i, layer in enumerate(run_function):
                if not (isinstance(layer, torch.nn.Module) and hasattr(layer, "set_state_dict")):
                    continue
                pattern = f"layer_{start + i:02d}" + (f"-virtual_pp_stage_{vpp:02d}" if vpp is not None else "") + "-tensor_*-model_states.pdparams"
                files = sorted(glob.glob(os.path.join(path, pattern)))
                if not files:
                    continue
                from paddle_b200.distributed import fleet
                hcg = fleet.get_hybrid_communicate_group()
                mp_rank = hcg.get_model_parallel_rank() if hcg else 0
                mp_world = hcg.get_model_parallel_world_size() if hcg else 1
                layer.set_state_dict(_io.load(files[mp_rank * len(files) // mp_world]))
        if self._num_virtual_pipeline_stages > 1:
            for v, chunk in enumerate(self._model_chunks):
                load(chunk.get_run_function(), self._start_poss[v], v)
        else:
            load(self.run_function, self._start_pos)
