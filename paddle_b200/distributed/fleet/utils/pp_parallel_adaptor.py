"""Convert pipeline-parallel checkpoints between pipeline layouts (reference:
python/paddle/distributed/fleet/utils/pp_parallel_adaptor.py:24-603): pp degree, virtual-pp (interleave) degree and
segmentation method may all change; mp and sharding degrees stay.

Checkpoint layout (same as the reference): ``<root>/mp_XX_sharding_XX_pp_XX/{model.pdparams, model_state.pdopt,
meta_state.pdopt}``.  Parameter keys inside one stage are ``_layers.<j>.<...>`` (``j`` = global layer index), with
virtual pipeline stages ``_layers.<chunk start>.<j - chunk start>.<...>``, and ``_layers.shared_layers.<name>.<...>`` for
layers shared between stages (tied embeddings), which every stage carries.

Design (differs from the reference, which spills every layer to ``./tmp_layer_files``): all stages of one (mp, sharding)
column are read into memory as a list of per-layer records ``{params, opt, master_weights}``; the records are ordered by
global layer index, re-segmented for the destination layout and written back stage by stage.  Optimizer-state entries
follow their parameter through the ``StructuredToParameterName@@`` table of the saved state dict (tensor name ->
accumulators ``<tensor name>_<acc>``); tensor names are re-numbered per destination stage so that they match what a
freshly built model of that stage would call its parameters (``linear_0.w_0`` ...).
"""
from __future__ import annotations

import argparse
import math
import os
import re
import shutil
from collections import OrderedDict

from paddle_b200.framework import io as _io

_TABLE = "StructuredToParameterName@@"


class ParallelConfig:
    def __init__(self, mp: int, pp: int, vpp: int = 1, sharding: int = 1):
        self.mp, self.pp, self.vpp, self.sharding = mp, pp, vpp, sharding

    def pipe_parallel_group(self, i: int, j: int):
        """(mp, sharding, pp) coordinates of the stages of one pipeline."""
        return [(i, j, k) for k in range(self.pp)]


class LayerReNamingHelper:
    """Consecutive numbers for one kind of layer: every new old-name prefix takes the next index."""

    def __init__(self, template: str):
        self._template = template
        self._i = -1
        self._last = None

    def get_new_layer_name(self, old_layer_name: str):
        old = old_layer_name.split(".")[0]
        if self._last is None or old != self._last:
            self._i += 1
            self._last = old
        return self._template.format(self._i)


class LayerReNamingManager:
    KINDS = ("linear", "layer_norm", "embedding", "rms_norm", "column_parallel_linear", "row_parallel_linear",
             "vocab_parallel_embedding", "column_sequence_parallel_linear", "row_sequence_parallel_linear")

    def __init__(self):
        self._helpers = OrderedDict((k, LayerReNamingHelper(k + "_{}")) for k in sorted(self.KINDS, key=len, reverse=True))

    def start_layer(self):
        """A new pipeline layer begins: equal old names coming from different source stages are different layers."""
        for h in self._helpers.values():
            h._last = None

    def get_new_layer_name(self, old_name: str):
        for kind, helper in self._helpers.items():
            if re.match(rf"^{kind}_\d+", old_name):
                return helper.get_new_layer_name(old_name)
        return ""

    def get_new_param_name(self, old_name: str):
        parts = old_name.split(".")
        new = self.get_new_layer_name(parts[0])
        if not new:
            return old_name                      # unknown kind: keep the name (still unique inside its stage)
        parts[0] = new
        return ".".join(parts)


def _dir(root, coord):
    return "{}/mp_{:0>2d}_sharding_{:0>2d}_pp_{:0>2d}".format(root, *coord)


class PipeLineModelAdaptor:
    def __init__(self, src_parallel_config: ParallelConfig, dst_parallel_config: ParallelConfig,
                 transformer_layer_num: int, segment_method: str = "layer"):
        self._src = src_parallel_config
        self._dst = dst_parallel_config
        self._transformer_layer_num = transformer_layer_num
        self._segment_method = segment_method

    # ------------------------------------------------------------------ driver
    def apply(self, src_model_path: str, dst_model_path: str):
        assert self._src.mp == self._dst.mp and self._src.sharding == self._dst.sharding, \
            "only the pipeline layout (pp, vpp, segmentation) may change"
        for i in range(self._src.mp):
            for j in range(self._src.sharding):
                src_dirs = [_dir(src_model_path, c) for c in self._src.pipe_parallel_group(i, j)]
                layers = []
                for n, d in enumerate(src_dirs):
                    layers.extend(self.extract_layers(d, with_shared=(n == 0)))
                layers = self.sort_layers(layers)
                segments = self.segment_layers(layers, self._dst, self._segment_method)
                dst_dirs = [_dir(dst_model_path, c) for c in self._dst.pipe_parallel_group(i, j)]
                for seg, d in zip(segments, dst_dirs):
                    self.merge_layers(seg, d)
                for k, d in enumerate(dst_dirs):
                    meta = os.path.join(src_dirs[min(k, len(src_dirs) - 1)], "meta_state.pdopt")
                    if os.path.exists(meta):
                        shutil.copyfile(meta, os.path.join(d, "meta_state.pdopt"))

    def peek_model(self, model_dir: str):
        for i in range(self._src.mp):
            for j in range(self._src.sharding):
                for c in self._src.pipe_parallel_group(i, j):
                    d = _dir(model_dir, c)
                    print(f"peek partial model in {d}:")
                    self.peek_partial_model(d)

    def peek_partial_model(self, sub_dir: str):
        sd = _io.load(os.path.join(sub_dir, "model.pdparams"), keep_name_table=True)
        table = sd.get(_TABLE, {})
        for k in sd:
            if k != _TABLE:
                print(f"\t{k} -> {table.get(k, k)}")

    # ------------------------------------------------------------------ read
    @staticmethod
    def _extract_layer_name(param_name: str):
        m = re.search(r"^_layers((\.\d+)+|(\.shared_layers\.[^\.]+))", param_name)
        return m.group() if m else ""

    def extract_layers(self, dir: str, with_shared: bool):
        """-> [(layer name, record)], record = {"params": {key: tensor}, "names": {key: tensor name}, "opt": {...},
        "master_weights": {...}, "LR_Scheduler": ...}."""
        params = _io.load(os.path.join(dir, "model.pdparams"), keep_name_table=True)
        table = dict(params.pop(_TABLE, {}) or {})
        opt_path = os.path.join(dir, "model_state.pdopt")
        opt = _io.load(opt_path) if os.path.exists(opt_path) else {}
        shared_seen = any("_layers.shared_layers" in k for k in params)
        layers = OrderedDict()
        tname_to_layer = {}
        for k, v in params.items():
            layer = self._extract_layer_name(k)
            assert layer, f"{k} is not a pipeline-layer parameter key"
            # a stage that carries the shared embedding also sees it through _layers.0: keep one copy
            if shared_seen and "shared_layers" not in layer and ("word_embeddings" in k or "position_embeddings" in k
                                                                 or "embed_tokens" in k):
                continue
            rec = layers.setdefault(layer, {"params": OrderedDict(), "names": OrderedDict(), "opt": OrderedDict(),
                                            "master_weights": OrderedDict()})
            rec["params"][k] = v
            tname = table.get(k, k)
            rec["names"][k] = tname
            tname_to_layer[tname] = layer
        tnames = sorted(tname_to_layer, key=len, reverse=True)       # longest prefix first
        for k, v in opt.items():
            if k in ("master_weights", "LR_Scheduler"):
                continue
            owner = next((t for t in tnames if k.startswith(t + "_")), None)
            if owner is not None:
                layers[tname_to_layer[owner]]["opt"][k] = v
        for k, v in (opt.get("master_weights") or {}).items():
            if k in tname_to_layer:
                layers[tname_to_layer[k]]["master_weights"][k] = v
        if "LR_Scheduler" in opt:
            for rec in layers.values():
                rec["LR_Scheduler"] = opt["LR_Scheduler"]
        return [(name, rec) for name, rec in layers.items() if with_shared or "shared_layers" not in name]

    # ------------------------------------------------------------------ order + split
    def sort_layers(self, layers: list):
        def order(elem):
            name = elem[0]
            if "shared_layers" in name:
                return (-1.0, 0.0)
            nums = [int(x) for x in name[len("_layers."):].split(".")]
            # virtual-pp keys are (chunk start, offset): the global index is their sum
            return (float(sum(nums)), 0.0)
        layers = sorted(layers, key=order)
        out = []
        for e in layers:
            if out and order(e) == order(out[-1]) and "shared_layers" not in e[0]:
                continue
            if out and e[0] == out[-1][0]:
                continue
            out.append(e)
        return out

    def segment_layers(self, layers: list, config: ParallelConfig, segment_method: str = "layer"):
        """-> per destination pp rank: [([new key prefixes], record)]."""
        shared = [e for e in layers if "shared_layers" in e[0]]
        body = [e for e in layers if "shared_layers" not in e[0]]
        n, stages = len(body), config.pp * config.vpp

        def by_layer():
            w = [0] * n
            heavy = range(1, 1 + self._transformer_layer_num) if self._transformer_layer_num else range(1, n - 1)
            assert not self._transformer_layer_num or self._transformer_layer_num < n
            for i in heavy:
                w[i] = 1
            part = sum(w) // stages
            assert part > 0, f"{sum(w)} transformer layers cannot fill {stages} pipeline stages"
            bounds, acc, b = [0] * (stages + 1), 0, 1
            for idx, x in enumerate(w):
                acc += x
                if acc == part and b < stages:
                    bounds[b] = idx + 1
                    b += 1
                    acc = 0
            bounds[stages] = n
            return bounds

        def uniform():
            bounds = [0] * (stages + 1)
            part, extra = math.floor(n / stages), n % stages
            for i in range(1, stages):
                bounds[i] = int(min(bounds[i - 1] + part + (1 if i > stages - extra else 0), n))
            bounds[stages] = n
            return bounds

        bounds = uniform() if segment_method == "uniform" else by_layer()
        segments = [[] for _ in range(config.pp)]
        for s in range(stages):
            start, end = bounds[s], bounds[s + 1]
            for j in range(start, end):
                prefix = f"_layers.{start}.{j - start}" if config.vpp > 1 else f"_layers.{j}"
                segments[s % config.pp].append(([prefix], body[j][1], body[j][0]))
        for rank in range(config.pp):                 # shared layers travel with every stage
            segments[rank] = [([e[0]], e[1], e[0]) for e in shared] + segments[rank]
        return segments

    # ------------------------------------------------------------------ write
    def merge_layers(self, layers_segment: list, save_dir: str):
        params, table, opt, master = OrderedDict(), OrderedDict(), OrderedDict(), OrderedDict()
        renamer = LayerReNamingManager()
        lr = None
        for prefixes, rec, old_layer in layers_segment:
            renamer.start_layer()
            tname_map = {}
            for k, v in rec["params"].items():
                old_t = rec["names"][k]
                new_t = tname_map.setdefault(old_t, renamer.get_new_param_name(old_t))
                v.name = new_t                     # paddle.save writes the name table from the tensors' names
                for prefix in prefixes:
                    new_k = prefix + k[len(old_layer):]
                    params[new_k] = v
                    table[new_k] = new_t
            olds = sorted(tname_map, key=len, reverse=True)
            for k, v in rec["opt"].items():
                owner = next((t for t in olds if k.startswith(t + "_")), None)
                opt[(tname_map[owner] + k[len(owner):]) if owner else k] = v
            for k, v in rec["master_weights"].items():
                master[tname_map.get(k, k)] = v
            lr = rec.get("LR_Scheduler", lr)
        os.makedirs(save_dir, exist_ok=True)
        _io.save(params, os.path.join(save_dir, "model.pdparams"))
        if master:
            opt["master_weights"] = master
        if lr is not None:
            opt["LR_Scheduler"] = lr
        _io.save(opt, os.path.join(save_dir, "model_state.pdopt"))


def parse_args():
    p = argparse.ArgumentParser(prog="model converter", description="convert a pipeline-parallel checkpoint to another layout")
    p.add_argument("--src_path", type=str, default="./output/epoch_0_step_30")
    p.add_argument("--dst_path", type=str, default="./test_adapt")
    p.add_argument("--src_mp", type=int, default=2)
    p.add_argument("--src_pp", type=int, default=2)
    p.add_argument("--src_vp", type=int, default=2)
    p.add_argument("--dst_mp", type=int, default=None)
    p.add_argument("--dst_pp", type=int, default=None)
    p.add_argument("--dst_vp", type=int, default=2)
    p.add_argument("--sharding", type=int, default=1)
    p.add_argument("--method", type=str, default="adapt_model", choices=["adapt_model", "peek_model"])
    p.add_argument("--segment_method", type=str, default="layer", choices=["layer", "uniform"])
    p.add_argument("--transformer_layer_num", type=int, default=0)
    args = p.parse_args()
    args.dst_mp = args.src_mp if args.dst_mp is None else args.dst_mp
    args.dst_pp = args.src_pp if args.dst_pp is None else args.dst_pp
    assert args.src_mp == args.dst_mp, f"src mp {args.src_mp} dst mp {args.dst_mp}"
    return args


def adaptor_from_args(args):
    return PipeLineModelAdaptor(ParallelConfig(args.src_mp, args.src_pp, args.src_vp, args.sharding),
                                ParallelConfig(args.dst_mp, args.dst_pp, args.dst_vp, args.sharding),
                                args.transformer_layer_num, args.segment_method)


def main():
    args = parse_args()
    adaptor = adaptor_from_args(args)
    if args.method == "peek_model":
        adaptor.peek_model(args.dst_path)
    else:
        adaptor.apply(args.src_path, args.dst_path)


if __name__ == "__main__":
    main()
