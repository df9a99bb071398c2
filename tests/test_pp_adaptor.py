"""pp_parallel_adaptor: a pp=2 checkpoint (embedding + 4 blocks + head, AdamW state with master weights) converted to
pp=4, to pp=2 x vpp=2 and back to pp=2 keeps every tensor and every optimizer entry attached to the same layer."""
import os

import torch

import paddle_b200 as paddle
from paddle_b200.distributed.fleet.utils.pp_parallel_adaptor import ParallelConfig, PipeLineModelAdaptor


def _make_stage(root, pp_rank, layer_ids, counters):
    d = os.path.join(root, f"mp_00_sharding_00_pp_{pp_rank:02d}")
    os.makedirs(d, exist_ok=True)
    params, opt, master = {}, {}, {}
    for j in layer_ids:
        kind = "embedding" if j == 0 else "linear"
        n = counters.setdefault((pp_rank, kind), 0)
        counters[(pp_rank, kind)] = n + 1
        tname = f"{kind}_{n}.w_0"
        w = torch.full((4, 4), float(j))
        w.name = tname
        params[f"_layers.{j}.weight"] = w
        opt[f"{tname}_moment1_0"] = torch.full((4, 4), j + 0.25)
        opt[f"{tname}_moment2_0"] = torch.full((4, 4), j + 0.5)
        opt[f"{tname}_beta1_pow_acc_0"] = torch.tensor([0.9])
        master[tname] = torch.full((4, 4), j + 0.125)
    opt["master_weights"] = master
    opt["LR_Scheduler"] = {"last_epoch": 7, "last_lr": 0.01}
    paddle.save(params, os.path.join(d, "model.pdparams"))
    paddle.save(opt, os.path.join(d, "model_state.pdopt"))
    paddle.save({"epoch": 1}, os.path.join(d, "meta_state.pdopt"))


def _read(root, pp):
    """{global layer index: (weight value, moment1 value, master value)} from every stage, through the name table."""
    out = {}
    for r in range(pp):
        d = os.path.join(root, f"mp_00_sharding_00_pp_{r:02d}")
        sd = paddle.load(os.path.join(d, "model.pdparams"), keep_name_table=True)
        table = sd.pop("StructuredToParameterName@@")
        opt = paddle.load(os.path.join(d, "model_state.pdopt"))
        assert opt["LR_Scheduler"]["last_epoch"] == 7
        assert os.path.exists(os.path.join(d, "meta_state.pdopt"))
        names = set(table.values())
        assert len(names) == len(table), "tensor names must be unique inside a stage"
        for k, v in sd.items():
            idx = sum(int(x) for x in k.split(".")[1:-1])
            t = table[k]
            out[idx] = (float(v[0, 0]), float(opt[f"{t}_moment1_0"][0, 0]), float(opt["master_weights"][t][0, 0]), k, r)
    return out


def test_convert_pp2_to_pp4_to_vpp_and_back(tmp_path):
    src = str(tmp_path / "pp2")
    counters = {}
    _make_stage(src, 0, [0, 1, 2], counters)          # embedding + blocks 1, 2
    _make_stage(src, 1, [3, 4, 5], counters)          # blocks 3, 4 + head
    want = {j: (float(j), j + 0.25, j + 0.125) for j in range(6)}

    pp4 = str(tmp_path / "pp4")
    PipeLineModelAdaptor(ParallelConfig(1, 2), ParallelConfig(1, 4), transformer_layer_num=4).apply(src, pp4)
    got = _read(pp4, 4)
    assert {j: v[:3] for j, v in got.items()} == want
    assert [got[j][4] for j in range(6)] == [0, 0, 1, 2, 3, 3]        # one block per stage, embedding first, head last

    vpp = str(tmp_path / "pp2v2")
    PipeLineModelAdaptor(ParallelConfig(1, 4), ParallelConfig(1, 2, vpp=2), transformer_layer_num=4).apply(pp4, vpp)
    got = _read(vpp, 2)
    assert {j: v[:3] for j, v in got.items()} == want
    assert got[2][3].count(".") == 3 and [got[j][4] for j in range(6)] == [0, 0, 1, 0, 1, 1]   # chunks alternate between ranks

    back = str(tmp_path / "back")
    PipeLineModelAdaptor(ParallelConfig(1, 2, vpp=2), ParallelConfig(1, 2), transformer_layer_num=4).apply(vpp, back)
    got = _read(back, 2)
    assert {j: v[:3] for j, v in got.items()} == want
    assert [got[j][4] for j in range(6)] == [0, 0, 0, 1, 1, 1]
