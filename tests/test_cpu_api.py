"""CPU tests of the Paddle API surface on torch (no GPU, single process)."""
import os
import pickle

import numpy as np
import pytest
import torch

import paddle_b200 as paddle
from paddle_b200 import nn


def test_alias_is_same_module():
    import paddle as p2
    import paddle.distributed.fleet as f2
    assert p2 is paddle and f2 is paddle.distributed.fleet
    from paddle.distributed.checkpoint.metadata import Metadata
    assert Metadata.__module__ == "paddle.distributed.checkpoint.metadata"


def test_tensor_api_basics():
    x = paddle.to_tensor([[1.0, 2.0], [3.0, 4.0]])
    assert x.stop_gradient and x.dtype == paddle.float32
    x.stop_gradient = False
    y = paddle.matmul(x, x, transpose_y=True).sum()
    y.backward()
    assert x.grad is not None
    assert paddle.split(paddle.arange(12).reshape([3, 4]), 2, axis=1)[0].shape == (3, 2)
    assert paddle.concat([x, x], axis=0).shape == (4, 2)
    assert x.astype("float16").dtype == paddle.float16
    assert x.transpose([1, 0]).shape == (2, 2)
    assert paddle.max(x, axis=1).shape == (2,)
    bf = paddle.ones([2, 2], dtype="bfloat16")
    assert bf.numpy().dtype == np.uint16
    assert isinstance(x.place, paddle.CPUPlace().__class__.__mro__[1])
    x.clear_grad()
    assert x.grad is None


def test_layer_state_dict_names_and_roundtrip(tmp_path):
    with paddle.utils.unique_name.guard():
        m = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 2))
    sd = m.state_dict()
    assert list(sd) == ["0.weight", "0.bias", "2.weight", "2.bias"]
    assert sd["0.weight"].name == "linear_0.w_0" and sd["2.bias"].name == "linear_1.b_0"
    assert tuple(sd["0.weight"].shape) == (4, 8)          # paddle layout [in, out]
    path = str(tmp_path / "m.pdparams")
    paddle.save(sd, path)
    raw = pickle.load(open(path, "rb"))
    assert isinstance(raw["0.weight"], np.ndarray) and raw["StructuredToParameterName@@"]["0.weight"] == "linear_0.w_0"
    with paddle.utils.unique_name.guard():
        m2 = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 2))
    missing, unexpected = m2.set_state_dict(paddle.load(path))
    assert not missing and not unexpected
    x = paddle.randn([3, 4])
    torch.testing.assert_close(m(x), m2(x))


def test_save_load_bf16_and_nested(tmp_path):
    obj = {"a": paddle.ones([3], dtype="bfloat16"), "n": {"b": paddle.arange(4)}, "k": 3}
    p = str(tmp_path / "o.pd")
    paddle.save(obj, p)
    back = paddle.load(p)
    assert back["a"].dtype == paddle.bfloat16 and back["k"] == 3 and back["n"]["b"].tolist() == [0, 1, 2, 3]
    lst = [paddle.ones([2]), {"z": paddle.zeros([1])}]
    paddle.save(lst, p)
    b2 = paddle.load(p)
    assert b2[0].tolist() == [1, 1] and b2[1]["z"].tolist() == [0]


def test_hooks_and_containers():
    l = nn.Linear(2, 2)
    seen = []
    h1 = l.register_forward_pre_hook(lambda layer, inp: seen.append("pre"))
    h2 = l.register_forward_post_hook(lambda layer, inp, out: seen.append("post"))
    l(paddle.ones([1, 2]))
    h1.remove(); h2.remove()
    l(paddle.ones([1, 2]))
    assert seen == ["pre", "post"]
    ll = nn.LayerList([nn.Linear(2, 2) for _ in range(3)])
    assert len(ll) == 3 and len(ll.parameters()) == 6
    assert len(list(ll.named_sublayers())) == 3


@pytest.mark.parametrize("mp", [False, True])
def test_adamw_matches_torch(mp):
    torch.manual_seed(0)
    dt = torch.bfloat16 if mp else torch.float32
    ps = [paddle.framework.Parameter(torch.randn(17, 5).to(dt)), paddle.framework.Parameter(torch.randn(9).to(dt))]
    rs = [p.detach().float().clone().requires_grad_(True) for p in ps]
    opt = paddle.optimizer.AdamW(1e-2, parameters=ps, weight_decay=0.05, beta2=0.98, multi_precision=mp)
    ref = torch.optim.AdamW(rs, lr=1e-2, weight_decay=0.05, betas=(0.9, 0.98), eps=1e-8)
    for _ in range(5):
        for p, r in zip(ps, rs):
            g = torch.randn_like(r)
            p.grad, r.grad = g.to(dt), g.to(dt).float()
        opt.step(); ref.step()
    for p, r in zip(ps, rs):
        tgt = opt._master_weights[p.name] if mp else p
        torch.testing.assert_close(tgt.detach().float(), r.detach(), rtol=1e-5, atol=1e-6)
    sd = opt.state_dict()
    assert f"{ps[0].name}_moment1_0" in sd and f"{ps[0].name}_beta1_pow_acc_0" in sd
    assert abs(float(sd[f"{ps[0].name}_beta1_pow_acc_0"]) - 0.9 ** 6) < 1e-6
    if mp:
        assert ps[0].name in sd["master_weights"]


def test_optimizer_state_dict_roundtrip():
    def make():
        with paddle.utils.unique_name.guard():
            m = nn.Linear(3, 3)
        return m, paddle.optimizer.AdamW(1e-2, parameters=m.parameters())
    m1, o1 = make()
    x = paddle.randn([4, 3])
    for _ in range(3):
        m1(x).sum().backward(); o1.step(); o1.clear_grad()
    m2, o2 = make()
    m2.set_state_dict(m1.state_dict())
    o2.set_state_dict({k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in o1.state_dict().items()})
    for m, o in ((m1, o1), (m2, o2)):
        m(x).sum().backward(); o.step(); o.clear_grad()
    torch.testing.assert_close(m1.weight.detach(), m2.weight.detach())


def test_momentum_and_lr_scheduler():
    p = paddle.framework.Parameter(torch.ones(4))
    sched = paddle.optimizer.lr.StepDecay(0.1, step_size=2, gamma=0.5)
    opt = paddle.optimizer.Momentum(sched, momentum=0.9, parameters=[p])
    r = torch.ones(4, requires_grad=True)
    ro = torch.optim.SGD([r], lr=0.1, momentum=0.9)
    for i in range(4):
        p.grad = torch.full((4,), 0.5); r.grad = torch.full((4,), 0.5)
        for g in ro.param_groups:
            g["lr"] = sched()
        opt.step(); ro.step(); sched.step()
    torch.testing.assert_close(p.detach(), r.detach())
    assert abs(sched() - 0.025) < 1e-9
    cos = paddle.optimizer.lr.CosineAnnealingDecay(1.0, T_max=10)
    w = paddle.optimizer.lr.LinearWarmup(cos, warmup_steps=2, start_lr=0.0, end_lr=1.0)
    vals = []
    for _ in range(4):
        vals.append(w()); w.step()
    assert vals[0] == 0.0 and abs(vals[1] - 0.5) < 1e-9 and abs(vals[2] - 1.0) < 1e-9


def test_global_norm_clip():
    ps = [paddle.framework.Parameter(torch.ones(10)), paddle.framework.Parameter(torch.ones(5))]
    for p in ps:
        p.grad = torch.full_like(p, 2.0)
    clip = nn.ClipGradByGlobalNorm(1.0)
    clip([(p, p.grad) for p in ps])
    total = torch.sqrt(sum(p.grad.pow(2).sum() for p in ps))
    assert abs(float(total) - 1.0) < 1e-5


def test_grad_scaler_dynamic():
    p = paddle.framework.Parameter(torch.ones(3))
    opt = paddle.optimizer.SGD(0.1, parameters=[p])
    sc = paddle.amp.GradScaler(init_loss_scaling=8.0, incr_every_n_steps=2, decr_every_n_nan_or_inf=1, incr_ratio=2.0, decr_ratio=0.5)
    loss = (p * 2).sum()
    sc.scale(loss).backward()
    sc.step(opt); sc.update()
    torch.testing.assert_close(p.detach(), torch.full((3,), 0.8))
    opt.clear_grad()
    p.grad = torch.tensor([float("inf"), 0, 0])
    sc.step(opt); sc.update()
    torch.testing.assert_close(p.detach(), torch.full((3,), 0.8))
    assert float(sc.get_loss_scaling()) == 4.0


def test_amp_autocast_and_decorate():
    m = nn.Linear(4, 4)
    with paddle.amp.auto_cast(level="O1", dtype="bfloat16"):
        y = m(paddle.randn([2, 4]))
    assert y.dtype in (torch.bfloat16, torch.float32)
    opt = paddle.optimizer.AdamW(1e-3, parameters=m.parameters())
    m, opt = paddle.amp.decorate(m, opt, level="O2", dtype="bfloat16")
    assert m.weight.dtype == torch.bfloat16 and opt._multi_precision


def test_flags():
    paddle.set_flags({"FLAGS_check_nan_inf": True})
    assert paddle.get_flags("FLAGS_check_nan_inf")["FLAGS_check_nan_inf"] is True
    paddle.set_flags({"FLAGS_check_nan_inf": False})
    with pytest.raises(ValueError):
        paddle.set_flags({"FLAGS_not_a_flag": 1})


def test_recompute_matches_plain():
    from paddle_b200.distributed.fleet.recompute import recompute, recompute_sequential
    torch.manual_seed(0)
    blk = nn.Sequential(nn.Linear(8, 8), nn.GELU(), nn.Dropout(0.5), nn.Linear(8, 8))
    x = paddle.randn([4, 8]); x.stop_gradient = False
    paddle.seed(7)
    y1 = blk(x); y1.sum().backward()
    g1 = [p.grad.clone() for p in blk.parameters()]; gx1 = x.grad.clone()
    blk.clear_gradients(); x.clear_grad()
    paddle.seed(7)
    y2 = recompute(blk, x); y2.sum().backward()
    torch.testing.assert_close(y1, y2)
    torch.testing.assert_close(gx1, x.grad)
    for a, p in zip(g1, blk.parameters()):
        torch.testing.assert_close(a, p.grad)
    blk.clear_gradients()
    y3 = recompute_sequential({"segments": 2}, [blk[0], blk[1], blk[3]], x)
    assert y3.shape == (4, 8)


def test_topology_and_strategy():
    from paddle_b200.distributed.fleet.base.topology import CommunicateTopology
    from paddle_b200.distributed import fleet
    topo = CommunicateTopology(["data", "pipe", "sharding", "sep", "model"], [2, 2, 1, 1, 2])
    assert topo.world_size() == 8
    assert topo.get_rank(data=1, pipe=0, sharding=0, sep=0, model=1) == 5
    assert topo.get_comm_list("model") == [[0, 1], [2, 3], [4, 5], [6, 7]]
    assert topo.get_comm_list("pipe") == [[0, 2], [1, 3], [4, 6], [5, 7]]
    assert topo.get_axis_list("data", 0) == [0, 1, 2, 3]
    assert sorted(map(tuple, topo.get_fused_ranks(["data", "model"]))) == [(0, 1, 4, 5), (2, 3, 6, 7)]
    st = fleet.DistributedStrategy()
    st.hybrid_configs = {"dp_degree": 2, "mp_degree": 4, "pp_configs": {"dp_comm_overlap": True}}
    assert st.hybrid_configs["mp_degree"] == 4 and st.hybrid_configs["pp_configs"]["dp_comm_overlap"] is True
    assert st.hybrid_configs["mp_configs"]["sync_param"] is True and st.fuse_grad_size_in_MB == 32
    with pytest.raises(ValueError):
        st.hybrid_configs = {"bogus": 1}
    st.pipeline_configs = {"accumulate_steps": 4, "micro_batch_size": 2}
    assert st.pipeline_configs["schedule_mode"] == "1F1B"


def test_segment_layers():
    from paddle_b200.distributed.fleet.meta_parallel.parallel_layers.pp_layers import LayerDesc, SegmentLayers
    descs = [LayerDesc(nn.Embedding, 10, 4)] + [LayerDesc(nn.Linear, 4, 4) for _ in range(8)] + [LayerDesc(nn.LayerNorm, 4)]
    assert SegmentLayers(descs, 4, "uniform").do_segment() == [0, 2, 4, 7, 10]
    assert SegmentLayers(descs, 4, "layer:Linear").do_segment() == [0, 3, 5, 7, 10]
    assert SegmentLayers(descs, 2, "layer:Linear", num_virtual_pipeline_stage=2).do_segment() == [0, 3, 5, 7, 10]


def test_bucket_assignment_matches_reference_algorithm():
    from paddle_b200.distributed.reducer import _assign_py, assign_group_by_size
    ts = [torch.empty(n, dtype=dt) for n, dt in [(300, torch.float32), (200, torch.float32), (100, torch.float16),
                                                  (5000, torch.float32), (10, torch.float32), (10, torch.float16)]]
    got = assign_group_by_size(ts, [False] * 6, [1024, 4096])
    exp = _assign_py([t.numel() * t.element_size() for t in ts], [0, 0, 2, 0, 0, 2], [False] * 6, [1024, 4096])
    assert got == exp == [[0], [1, 3], [2, 5], [4]]


def test_dist_checkpoint_single_process_reshard(tmp_path):
    from paddle_b200.distributed import checkpoint as ckpt
    w = torch.arange(32, dtype=torch.float32).reshape(4, 8)
    ckpt.save_state_dict({"w": w.clone(), "nested": {"b": torch.ones(3)}}, str(tmp_path))
    tgt = torch.zeros(4, 4)
    tgt._dist_meta = ((4, 8), (0, 4))      # right half of the global tensor
    sd = {"w": tgt, "nested": {"b": torch.zeros(3)}}
    ckpt.load_state_dict(sd, str(tmp_path))
    torch.testing.assert_close(tgt, w[:, 4:])
    assert sd["nested"]["b"].tolist() == [1, 1, 1]
    files = os.listdir(tmp_path)
    assert "0.metadata" in files and "0_0.distcp" in files


def test_launch_arg_parsing_and_profiler_scheduler():
    from paddle_b200.distributed.launch.main import parse_args
    a = parse_args(["--nnodes", "2", "--master", "127.0.0.1:8090", "--devices", "0,1", "--max_restart", "5", "train.py", "--lr", "1"])
    assert a.nnodes == "2" and a.devices == "0,1" and a.training_script_args == ["--lr", "1"] and a.max_restart == 5
    from paddle_b200.profiler import ProfilerState, make_scheduler
    s = make_scheduler(closed=1, ready=1, record=2, repeat=1)
    assert [s(i) for i in range(5)] == [ProfilerState.CLOSED, ProfilerState.READY, ProfilerState.RECORD,
                                        ProfilerState.RECORD_AND_RETURN, ProfilerState.CLOSED]


def test_profiler_timer_only_and_record_event():
    prof = paddle.profiler.Profiler(timer_only=True)
    prof.start()
    for _ in range(3):
        with paddle.profiler.RecordEvent("step"):
            paddle.ones([4]).sum()
        prof.step(num_samples=8)
    info = prof.step_info(unit="tokens")
    prof.stop()
    assert "batch_cost" in info and "ips" in info


def test_cpu_op_fallbacks_match_reference_math():
    from paddle_b200 import ops
    x = torch.randn(5, 16)
    w = torch.randn(16)
    torch.testing.assert_close(ops.rms_norm(x, w, 1e-6), x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-6) * w)
    logits, lab = torch.randn(7, 11), torch.randint(0, 11, (7,))
    torch.testing.assert_close(ops.softmax_cross_entropy(logits, lab), torch.nn.functional.cross_entropy(logits, lab, reduction="none"))
    stats = torch.stack([ops.ce_local_stats(logits[:, :5].contiguous(), lab, 0), ops.ce_local_stats(logits[:, 5:].contiguous(), lab, 5)])
    loss, lse = ops.ce_combine(stats, lab)
    torch.testing.assert_close(loss, torch.nn.functional.cross_entropy(logits, lab, reduction="none"), rtol=1e-5, atol=1e-5)
    q = torch.randn(1, 8, 2, 16)
    o, _, _ = ops.fused_rope(q, None, None, use_neox_rotary_style=True)
    assert o.shape == q.shape and torch.allclose(o[:, 0], q[:, 0], atol=1e-6)     # position 0 is the identity


def test_models_train_on_cpu():
    from paddle_b200.models import GPTConfig, GPTForCausalLM, MoEConfig, MoEForCausalLM
    paddle.seed(0)
    for m, V in ((GPTForCausalLM(GPTConfig(vocab_size=128, hidden_size=64, num_attention_heads=4)), 128),
                 (MoEForCausalLM(MoEConfig.tiny()), 512)):
        opt = paddle.optimizer.AdamW(3e-3, parameters=m.parameters())
        ids = paddle.randint(0, V, [2, 17])
        first = last = None
        for _ in range(6):
            loss = m(ids[:, :-1], labels=ids[:, 1:])
            loss.backward(); opt.step(); opt.clear_grad()
            first = first if first is not None else float(loss)
            last = float(loss)
        assert last < first


def test_hapi_model_fit(tmp_path):
    net = nn.Sequential(nn.Flatten(), nn.Linear(16, 4))
    model = paddle.Model(net)
    model.prepare(paddle.optimizer.SGD(0.1, parameters=net.parameters()), nn.CrossEntropyLoss(), paddle.metric.Accuracy())
    ds = paddle.io.TensorDataset([torch.randn(32, 4, 4), torch.randint(0, 4, (32, 1))])
    model.fit(ds, batch_size=8, epochs=1, verbose=0, save_dir=str(tmp_path))
    res = model.evaluate(ds, batch_size=8, verbose=0)
    assert "loss" in res and "acc" in res and os.path.exists(tmp_path / "final.pdparams")


def test_distributed_batch_sampler_partitions():
    ds = list(range(10))
    seen = []
    for r in range(2):
        s = paddle.io.DistributedBatchSampler(ds, batch_size=2, num_replicas=2, rank=r)
        seen.append([i for b in s for i in b])
    assert sorted(seen[0] + seen[1]) == list(range(10)) and len(seen[0]) == len(seen[1]) == 5


def test_auto_parallel_single_process():
    import paddle_b200.distributed as dist
    mesh = dist.ProcessMesh([0], dim_names=["x"])
    t = dist.shard_tensor(paddle.arange(8).reshape([2, 4]), mesh, [dist.Shard(1)])
    assert t.shape == (2, 4) and t.placements == [dist.Shard(1)]
    r = dist.reshard(t, mesh, [dist.Replicate()])
    assert dist.unshard_dtensor(r).tolist() == [[0, 1, 2, 3], [4, 5, 6, 7]]


def test_elastic_manager_state_machine(tmp_path):
    from paddle_b200.distributed.fleet.elastic.manager import ElasticManager, ElasticStatus, FileStore

    class A:
        nnodes = "1:2"; host = "10.0.0.1"; job_id = "t"; elastic_timeout = 1
    store = FileStore(str(tmp_path))
    m = ElasticManager(A(), store)
    assert m.wait() and m.hosts == ["10.0.0.1"]
    store.put("/paddle/t/nodes/10.0.0.2", "10.0.0.2", ttl=10)
    assert m.watch() == ElasticStatus.RESTART
    m.exit()


def test_class_method_surface_of_layer_optimizer_scaler():
    """Methods the reference's Layer / Optimizer / LRScheduler / GradScaler / Tensor expose beyond the common ones."""
    net = paddle.nn.Linear(3, 2)
    seen = []
    handle = net.register_state_dict_hook(lambda sd: seen.append(sorted(sd)) or sd.__setitem__("extra", torch.zeros(1)))
    sd = net.state_dict()
    assert seen == [["bias", "weight"]] and "extra" in sd
    handle.remove()
    assert "extra" not in net.state_dict() and "extra" not in net.state_dict(use_hook=False)
    with pytest.raises(ValueError):
        net.backward()
    opt = paddle.optimizer.Momentum(0.1, parameters=net.parameters(), weight_decay=0.5)
    x = torch.randn(4, 3)
    net(x).sum().backward()
    pairs = opt.backward(None)
    assert [p is q for (p, _), q in zip(pairs, net.parameters())] == [True, True]
    assert len(opt.backward(None, parameters=[net.bias])) == 1
    reg = opt.append_regularization_ops(pairs)
    torch.testing.assert_close(reg[0][1], pairs[0][1] + 0.5 * net.weight.detach())
    l1 = opt.append_regularization_ops(pairs, paddle.regularizer.L1Decay(0.1))
    torch.testing.assert_close(l1[1][1], pairs[1][1] + 0.1 * torch.sign(net.bias.detach()))
    net.weight.regularizer = paddle.regularizer.L2Decay(2.0)             # the parameter's own regulariser wins
    torch.testing.assert_close(opt.append_regularization_ops(pairs)[0][1], pairs[0][1] + 2.0 * net.weight.detach())
    net.weight.regularizer = None
    before = net.weight.detach().clone()
    _, returned = opt.minimize(None)
    assert len(returned) == 2 and not torch.equal(before, net.weight.detach())
    assert any("velocity" in n for n in opt.get_opti_var_name_list())
    plain = paddle.optimizer.SGD(0.5, parameters=net.parameters())
    w0 = net.weight.detach().clone()
    plain.apply_gradients([(net.weight, torch.ones_like(net.weight)), (net.bias, None)])
    torch.testing.assert_close(net.weight.detach(), w0 - 0.5)
    sched = paddle.optimizer.lr.StepDecay(0.1, 2)
    sched.state_keys()
    assert sched.keys == ["last_epoch", "last_lr"]
    scaler = paddle.amp.GradScaler(init_loss_scaling=8.0)
    scaler.set_incr_ratio(3.0)
    scaler.set_decr_ratio(0.25)
    scaler.set_incr_every_n_steps(7)
    scaler.set_decr_every_n_nan_or_inf(2)
    assert (scaler.get_incr_ratio(), scaler.get_decr_ratio(), scaler.get_incr_every_n_steps(), scaler.get_decr_every_n_nan_or_inf()) == (3.0, 0.25, 7, 2)
    assert scaler.state_dict()["incr_ratio"] == 3.0
    with pytest.raises(AssertionError):
        scaler.set_incr_ratio(0.5)
    t = torch.tensor([1.0, -2.0])
    assert t.apply(lambda v: v * 2).tolist() == [2.0, -4.0] and t.tolist() == [1.0, -2.0]
    assert t.apply_(lambda v: v.abs()) is t and t.tolist() == [1.0, 2.0]
    v0 = t.inplace_version
    t.add_(1)
    assert t.inplace_version == v0 + 1
    with pytest.raises(RuntimeError):
        torch.ones(2, requires_grad=True).apply(lambda v: v)


def test_auto_cast_custom_lists():
    x = torch.randn(4, 8)
    lin = paddle.nn.Linear(8, 8)
    with paddle.amp.auto_cast(dtype="bfloat16"):
        y = lin(x)
        assert y.dtype == torch.bfloat16
        assert paddle.amp.amp_state()["enable"]
    with paddle.amp.auto_cast(dtype="bfloat16", custom_black_list=["matmul"], custom_white_list=["exp", "softmax"]):
        y = lin(x)
        assert y.dtype == torch.float32 and torch.matmul(x, x.t()).dtype == torch.float32          # black: stays float32
        assert torch.exp(y).dtype == torch.bfloat16 and torch.softmax(y, -1).dtype == torch.bfloat16  # white: low precision
        assert (x + x).dtype == torch.float32                                                       # unlisted: untouched
    with paddle.amp.auto_cast(dtype="bfloat16", custom_black_list=["elementwise_add"]):
        b = lin(x)
        assert (b + b).dtype == torch.float32
    with paddle.amp.auto_cast(enable=False, custom_black_list=["matmul"]):
        assert lin(x).dtype == torch.float32 and not paddle.amp.amp_state()["enable"]
    assert lin(x).dtype == torch.float32 and paddle.amp.amp_state()["black"] == set()
    with pytest.raises(ValueError):
        with paddle.amp.auto_cast(custom_white_list=["exp"], custom_black_list=["exp"]):
            pass
    with pytest.raises(ValueError):
        with paddle.amp.auto_cast(level="O3"):
            pass
    with pytest.raises(ValueError):
        with paddle.amp.auto_cast(dtype="float32"):
            pass
