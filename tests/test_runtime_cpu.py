"""Training runtime on CPU (gloo): config resolution, LR schedules, loss scaling, optimizers, ZeRO 0/1/2/3 against a
single-process fp32 reference, checkpoint resume, pipeline parallelism. Mirrors the reference's
tests/unit/runtime/{test_ds_config_dict,test_lr_schedulers,zero/test_zero,half_precision,pipe}."""
import math

import pytest
import torch

from tests.dist_utils import run_distributed
from tests.runtime_helper import pipe_reference, pipe_worker, reference_run, zero_worker

pytestmark = pytest.mark.timeout(600)


def test_batch_size_resolution():
    from hcache_deepspeed_b200.runtime.config import DeepSpeedConfig, DeepSpeedConfigError
    c = DeepSpeedConfig({"train_batch_size": 32, "gradient_accumulation_steps": 2}, world_size=4)
    assert (c.train_batch_size, c.train_micro_batch_size_per_gpu, c.gradient_accumulation_steps) == (32, 4, 2)
    c = DeepSpeedConfig({"train_micro_batch_size_per_gpu": 3}, world_size=2)
    assert (c.train_batch_size, c.gradient_accumulation_steps) == (6, 1)
    c = DeepSpeedConfig({"train_batch_size": 16, "train_micro_batch_size_per_gpu": 2}, world_size=2)
    assert c.gradient_accumulation_steps == 4
    with pytest.raises(DeepSpeedConfigError):
        DeepSpeedConfig({"train_batch_size": 10, "train_micro_batch_size_per_gpu": 3, "gradient_accumulation_steps": 1}, world_size=2)
    with pytest.raises(DeepSpeedConfigError):
        DeepSpeedConfig({"steps_per_print": 1}, world_size=1)
    with pytest.raises(DeepSpeedConfigError):
        DeepSpeedConfig({"train_batch_size": 4, "fp16": {"enabled": True}, "bf16": {"enabled": True}}, world_size=1)
    c = DeepSpeedConfig({"train_batch_size": 4, "zero_optimization": {"stage": 3, "stage3_max_live_parameters": 5, "cpu_offload": True}},
                        world_size=1)
    assert c.zero_config.max_live_parameters == 5 and c.zero_config.overlap_comm and str(c.zero_config.offload_optimizer.device).endswith("cpu")


def test_lr_schedules():
    from hcache_deepspeed_b200.runtime.lr_schedules import LRRangeTest, OneCycle, WarmupCosineLR, WarmupDecayLR, WarmupLR
    p = [torch.nn.Parameter(torch.zeros(1))]
    opt = torch.optim.SGD(p, lr=0.1)
    s = WarmupLR(opt, warmup_min_lr=0.0, warmup_max_lr=0.1, warmup_num_steps=10, warmup_type="linear")
    lrs = []
    for _ in range(15):
        s.step()
        lrs.append(opt.param_groups[0]["lr"])
    assert lrs[0] == 0.0 and abs(lrs[5] - 0.05) < 1e-9 and lrs[10] == pytest.approx(0.1) and lrs[14] == pytest.approx(0.1)
    opt = torch.optim.SGD(p, lr=0.1)
    s = WarmupDecayLR(opt, total_num_steps=20, warmup_max_lr=0.1, warmup_num_steps=10, warmup_type="log")
    lrs = []
    for _ in range(21):
        s.step()
        lrs.append(opt.param_groups[0]["lr"])
    assert lrs[0] == 0.0 and lrs[9] == pytest.approx(0.1 * math.log(10) / math.log(10)) and lrs[15] == pytest.approx(0.05) and lrs[20] == 0.0
    opt = torch.optim.SGD(p, lr=0.1)
    s = LRRangeTest(opt, lr_range_test_min_lr=0.01, lr_range_test_step_size=5, lr_range_test_step_rate=1.0, lr_range_test_staircase=True)
    lrs = []
    for _ in range(11):
        s.step()
        lrs.append(opt.param_groups[0]["lr"])
    assert lrs[0] == pytest.approx(0.01) and lrs[4] == pytest.approx(0.02) and lrs[9] == pytest.approx(0.03)
    opt = torch.optim.SGD(p, lr=0.1, momentum=0.9)
    s = OneCycle(opt, cycle_min_lr=0.01, cycle_max_lr=0.1, cycle_first_step_size=10, decay_step_size=5, decay_lr_rate=0.5)
    lrs = []
    for _ in range(30):
        s.step()
        lrs.append(opt.param_groups[0]["lr"])
    assert max(lrs) == pytest.approx(0.1, rel=1e-6) and lrs.index(max(lrs)) == 9 and lrs[19] == pytest.approx(0.01, abs=1e-9)
    assert lrs[29] < 0.01
    opt = torch.optim.SGD(p, lr=0.2)
    s = WarmupCosineLR(opt, total_num_steps=20, warmup_num_steps=5, warmup_type="linear", cos_min_ratio=0.1)
    lrs = []
    for _ in range(21):
        s.step()
        lrs.append(opt.param_groups[0]["lr"])
    assert lrs[0] == 0.0 and max(lrs) <= 0.2 + 1e-9 and lrs[20] == pytest.approx(0.02, rel=1e-6)
    sd = s.state_dict()
    s.load_state_dict(sd)
    assert s.last_batch_iteration == 20


def test_dynamic_loss_scaler():
    from hcache_deepspeed_b200.runtime.fp16.loss_scaler import DynamicLossScaler
    s = DynamicLossScaler(init_scale=2 ** 8, scale_window=2, delayed_shift=2, min_scale=1)
    s.update_scale(True)   # first overflow only burns a hysteresis credit
    assert s.cur_scale == 256
    s.update_scale(True)
    assert s.cur_scale == 128
    s.update_scale(False)
    s.update_scale(False)
    assert s.cur_scale == 256
    s = DynamicLossScaler(init_scale=2, min_scale=1, delayed_shift=1)
    s.update_scale(True)
    with pytest.raises(Exception):
        s.update_scale(True)


def test_cpu_optimizers_match_torch():
    from hcache_deepspeed_b200.ops.adam import DeepSpeedCPUAdam, FusedAdam
    from hcache_deepspeed_b200.ops.adagrad import DeepSpeedCPUAdagrad
    from hcache_deepspeed_b200.ops.lamb import FusedLamb
    from hcache_deepspeed_b200.ops.lion import DeepSpeedCPULion, FusedLion
    torch.manual_seed(0)

    def run(a_cls, b_cls, akw, bkw, n=70001, steps=4):
        w0 = torch.randn(n)
        a, b = torch.nn.Parameter(w0.clone()), torch.nn.Parameter(w0.clone())
        oa, ob = a_cls([a], **akw), b_cls([b], **bkw)
        for _ in range(steps):
            g = torch.randn(n)
            a.grad, b.grad = g.clone(), g.clone()
            oa.step()
            ob.step()
        return float((a - b).abs().max())

    assert run(DeepSpeedCPUAdam, torch.optim.AdamW, dict(lr=1e-2, weight_decay=0.1), dict(lr=1e-2, weight_decay=0.1)) < 1e-5
    assert run(DeepSpeedCPUAdam, torch.optim.Adam, dict(lr=1e-2, weight_decay=0.1, adamw_mode=False), dict(lr=1e-2, weight_decay=0.1)) < 1e-5
    assert run(FusedAdam, torch.optim.AdamW, dict(lr=1e-2, weight_decay=0.1), dict(lr=1e-2, weight_decay=0.1)) < 1e-5
    assert run(DeepSpeedCPUAdagrad, torch.optim.Adagrad, dict(lr=1e-2), dict(lr=1e-2, eps=1e-10)) < 1e-6
    assert run(DeepSpeedCPULion, FusedLion, dict(lr=1e-3, weight_decay=0.1), dict(lr=1e-3, weight_decay=0.1)) < 1e-5
    p = torch.nn.Parameter(torch.randn(1000))
    sh = torch.zeros(1000, dtype=torch.bfloat16)
    o = DeepSpeedCPUAdam([p], lr=1e-2)
    p.grad = torch.randn(1000)
    o.step(fp16_param_groups=[[sh]])
    assert torch.equal(sh, p.data.to(torch.bfloat16))
    w = torch.nn.Parameter(torch.randn(50, 20))
    o = FusedLamb([w], lr=1e-2, weight_decay=0.01)
    w.grad = torch.randn(50, 20)
    before = w.detach().clone()
    o.step()
    assert 0.01 <= o.get_lamb_coeffs()[0] <= 10 and not torch.equal(before, w.detach())


def test_partition_helpers():
    from hcache_deepspeed_b200.runtime.utils import partition_balanced, partition_uniform
    assert partition_uniform(10, 3) == [0, 4, 7, 10]
    assert partition_uniform(2, 4) == [0, 1, 2, 2, 2]
    parts = partition_balanced([1, 1, 1, 1, 10, 1, 1, 1], 3)
    w = [1, 1, 1, 1, 10, 1, 1, 1]
    assert parts[0] == 0 and parts[-1] == 8 and max(sum(w[parts[i]:parts[i + 1]]) for i in range(3)) == 10


def test_zero_init_and_gathered_parameters():
    from hcache_deepspeed_b200.runtime.zero import GatheredParameters, Init, ZeroParamStatus
    with Init():
        m = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))
    p = m[0].weight
    assert p.ds_status == ZeroParamStatus.NOT_AVAILABLE and p.numel() == 0 and tuple(p.ds_shape) == (16, 8)
    assert float(p.ds_tensor.abs().sum()) > 0  # initialisers ran on the full tensor before partitioning
    with GatheredParameters(m.parameters(), modifier_rank=0):
        assert tuple(m[0].weight.shape) == (16, 8)
        m[0].weight.data.fill_(1.0)
    assert p.ds_status == ZeroParamStatus.NOT_AVAILABLE and float(p.ds_tensor.min()) == 1.0
    assert tuple(torch.nn.Linear(2, 2).weight.shape) == (2, 2)  # constructors restored outside the context


def test_activation_checkpointing_matches_plain():
    from hcache_deepspeed_b200.runtime.activation_checkpointing import checkpointing
    checkpointing.configure(None, checkpoint_in_cpu=True)
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(8, 8), torch.nn.Dropout(0.5), torch.nn.Tanh(), torch.nn.Linear(8, 8))
    x = torch.randn(4, 8, requires_grad=True)
    torch.manual_seed(5)
    y = net(x).sum()
    y.backward()
    g_plain, gx_plain = [p.grad.clone() for p in net.parameters()], x.grad.clone()
    for p in net.parameters():
        p.grad = None
    x.grad = None
    torch.manual_seed(5)
    y2 = checkpointing.checkpoint(net, x).sum()
    y2.backward()
    assert torch.allclose(y, y2)
    for a, p in zip(g_plain, net.parameters()):
        assert torch.allclose(a, p.grad, atol=1e-6)
    assert torch.allclose(gx_plain, x.grad, atol=1e-6)
    checkpointing.configure(None)


def test_engine_single_process_fp32_matches_torch():
    import hcache_deepspeed_b200 as ds
    from tests.runtime_helper import TinyNet, make_batches
    want, _ = reference_run(steps=5, global_bs=8, clip=0.5)
    torch.manual_seed(1234)
    engine, _, _, sched = ds.initialize(model=TinyNet(), config={
        "train_batch_size": 8, "gradient_accumulation_steps": 2, "gradient_clipping": 0.5,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.0}},
        "scheduler": {"type": "WarmupLR", "params": {"warmup_min_lr": 1e-2, "warmup_max_lr": 1e-2, "warmup_num_steps": 2}}})
    got = []
    for x, y in make_batches(5, 8):
        tot = 0.0
        for a in range(2):
            loss = engine(x[a * 4:(a + 1) * 4], y[a * 4:(a + 1) * 4])
            engine.backward(loss)
            engine.step()
            tot += float(loss) / 2
        got.append(tot)
    assert engine.global_steps == 5 and sched is not None and engine.get_global_grad_norm() > 0
    assert got == pytest.approx(want, rel=2e-4, abs=2e-5)


@pytest.mark.parametrize("stage,gas,offload", [(0, 1, False), (1, 2, False), (2, 2, False), (2, 1, True), (3, 2, False), (3, 1, True)])
def test_zero_stages_match_reference_2ranks(stage, gas, offload, tmp_path):
    steps, world = 6, 2
    want, want_state = reference_run(steps=steps, global_bs=4 * gas * world)
    ckpt = str(tmp_path / "ckpt") if stage in (1, 2, 3) and not offload else None
    res = run_distributed(zero_worker, world, stage, steps, gas, offload, "fp32", ckpt)
    for r in res:
        assert r["global_steps"] == steps
        for k, v in want_state.items():
            torch.testing.assert_close(r["state"][k], v, rtol=2e-4, atol=2e-5)
        if ckpt:
            # the resumed run's losses equal the uninterrupted run's second half (per rank, local micro-batches)
            assert r["resumed_losses"] == pytest.approx(r["losses"][steps // 2:], rel=1e-4, abs=1e-5)
    mean = [sum(r["losses"][i] for r in res) / world for i in range(steps)]
    assert mean == pytest.approx(want, rel=5e-4, abs=5e-5)


def test_zero3_leaf_module_2ranks(monkeypatch):
    """A module flagged with set_z3_leaf_modules gathers its whole subtree from one hook; results are unchanged."""
    monkeypatch.setenv("HDS_TEST_Z3_LEAF", "1")
    steps, world = 4, 2
    want, want_state = reference_run(steps=steps, global_bs=4 * world)
    res = run_distributed(zero_worker, world, 3, steps, 1, False, "fp32", None)
    for r in res:
        for k, v in want_state.items():
            torch.testing.assert_close(r["state"][k], v, rtol=2e-4, atol=2e-5)


@pytest.mark.parametrize("distance", [0, 100, 10 ** 9])
def test_zero3_max_reuse_distance_2ranks(monkeypatch, distance):
    """stage3_max_reuse_distance: a parameter the recorded order needs again within that many fetched elements is not
    released in between (ref: partitioned_param_coordinator.py release_sub_module). Results never change."""
    monkeypatch.setenv("HDS_TEST_Z3_REUSE", str(distance))
    steps, world = 4, 2
    want, want_state = reference_run(steps=steps, global_bs=4 * world)
    res = run_distributed(zero_worker, world, 3, steps, 1, False, "fp32", None)
    kept = [r["offload_stats"]["kept_for_reuse"] for r in res]
    for r in res:
        for k, v in want_state.items():
            torch.testing.assert_close(r["state"][k], v, rtol=2e-4, atol=2e-5)
    if distance == 0:
        assert kept == [0, 0]
    else:
        assert all(k > 0 for k in kept)                 # the head's weights at least survive from forward into backward
    test_zero3_max_reuse_distance_2ranks.kept = getattr(test_zero3_max_reuse_distance_2ranks, "kept", {})
    test_zero3_max_reuse_distance_2ranks.kept[distance] = kept[0]
    seen = test_zero3_max_reuse_distance_2ranks.kept
    if 100 in seen and 10 ** 9 in seen:
        assert seen[100] < seen[10 ** 9]                # a short distance keeps fewer parameters than an unbounded one


def test_zero3_nvme_optimizer_swapper_2ranks(tmp_path):
    """ZeRO-Infinity: fp32 masters + Adam moments live in files between steps (pipelined swapper); same training
    result, checkpoint resume included."""
    steps, world = 6, 2
    want, want_state = reference_run(steps=steps, global_bs=4 * world)
    res = run_distributed(zero_worker, world, 3, steps, 1, "nvme", "fp32", str(tmp_path / "ckpt"))
    for r in res:
        for k, v in want_state.items():
            torch.testing.assert_close(r["state"][k], v, rtol=2e-4, atol=2e-5)
        assert r["resumed_losses"] == pytest.approx(r["losses"][steps // 2:], rel=1e-4, abs=1e-5)
        assert r["swap_stats"]["bytes_read"] > 0 and r["swap_stats"]["bytes_written"] > 0
        assert all(n == 0 for n in r["masters_resident"])      # nothing resident between steps


@pytest.mark.parametrize("stage", [3, 1])
def test_schedule_compiler_2ranks(monkeypatch, stage):
    """engine.compile(): after the profiled first step the passes apply a prefetch plan, persistent parameters and
    optimizer-state offload; training results do not change."""
    monkeypatch.setenv("HDS_TEST_Z3_COMPILE", "1")
    steps, world = 5, 2
    want, want_state = reference_run(steps=steps, global_bs=4 * world)
    res = run_distributed(zero_worker, world, stage, steps, 1, False, "fp32", None)
    for r in res:
        for k, v in want_state.items():
            torch.testing.assert_close(r["state"][k], v, rtol=2e-4, atol=2e-5)
        rep = r["compile_report"]
        assert rep is not None and rep["adam_state_bytes"] > 0
        if stage == 3:
            assert rep["events"] > 0 and rep["prefetch_launch_points"] > 0
            assert rep["persistent_params"] > 0 and rep["persistent_bytes"] <= 2 * 3000
            assert r["persistent"] == rep["persistent_params"]
            assert r["prefetch_stats"]["prefetched"] > 0 and r["prefetch_stats"]["trace_invalidations"] == 0


@pytest.mark.parametrize("stage,dtype", [(0, "bf16"), (2, "bf16"), (2, "fp16"), (0, "fp16")])
def test_mixed_precision_trains(stage, dtype):
    res = run_distributed(zero_worker, 2, stage, 12, 1, False, dtype, None)
    for r in res:
        assert all(math.isfinite(l) for l in r["losses"])
    mean = [sum(r["losses"][i] for r in res) / 2 for i in range(12)]
    assert sum(mean[-3:]) < sum(mean[:3])  # it learns
    for k in res[0]["state"]:
        torch.testing.assert_close(res[0]["state"][k], res[1]["state"][k], rtol=0, atol=0)  # replicas stay in sync


@pytest.mark.parametrize("world,stages", [(2, 2), (4, 2)])
def test_pipeline_matches_sequential(world, stages):
    steps, mbs = 4, 4
    want = pipe_reference(steps, mbs, world // stages)
    res = run_distributed(pipe_worker, world, stages, steps, mbs)
    for r in res:
        assert r["losses"] == pytest.approx(want, rel=1e-3, abs=1e-4)
        assert r["parts"][0] == 0 and r["parts"][-1] == 5 and math.isfinite(r["eval"])
    assert {r["stage"] for r in res} == set(range(stages))


def test_pipe_schedules_pair_up_under_blocking_p2p():
    """Every send of the 1F1B / inference schedules meets its receive when point-to-point calls block until matched
    (rendezvous semantics), for every stage count and micro-batch count; no buffer is overwritten by a forward before its
    backward ran; each stage runs every micro-batch forward and backward exactly once."""
    from hcache_deepspeed_b200.runtime.pipe import schedule as S
    match = {"SendActivation": "RecvActivation", "RecvActivation": "SendActivation", "SendGrad": "RecvGrad", "RecvGrad": "SendGrad"}

    def completes(cls, M, st):
        prog = [[c for step in cls(M, st, i).steps() for c in step] for i in range(st)]
        pc, moved = [0] * st, True
        while moved:
            moved = False
            for i in range(st):
                while pc[i] < len(prog[i]):
                    n = prog[i][pc[i]].name
                    if n in match:
                        peer = i + 1 if n in ("SendActivation", "RecvGrad") else i - 1
                        if pc[peer] < len(prog[peer]) and prog[peer][pc[peer]].name == match[n]:
                            pc[i] += 1
                            pc[peer] += 1
                            moved = True
                            continue
                        break
                    pc[i] += 1
                    moved = True
        return all(pc[i] == len(prog[i]) for i in range(st))

    for st in range(1, 8):
        for M in range(1, 12):
            assert completes(S.TrainSchedule, M, st) and completes(S.InferenceSchedule, M, st), (M, st)
            for i in range(st):
                sched = S.TrainSchedule(M, st, i)
                cmds = [c for step in sched.steps() for c in step]
                assert sum(c.name == "ForwardPass" for c in cmds) == M == sum(c.name == "BackwardPass" for c in cmds)
                assert cmds[-1].name == "OptimizerStep"
                busy = set()
                for c in cmds:
                    if c.name == "ForwardPass":
                        assert c.buffer_id not in busy and c.buffer_id < sched.num_pipe_buffers()
                        busy.add(c.buffer_id)
                    elif c.name == "BackwardPass":
                        busy.discard(c.buffer_id)


def test_zero3_frozen_params_checkpoint_and_init_broadcast_2ranks(tmp_path):
    from tests.runtime_helper import zero3_frozen_worker
    res = run_distributed(zero3_frozen_worker, 2, str(tmp_path / "ckpt"))
    for r in res:
        assert r["init_matches_rank0"], "zero.Init kept rank-local initial values instead of rank 0's"
        assert r["frozen_unchanged"]
        for key in ("full_load", "module_only_load"):
            for k, v in r["trained"].items():
                torch.testing.assert_close(r[key][k], v, rtol=0, atol=0, msg=f"{key}: {k}")
    shards = res[0]["fp32_from_shards"]
    for k, v in res[0]["trained"].items():
        torch.testing.assert_close(shards[k], v, rtol=1e-6, atol=1e-7)


class _ChatLoRALinear(torch.nn.Module):
    """DeepSpeed-Chat's LinearLayer_LoRA layout: frozen base weight + right [in, r] / left [r, out] adapters."""

    def __init__(self, base: torch.nn.Linear, r=4, scaling=2.0):
        super().__init__()
        self.weight = base.weight
        self.weight.requires_grad = False
        self.bias = base.bias
        self.lora_right_weight = torch.nn.Parameter(torch.randn(base.in_features, r) * 0.05)
        self.lora_left_weight = torch.nn.Parameter(torch.randn(r, base.out_features) * 0.05)
        self.lora_scaling = scaling / r

    def forward(self, x):
        return torch.nn.functional.linear(x, self.weight, self.bias) + (x @ self.lora_right_weight @ self.lora_left_weight) * self.lora_scaling


@pytest.mark.parametrize("stage", [0, 3])
def test_hybrid_engine_lora_generate_tracks_training(stage):
    """RLHF loop on one rank: generate -> train the LoRA adapters -> generate again. The injected inference copy must see
    the fused adapters of the CURRENT step (ref: runtime/hybrid_engine.py:132-168), the training weights stay un-fused,
    ZeRO-3 parameters are gathered for the refresh only."""
    transformers = pytest.importorskip("transformers")
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.hybrid_engine import DeepSpeedHybridEngine
    torch.manual_seed(0)
    cfg = transformers.LlamaConfig(hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
                                   num_key_value_heads=2, vocab_size=97, max_position_embeddings=256)
    model = transformers.LlamaForCausalLM(cfg).float()
    for layer in model.model.layers:
        layer.self_attn.q_proj = _ChatLoRALinear(layer.self_attn.q_proj)
        layer.mlp.down_proj = _ChatLoRALinear(layer.mlp.down_proj)
    ds_cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 5e-2}},
              "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0},
              "hybrid_engine": {"enabled": True, "max_out_tokens": 32}, "steps_per_print": 1000}
    engine, _, _, _ = ds.initialize(model=model, config=ds_cfg, model_parameters=[p for p in model.parameters() if p.requires_grad])
    assert isinstance(engine, DeepSpeedHybridEngine) and len(engine._lora_layers) == 4
    prompt = torch.randint(1, 97, (2, 7))

    def reference_generate():
        """Greedy decoding with the training module itself (its forward applies the adapters un-fused)."""
        ctx = engine._gather_context()
        with ctx, torch.no_grad():
            ids = prompt.clone()
            for _ in range(6):
                nxt = engine.module(ids).logits[:, -1].argmax(-1, keepdim=True)
                ids = torch.cat([ids, nxt], dim=1)
        return ids

    engine.eval()
    out0 = engine.generate(prompt, max_new_tokens=6, do_sample=False)
    assert engine._inference_engine.injected and torch.equal(out0[:, :prompt.shape[1]], prompt)
    assert torch.equal(out0, reference_generate())
    first_engine = engine._inference_engine
    engine.generate(prompt, max_new_tokens=2, do_sample=False)
    assert engine._inference_engine is first_engine                 # nothing moved: no refresh
    base_before = [m.weight.detach().clone() for m, _ in engine._lora_layers] if stage == 0 else None

    engine.train()
    for _ in range(3):
        ids = torch.randint(1, 97, (2, 12))
        loss = engine(input_ids=ids, labels=ids).loss
        engine.backward(loss)
        engine.step()
    engine.eval()
    out1 = engine.generate(prompt, max_new_tokens=6, do_sample=False)
    assert engine._inference_engine is not first_engine             # the step counter moved: weights refreshed
    assert torch.equal(out1, reference_generate())
    assert not engine.is_lora_fused
    if stage == 0:
        for (m, _), w in zip(engine._lora_layers, base_before):
            torch.testing.assert_close(m.weight.detach(), w, rtol=0, atol=1e-6)       # base weights came back un-fused
    stats = engine.latency_stats()
    assert stats["generate_s"] > 0 and stats["iterations"] >= 1
    engine.release_inference_engine()
    assert engine._inference_engine is None
    engine.retake_inference_cache()
    assert engine._inference_engine is not None


def test_zero_init_around_huggingface_constructor_2ranks():
    pytest.importorskip("transformers")
    from tests.runtime_helper import zero_init_hf_worker
    for r in run_distributed(zero_init_hf_worker, 2):
        assert r["n_zero"] > 10 and r["worst"] == 0.0            # every rank holds rank 0's HF initialisation
        assert 0.015 < r["wte_std"] < 0.025                      # initializer_range 0.02, not nn.Embedding's N(0, 1)
        assert r["pretrained_worst"] == 0.0 and r["pretrained_logits"] < 1e-5


def test_amp_block_runs_the_forward_under_bf16_autocast():
    """``"amp": {"enabled": true}``: fp32 parameters, forward under torch.autocast(bfloat16) (apex-free counterpart of the
    reference's amp path); training still converges; O0 keeps fp32; fp16 / ZeRO combinations are refused."""
    import hcache_deepspeed_b200 as ds
    seen = {}

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.b = torch.nn.Linear(8, 16), torch.nn.Linear(16, 1)

        def forward(self, x, y):
            h = self.a(x)
            seen["dtype"] = h.dtype
            return (self.b(torch.relu(h)).float() - y).pow(2).mean()

    base = {"train_micro_batch_size_per_gpu": 8, "optimizer": {"type": "Adam", "params": {"lr": 0.05}}, "steps_per_print": 1000}
    torch.manual_seed(0)
    x, y = torch.randn(8, 8), torch.randn(8, 1)
    engine, _, _, _ = ds.initialize(model=Net(), config=dict(base, amp={"enabled": True, "opt_level": "O1"}))
    losses = []
    for _ in range(30):
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
        losses.append(float(loss.detach()))
    assert seen["dtype"] == torch.bfloat16 and next(engine.module.parameters()).dtype == torch.float32
    assert losses[-1] < 0.5 * losses[0]
    engine, _, _, _ = ds.initialize(model=Net(), config=dict(base, amp={"enabled": True, "opt_level": "O0"}))
    engine(x, y)
    assert seen["dtype"] == torch.float32
    with pytest.raises(Exception):
        ds.initialize(model=Net(), config=dict(base, amp={"enabled": True}, bf16={"enabled": True}))
    with pytest.raises(NotImplementedError):
        ds.initialize(model=Net(), config=dict(base, amp={"enabled": True, "dtype": "fp16"}))


def test_sparse_gradients_allgather_rows_2ranks():
    from tests.runtime_helper import sparse_gradients_worker
    res = run_distributed(sparse_gradients_worker, 2, False)       # dense autograd gradient, exchanged row-sparse
    for r in res:
        assert 0 < r["stats"]["rows_sent"] <= 24 < r["stats"]["dense_rows"]
        torch.testing.assert_close(r["w"], r["w_dense"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(res[0]["w"], res[1]["w"], rtol=0, atol=0)
    res2 = run_distributed(sparse_gradients_worker, 2, True)        # nn.Embedding(sparse=True): the gradient stays sparse
    torch.testing.assert_close(res2[0]["w"], res2[1]["w"], rtol=0, atol=0)
    torch.testing.assert_close(res2[0]["w"], res[0]["w"], rtol=1e-5, atol=1e-6)


def test_moe_checkpoint_per_global_expert_reloads_under_other_ep_degree(tmp_path):
    from tests.runtime_helper import moe_checkpoint_worker
    d = str(tmp_path / "ck")
    saved = run_distributed(moe_checkpoint_worker, 2, d, "save")
    files = saved[0]["files"]
    assert [f for f in files if f.startswith("layer_0_expert_")] == [f"layer_0_expert_{e}_mp_rank_00_model_states.pt" for e in range(4)]
    assert "expp_rank_0_mp_rank_00_optim_states.pt" in files and "expp_rank_1_mp_rank_00_optim_states.pt" in files
    for r in saved:
        assert r["losses"][-1] < r["losses"][0] and r["opt_entries"] > 0
        torch.testing.assert_close(r["again"], r["pred"], rtol=0, atol=0)        # same-degree reload is exact
    one = run_distributed(moe_checkpoint_worker, 1, d, "load")[0]               # ep 2 -> ep 1: all four experts on one rank
    for r in saved:
        for gid, tensors in r["experts"].items():
            for a, b in zip(one["experts"][gid], tensors):
                torch.testing.assert_close(a, b, rtol=0, atol=0)
    # token capacity is ample, so routing is the same and the predictions agree with the 2-rank run
    torch.testing.assert_close(one["pred"], saved[0]["pred"], rtol=1e-5, atol=1e-6)


def test_pipeline_layer_checkpoints_reload_under_other_stage_count(tmp_path):
    from tests.runtime_helper import pipe_checkpoint_worker
    d = str(tmp_path / "ck")
    saved = run_distributed(pipe_checkpoint_worker, 2, d, "save")
    layer_files = [f for f in saved[0]["files"] if f.startswith("layer_")]
    assert layer_files == [f"layer_{i:02d}-model_00-model_states.pt" for i in range(6)]
    for r in saved:
        assert r["losses"][-1] < 0.6 * r["losses"][0]
        assert r["emb_moved"] > 1e-3                       # the checkpointed first stage still trains its embedding
    assert saved[0]["emb_sum"] == pytest.approx(saved[1]["emb_sum"], rel=1e-6)      # tied weight identical on both stages
    one = run_distributed(pipe_checkpoint_worker, 1, d, "load")[0]                  # 2 stages -> 1 stage
    two = run_distributed(pipe_checkpoint_worker, 2, d, "load")
    assert one["eval"] == pytest.approx(saved[0]["eval"], rel=1e-5) and two[0]["eval"] == pytest.approx(saved[0]["eval"], rel=1e-5)
    assert one["emb_sum"] == pytest.approx(saved[0]["emb_sum"], rel=1e-6)


def test_zero3_parameter_partitions_on_nvme_2ranks(tmp_path):
    from tests.runtime_helper import zero3_nvme_param_worker
    host = run_distributed(zero3_nvme_param_worker, 2, "cpu", str(tmp_path))
    nvme = run_distributed(zero3_nvme_param_worker, 2, "nvme", str(tmp_path))
    for h, n in zip(host, nvme):
        assert n["losses"] == h["losses"] and n["losses"][-1] < n["losses"][0]        # bit-identical to host offload
        assert h["nvme_params"] == [] and n["nvme_params"] == ["emb.weight", "out.weight"]
        assert len(n["files"]) == 2 and all(f.endswith("_param.tensor.swp") for f in n["files"])
        assert n["free_buffers"] == 4 and n["host_slices"] == [1, 1]                  # nothing staged between steps
        assert n["after_reload"] < n["losses"][0]


def test_zeropp_quantized_weights_and_gradients_2ranks():
    from tests.runtime_helper import zeropp_worker
    base = run_distributed(zeropp_worker, 2, False, False)
    qpp = run_distributed(zeropp_worker, 2, True, True)
    for b, q in zip(base, qpp):
        assert b["flagged"] == 0 and b["qgz"]["calls"] == 0
        assert q["flagged"] == 6 and q["qgz"]["calls"] > 0
        assert q["qgz"]["bytes_sent"] < 0.3 * q["qgz"]["bytes_unquantized"]            # int8 + scales vs fp32
        assert q["losses"][-1] < 0.2 * q["losses"][0]                                   # still trains
        assert abs(q["losses"][-1] - b["losses"][-1]) < 0.1 * b["losses"][0]            # and tracks the exact run


def test_mics_shard_groups_4ranks(monkeypatch):
    """MiCS (ref: runtime/zero/mics.py): ZeRO-3 partitions inside shard groups of 2, gradients all-reduced across the two
    replica groups; same training result as plain data parallelism."""
    monkeypatch.setenv("HDS_TEST_MICS", "2")
    steps, world = 4, 4
    want, want_state = reference_run(steps=steps, global_bs=4 * world)
    res = run_distributed(zero_worker, world, 3, steps, 1, False, "fp32", None)
    for r in res:
        assert r["mics"][:3] == ("MiCS_Optimizer", 2, 2) and 0 <= r["mics"][3] < 2      # sliced over 2 ranks, not 4
        for k, v in want_state.items():
            torch.testing.assert_close(r["state"][k], v, rtol=2e-4, atol=2e-5)


def test_partition_activations_across_model_parallel_ranks():
    from tests.runtime_helper import partition_activations_worker
    for r in run_distributed(partition_activations_worker, 2):
        for k, v in r.items():
            assert v["err"] < 1e-6 and v["loss"] < 1e-6, (k, v)
            assert v["saved"] and all(part * 2 == full for full, part in v["saved"]), (k, v)      # half of each input kept per rank


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_safe_accessors_of_partitioned_tensors_2ranks(stage):
    from tests.runtime_helper import safe_accessors_worker
    for r in run_distributed(safe_accessors_worker, 2, stage):
        assert r["grad_err"] < 1e-6 and r["frozen_by_zero_grad"] == 0.0, r
        assert r["other_param_err"] < 1e-6 and r["exp_avg_err"] < 1e-7, r
        assert r["set_param"] == 0.0 and r["set_state"] == 0.0 and r["lp_follows"] == 0.0, r
        if stage == 3:
            assert r["local_numel"][0] == r["local_numel"][1] and r["local_set"] == (1.0, 2.0) and r["local_state"] == r["local_numel"][0]


def test_unwrap_model_for_generation_gathers_once():
    """``zero.unwrap_model_for_generation(engine)``: inside, every ZeRO-3 parameter is resident and forwards trigger no
    fetch / release; afterwards the hooks are back and training continues."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.zero import ZeroParamStatus, unwrap_model_for_generation
    from hcache_deepspeed_b200.runtime.zero import partition_parameters as PP
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1))
    engine, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "steps_per_print": 1000,
        "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}})
    x = torch.randn(2, 8)
    want = engine(x).detach().clone()
    assert all(p.ds_status == ZeroParamStatus.NOT_AVAILABLE for p in model.parameters())
    calls = {"n": 0}
    orig = PP.all_gather_params

    def counting(params, async_op=False):
        calls["n"] += sum(1 for p in params if PP.is_zero_param(p) and p.ds_status != ZeroParamStatus.AVAILABLE)
        return orig(params, async_op=async_op)

    with unwrap_model_for_generation(engine):
        PP.all_gather_params = counting
        try:
            assert all(p.ds_status == ZeroParamStatus.AVAILABLE for p in model.parameters())
            with torch.no_grad():
                for _ in range(3):
                    torch.testing.assert_close(model(x), want)
        finally:
            PP.all_gather_params = orig
        assert calls["n"] == 0 and all(p.ds_status == ZeroParamStatus.AVAILABLE for p in model.parameters())
    assert all(p.ds_status == ZeroParamStatus.NOT_AVAILABLE for p in model.parameters())
    loss = engine(x).pow(2).mean()
    engine.backward(loss)
    engine.step()
    assert float((engine(x).detach() - want).abs().max()) > 0


def test_engine_config_getters_follow_the_config():
    """The reference's one-line engine getters (engine.py:480-960), generated here from one table."""
    import hcache_deepspeed_b200 as ds
    cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}}, "steps_per_print": 1000,
           "zero_optimization": {"stage": 2, "reduce_bucket_size": 1234, "allgather_bucket_size": 4321, "overlap_comm": True,
                                 "contiguous_gradients": False, "round_robin_gradients": True},
           "flops_profiler": {"enabled": False, "profile_step": 7, "top_modules": 3, "module_depth": 2},
           "eigenvalue": {"enabled": False, "max_iter": 11, "tol": 0.5},
           "data_efficiency": {"enabled": True, "data_sampling": {"enabled": False},
                               "data_routing": {"enabled": True, "random_ltd": {"enabled": False}}},
           "compression_training": {"weight_quantization": {"shared_parameters": {"enabled": True, "quantize_groups": 4,
                                                                                   "quantization_type": "asymmetric"}}}}
    engine, _, _, _ = ds.initialize(model=torch.nn.Linear(4, 4), config=cfg)
    assert engine.zero_reduce_bucket_size() == 1234 and engine.zero_allgather_bucket_size() == 4321
    assert engine.zero_overlap_comm() is True and engine.zero_contiguous_gradients() is False and engine.zero_round_robin_gradients()
    assert engine.zero_optimization_partition_gradients() and not engine.zero_optimization_partition_weights()
    assert not engine.zero_use_cpu_optimizer() and engine.zero_partial_offload() == 1.0 and engine.mics_shard_size() == -1
    assert engine.flops_profiler_profile_step() == 7 and engine.flops_profiler_top_modules() == 3 and engine.flops_profiler_module_depth() == 2
    assert not engine.flops_profiler_enabled() and not engine.autotuning_enabled()
    assert engine.eigenvalue_max_iter() == 11 and engine.eigenvalue_tol() == 0.5 and not engine.eigenvalue_enabled()
    assert engine.data_efficiency_enabled() and not engine.data_sampling_enabled() and not engine.random_ltd_enabled()
    assert not engine.curriculum_learning_enabled() and engine.get_pld_theta() is None
    q = engine.quantize_training()
    assert q[1] is True and q[2] == 4 and q[5] == "asymmetric"
    assert engine.is_map_style_dataset([1, 2]) and not engine.is_iterable_style_dataset([1, 2])
    assert engine.autotuning_metric_path().endswith("autotuning_metric.json")
    assert engine.get_sequence_parallel_group() is engine.seq_parallel_group
    # every getter of the table is callable and none raises on a minimal config
    from hcache_deepspeed_b200.runtime.engine_accessors import _CONFIG_GETTERS
    bare, _, _, _ = ds.initialize(model=torch.nn.Linear(4, 4), config={"train_batch_size": 1})
    for name in _CONFIG_GETTERS:
        getattr(bare, name)()


@pytest.mark.parametrize("stage", [0, 1, 2, 3])
def test_engine_public_helpers_2ranks(stage):
    from tests.runtime_helper import engine_accessors_worker
    for r in run_distributed(engine_accessors_worker, 2, stage):
        a, b, mean = r["bucket"]
        assert a == mean and b == 10 * mean, r
        assert r["scalars"] == [5, 6] and r["sparse_rows"] == [0, 1], r
        assert r["sparse_dense"] == [0.5, 1.0], r
        assert "optim_states" in r["kinds"] and "lp_params" in r["kinds"] and r["kinds_after"] == 0, r
        assert abs(r["loss_pair"][0] - r["loss_pair"][1]) < 1e-7, r
        assert r["released"] and r["loaded"], r
        if stage >= 1:
            assert r["master_follows_load"] == 0.0 and r["lr"] == (0.125, 0.125) and r["overflow"] is False, r
            assert r["grad_shape_ok"] and r["norm_pair"][0] == pytest.approx(r["norm_pair"][1], rel=1e-5), r
            assert r["parts"] == [[0, 1, 2, 3], [4, 5, 6]] and r["scale"] == 4.0, r
            assert r["pinfo"] == ([4, 5], [3], 3) and r["lean_keys"] == ["exp_avg", "exp_avg_sq"], r


def test_pipeline_engine_small_public_methods_2stages(tmp_path, monkeypatch):
    from tests.runtime_helper import pipe_api_worker
    monkeypatch.setenv("HDS_TEST_PIPE_DIR", str(tmp_path))
    res = run_distributed(pipe_api_worker, 2)
    assert res[0]["extra"].keys() == {"l1"} and res[0]["extra"] == res[1]["extra"] and res[0]["extra"]["l1"] > 0
    for r in res:
        assert r["micro_batches"] == (3, 3) and r["interval"] == 1 and r["reloaded"], r
        assert r["loss_new_shape"] == res[0]["loss_new_shape"] and r["loss_ckpt"] == res[0]["loss_ckpt"]
        assert r["mem"] is None
    assert res[0]["files"] == ["layer_00-model_00-model_states.pt"] and res[1]["files"] == ["layer_02-model_00-model_states.pt"]


def test_zero_offload_public_step_names():
    """The reference's names for the fetch / release steps and persistence marking of ``DeepSpeedZeRoOffload``
    (parameter_offload.py:200-520), driven by hand on a ZeRO-3 engine."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.zero import ZeroParamStatus
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 2))
    engine, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "steps_per_print": 1000,
        "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}})
    off = engine.optimizer.parameter_offload
    lin = model[0]
    off.pre_sub_module_forward_function(lin)
    assert all(p.ds_status == ZeroParamStatus.AVAILABLE for p in lin.parameters())
    off.post_sub_module_forward_function(lin)
    assert all(p.ds_status == ZeroParamStatus.NOT_AVAILABLE for p in lin.parameters())
    off.pre_sub_module_backward_function(model[2])
    off.partition_all_parameters()
    assert all(p.ds_status == ZeroParamStatus.NOT_AVAILABLE for p in model.parameters())
    newly = off.mark_persistent_parameters(param_threshold=16, model_threshold=20)
    assert sorted(p.ds_numel for p in newly) == [2, 16] and all(p.ds_persist and p.ds_status == ZeroParamStatus.AVAILABLE for p in newly)
    x = torch.randn(2, 8)
    loss = engine(x).pow(2).mean()
    engine.backward(loss)
    engine.step()
    assert all(p.ds_status == ZeroParamStatus.AVAILABLE for p in newly)         # persistent ones stay resident through a step
    off.empty_partition_cache()
    n_hooks = len(off.hooks)
    off.destroy()
    off.setup_zero_stage3_hooks()
    assert len(off.hooks) == n_hooks


def test_runtime_utils_helpers_single_process():
    import math
    from hcache_deepspeed_b200.runtime import utils as U
    a, b = torch.nn.Parameter(torch.tensor([3.0, 0.0])), torch.nn.Parameter(torch.tensor([[4.0]]))
    a.grad, b.grad = torch.tensor([0.0, 6.0]), torch.tensor([[8.0]])
    assert U.get_weight_norm([a, b]) == pytest.approx(5.0) and U.get_weight_norm([a, b], norm_type=math.inf) == 4.0
    assert U.get_flattened_grad_norm([a, b]) == pytest.approx(10.0) and U.get_grad_zeros([a, b]) == 1
    assert U.get_flattened_grad_norm([a, b], grad_norm_mask=[[[1, 2]], []]) == pytest.approx(8.0)       # a's element 1 left out
    assert U.get_norm_with_moe_layers(3.0, None, {}) == pytest.approx(3.0) and U.get_norm_with_moe_layers(-1, None, {}) == -1
    t = torch.tensor([1.0, float("nan"), float("inf")])
    U.mask_nan_or_inf_with_val_inplace(t)
    assert t.tolist() == [1.0, -1.0, -1.0]
    nest = {"x": [torch.ones(2), (torch.zeros(1), 5)], "y": "s"}
    cp = U.copy_to_device(nest, "cpu", lambda t: True)
    assert U.compare_tensors_in_structures(nest, cp) and cp is not nest and cp["x"] is not nest["x"]
    assert U.move_to_device(nest, "cpu", lambda t: True)["x"][0] is nest["x"][0]
    assert not U.compare_tensors_in_structures(nest, {"x": [torch.ones(2), (torch.ones(1), 5)], "y": "s"})
    opt = U.DummyOptim([a, b])
    opt.zero_grad()
    assert a.grad is None and len(opt.param_groups[0]["params"]) == 2
    from transformers.pytorch_utils import Conv1D
    c = Conv1D(6, 4)                                    # weight stored [in=4, out=6]
    lin = U.TLinear(c)
    x = torch.randn(3, 4)
    torch.testing.assert_close(lin(x), c(x))
    calls = []
    U.graph_process(False, lambda v: calls.append(v), torch.ones(1))       # CPU tensors: plain call
    assert len(calls) == 1 and U.get_ma_status() == 0 and U.memory_status("x") is None
    with U.noop_context():
        pass


def test_partitioned_tensor_and_group_names_2ranks():
    from tests.runtime_helper import runtime_utils_worker
    res = run_distributed(runtime_utils_worker, 2)
    for rank, r in enumerate(res):
        assert r["even"] == (True, True, 12, [4, 6]) and r["uneven"][:2] == (True, True) and r["uneven"][3] == [3, 5], r
        assert r["gathered"] == [1.0] * 4 + [2.0] * 4 and r["gathered_sharded"] == [0.0, 1.0, 2.0, 3.0, 10.0, 11.0, 12.0, 13.0], r
        assert r["mpu"] == (1, 0, 2, rank, rank) and r["mpu_pinned"] == (2, rank, 0) and r["obj"] == {"a": 0}, r
    assert res[0]["uneven"][2] + res[1]["uneven"][2] == 15


def test_zero3_async_gather_handles_and_quantizer():
    """``param.all_gather(async_op=True)`` / ``all_gather_coalesced`` hand back handles to wait on; ``free_param``;
    the qwZ quantiser round trip (ref: partition_parameters.py:640-780, 1230)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.zero import ZeroParamStatus
    from hcache_deepspeed_b200.runtime.zero import partition_parameters as PP
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Linear(16, 2))
    want = [p.detach().clone() for p in model.parameters()]
    ds.initialize(model=model, config={"train_micro_batch_size_per_gpu": 1, "optimizer": {"type": "SGD", "params": {"lr": 0.1}},
                                       "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}, "steps_per_print": 1000})
    ps = list(model.parameters())
    assert all(p.ds_status == ZeroParamStatus.NOT_AVAILABLE for p in ps)
    h = ps[0].all_gather_coalesced(ps[:2])
    assert isinstance(h, PP.AllGatherCoalescedHandle)
    h.wait()
    h.wait()
    assert all(p.ds_status == ZeroParamStatus.AVAILABLE and torch.equal(p.data, w) for p, w in zip(ps[:2], want[:2]))
    assert isinstance(ps[0].all_gather_coalesced(ps[:2]), PP.NoGatherCoalescedHandle)
    h2 = ps[2].all_gather(async_op=True)
    PP.MultipleAllGatherHandles([h2, PP.NoGatherHandle(ps[0])]).wait()
    assert ps[2].ds_status == ZeroParamStatus.AVAILABLE and torch.equal(ps[2].data, want[2])
    PP.free_param(ps[2])
    assert ps[2].ds_status == ZeroParamStatus.NOT_AVAILABLE and ps[2].numel() == 0
    ps[2].all_gather()
    assert torch.equal(ps[2].data, want[2])
    q = PP.CUDAQuantizer(group_size=64)
    x = torch.randn(1000)
    codes, scale = q.quantize(x)
    back = q.dequantize(codes, scale).reshape(-1)[:1000]
    assert codes.dtype == torch.int8 and float((back - x).abs().max()) < float(x.abs().max()) / 100
    assert torch.nn.Linear in PP.get_all_subclasses(torch.nn.Module) and PP.InsertPostInitMethodToModuleSubClasses is PP.Init
    f = PP.zero_wrapper_for_fp_tensor_constructor(torch.ones, torch.bfloat16)
    assert f(3).dtype == torch.bfloat16 and f(3, dtype=torch.int32).dtype == torch.int32


def test_zero3_detects_external_parameters_automatically():
    """A forward that reads another module's partitioned parameter (``self.other.weight``) triggers the gather through the
    ``ZeROOrderedDict`` getter and registers the parameter as external - no manual ``register_external_parameter``
    (ref: parameter_offload.py:40-75)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.runtime.zero import ZeroParamStatus

    class Tied(torch.nn.Module):
        def __init__(self, owner):
            super().__init__()
            self.__dict__["owner"] = owner                  # not a sub-module: invisible to the hook installer

        def forward(self, h):
            return h @ self.owner.weight.t()                # reads the embedding's weight

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.emb = torch.nn.Embedding(11, 6)
            self.mid = torch.nn.Linear(6, 6)
            self.head = Tied(self.emb)

        def forward(self, ids):
            return self.head(torch.tanh(self.mid(self.emb(ids)))).pow(2).mean()

    def run(stage):
        torch.manual_seed(0)
        engine, _, _, _ = ds.initialize(model=Net(), config={
            "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "steps_per_print": 1000,
            "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0}})
        ids = torch.tensor([[1, 2, 3], [4, 5, 6]])
        losses = []
        for _ in range(3):
            loss = engine(ids)
            engine.backward(loss)
            engine.step()
            losses.append(float(loss))
        return engine, losses

    e3, l3 = run(3)
    _, l0 = run(0)
    assert l3 == pytest.approx(l0, rel=1e-5)
    w = e3.module.emb.weight
    assert id(w) in getattr(e3.module.head, "_external_params", {})
    assert w.ds_status == ZeroParamStatus.NOT_AVAILABLE      # released again after the step


def test_mics_hierarchical_all_gather_4ranks():
    from tests.runtime_helper import mics_hierarchy_worker
    want = [v for r in range(4) for v in (10.0 * r, 10.0 * r + 1)]
    for r in run_distributed(mics_hierarchy_worker, 4):
        assert r["hier"] and not r["flat"] and (r["intra"], r["inter"]) == (2, 2), r
        assert r["two_hop"] == want and r["one_hop"] == want, r


def test_zeropp_loco_error_feedback_2ranks():
    """``zeropp_loco_param``: the quantised gradient exchange keeps what quantisation lost per parameter and adds it back
    next time (ref: coalesced_collectives.py all_to_all_loco_quant_reduce, stage3 zeropp_loco_param)."""
    from tests.runtime_helper import zeropp_worker
    for r in run_distributed(zeropp_worker, 2, False, True, {"err_beta": 0.8, "reset_T": 16}):
        assert r["losses"][-1] < 0.5 * r["losses"][0] and r["qgz"]["calls"] > 0, r["losses"][-1]
        assert r["loco_buffers"] >= 2 and max(r["loco_ages"]) <= 16, r
        assert r["fed_back_err"] < 0.5 * r["one_shot_err"], r
