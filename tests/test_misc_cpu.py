"""Smaller subsystems: compression, autotuning, 1-bit communication/optimizers, NVMe swapping, Megatron checkpoint
re-sharding, sparse-attention layouts, Evoformer attention, accelerator / op-builder surfaces, data sampling."""
import math
import os
import subprocess
import types

import numpy as np
import pytest
import torch

from tests.dist_utils import run_distributed
from tests.parallel_helper import onebit_adam_worker, onebit_worker

pytestmark = pytest.mark.timeout(600)


class _MLP(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.fc1, self.fc2 = torch.nn.Linear(16, 32), torch.nn.Linear(32, 8)

    def forward(self, x):
        return self.fc2(torch.relu(self.fc1(x)))


def test_compression_schedule_and_clean():
    from hcache_deepspeed_b200.compression import compression_scheduler, init_compression, redundancy_clean
    from hcache_deepspeed_b200.compression.basic_layer import LinearLayer_Compress
    from hcache_deepspeed_b200.compression.compress import get_compression_config
    cfg = {"compression_training": {
        "weight_quantization": {"shared_parameters": {"enabled": True, "schedule_offset": 2, "quantize_groups": 1,
                                                      "quantization_type": "symmetric"},
                                "different_groups": {"g": {"params": {"start_bits": 8, "target_bits": 4, "quantization_period": 10},
                                                           "modules": ["fc"]}}},
        "row_pruning": {"shared_parameters": {"enabled": True, "schedule_offset": 0, "method": "l1"},
                        "different_groups": {"r": {"params": {"dense_ratio": 0.5}, "modules": ["fc1"], "related_modules": [["fc2"]]}}},
        "sparse_pruning": {"shared_parameters": {"enabled": True, "schedule_offset": 1, "method": "l1"},
                           "different_groups": {"s": {"params": {"dense_ratio": 0.25}, "modules": ["fc2"]}}}}}
    torch.manual_seed(0)
    m = _MLP()
    init_compression(m, cfg)
    assert isinstance(m.fc1, LinearLayer_Compress) and isinstance(m.fc2, LinearLayer_Compress)
    sch = compression_scheduler(m, get_compression_config(cfg))
    x = torch.randn(4, 16)
    sch.step()
    assert m.fc1.row_pruning_enabled and m.fc2.sparse_pruning_enabled and not m.fc1.weight_quantization_enabled
    y = m(x)
    assert int((m.fc1(x) == 0).all(dim=0).sum()) >= 16          # half of fc1's rows are masked
    sch.step()
    sch.step()
    assert m.fc1.weight_quantization_enabled
    m(x).sum().backward()
    assert m.fc1.weight.grad is not None                          # straight-through estimator
    redundancy_clean(m, cfg)
    assert m.fc1.weight.shape == (16, 16) and m.fc2.weight.shape == (8, 16) and m(x).shape == (4, 8)
    assert len(m.fc2.weight.unique()) <= 17 and float((m.fc2.weight == 0).float().mean()) >= 0.5


def test_engine_steps_the_compression_scheduler():
    """``compression_training`` in the engine config: techniques switch on at their schedule offsets as the engine takes
    optimizer steps (ref: engine.py _configure_compression_scheduler / _take_model_step)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.compression import init_compression
    comp = {"weight_quantization": {"shared_parameters": {"enabled": True, "schedule_offset": 3, "quantize_groups": 1,
                                                          "quantization_type": "symmetric"},
                                    "different_groups": {"g": {"params": {"start_bits": 8, "target_bits": 8, "quantization_period": 10},
                                                               "modules": ["fc1"]}}},
            "sparse_pruning": {"shared_parameters": {"enabled": True, "schedule_offset": 1, "method": "l1"},
                               "different_groups": {"s": {"params": {"dense_ratio": 0.5}, "modules": ["fc2"]}}}}
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "SGD", "params": {"lr": 0.01}}, "steps_per_print": 1000,
           "compression_training": comp}
    torch.manual_seed(0)
    m = _MLP()
    init_compression(m, cfg)
    engine, _, _, _ = ds.initialize(model=m, config=cfg)
    assert engine.compression_scheduler is not None
    seen = []
    for _ in range(4):
        loss = engine(torch.randn(4, 16)).pow(2).mean()
        engine.backward(loss)
        engine.step()
        seen.append((m.fc2.sparse_pruning_enabled, m.fc1.weight_quantization_enabled))
    assert seen == [(True, False), (True, False), (True, True), (True, True)]


def test_autotuner_picks_best_feasible_config(tmp_path):
    from hcache_deepspeed_b200.autotuning import Autotuner

    def run(cfg):
        z, mbs = cfg["zero_optimization"]["stage"], cfg["train_micro_batch_size_per_gpu"]
        return None if (z <= 1 and mbs > 4) else mbs * 10 - z * 3            # OOM for big batches without ZeRO-2+

    for tuner in ("gridsearch", "random", "model_based"):
        a = Autotuner({"fp16": {"enabled": True}, "autotuning": {"tuner_type": tuner, "num_tuning_micro_batch_sizes": 4,
                                                                  "tuner_early_stopping": 100, "search": "flat"}},
                      num_params=int(30e9), num_gpus=8, run_fn=run)
        assert a.feasible_stages() == [1, 2, 3]                              # 30B fp16 states: 600 GB unpartitioned
        best = a.tune()
        assert (best["zero_optimization"]["stage"], best["train_micro_batch_size_per_gpu"]) == (2, 8) and a.best_metric == 74
        assert os.path.exists(a.write_optimal_config(str(tmp_path / f"{tuner}.json")))


def test_autotuner_staged_search_walks_stages_and_micro_batch_sizes(tmp_path):
    """The reference's order of work: per ZeRO stage find the runnable micro-batch range (doubling + bisection around the
    out-of-memory point), measure a short list up to the plateau, stop walking stages when one does not improve (fast)."""
    from hcache_deepspeed_b200.autotuning import Autotuner
    oom_above = {1: 12, 2: 40, 3: 100}
    calls = []

    def run(cfg):
        z, mbs = cfg["zero_optimization"]["stage"], cfg["train_micro_batch_size_per_gpu"]
        calls.append((z, mbs))
        if mbs > oom_above[z]:
            return None
        return 1000.0 * mbs / (mbs + 8.0) - {1: 20.0, 2: 40.0, 3: 150.0}[z]   # saturating throughput, a cost per ZeRO stage

    a = Autotuner({"fp16": {"enabled": True}, "gradient_accumulation_steps": 2,
                   "autotuning": {"num_tuning_micro_batch_sizes": 4, "max_train_micro_batch_size_per_gpu": 256, "fast": True}},
                  num_params=int(30e9), num_gpus=8, run_fn=run, user_args=["--bs", "16", "--lr=0.5"])
    assert a.metric() == "throughput" and a.fast_enabled() and a.fp16_enabled() and a.mp_size() == 1
    assert a.get_gas_from_user_config() == 2 and a.get_val_from_user_args("$--bs") == 16 and a.get_val_from_user_args("$--lr") == 0.5
    assert a.get_instantiation_memory_required_per_gpu(3) < a.get_instantiation_memory_required_per_gpu(1) < a.get_gpu_memory_info()
    assert a.get_tuning_micro_batch_size_list(1, 13, 4) == [1, 5, 9, 13] and a.get_tuning_micro_batch_size_list(4, 4, 3) == [4]
    lo, hi = a.get_min_max_micro_batch_size(1, 1, 256)
    assert lo == 1 and 10 <= hi <= 12                                   # 8 runs, 16 does not: bisected to within ~12 %
    best = a.tune()
    recs = a.get_best_space_records()
    assert set(recs) == {"z1", "z2", "z3", "global"}
    z1, z2, z3 = recs["z1"][1], recs["z2"][1], recs["z3"][1]
    assert z2 > z1 and z2 > z3                                          # stage 2 wins: larger batches than 1, cheaper than 3
    assert best["zero_optimization"]["stage"] == 2 and 30 <= best["train_micro_batch_size_per_gpu"] <= 40
    assert best["gradient_accumulation_steps"] == 2 and a.best_metric == z2 == recs["global"][1]
    assert 16 <= a.get_plateau_mbs("z2") <= best["train_micro_batch_size_per_gpu"]   # (over all sizes probed in z2)
    assert all(m <= 256 for _, m in calls) and len(calls) < 60
    assert os.path.exists(a.write_optimal_config(str(tmp_path / "best.json")))
    # without fast mode the stage's other knobs are searched at the best micro-batch size
    b = Autotuner({"fp16": {"enabled": True}, "zero_optimization": {"stage": 1},
                   "autotuning": {"fast": False, "micro_batch_sizes": [2, 4], "tuner_num_trials": 4}},
                  num_params=int(30e9), num_gpus=8, run_fn=run)
    b.tune()
    knob_runs = [e for e, v, _ in b.records["z1"] if "reduce_bucket_size" in e["zero_optimization"]]
    assert len(knob_runs) == 4 and all(e["train_micro_batch_size_per_gpu"] == 4 for e in knob_runs)


def test_onebit_compressed_allreduce_two_ranks():
    res = run_distributed(onebit_worker, 2)
    assert res[0]["rel"] < 0.15 and res[0]["same"] == pytest.approx(res[1]["same"])


def test_onebit_adam_trains_and_freezes():
    res = run_distributed(onebit_adam_worker, 2)
    for r in res:
        assert r["frozen"] and r["allreduce_off"] and all(math.isfinite(l) for l in r["losses"])
    mean = [(res[0]["losses"][i] + res[1]["losses"][i]) / 2 for i in range(14)]
    assert mean[-1] < 0.7 * mean[0]
    assert float((res[0]["w"] - res[1]["w"]).abs().max()) < 0.05           # replicas stay close under 1-bit momentum


def test_nvme_swappers(tmp_path):
    from hcache_deepspeed_b200.io.aio import aio_handle
    from hcache_deepspeed_b200.runtime.swap_tensor import AsyncPartitionedParameterSwapper, AsyncTensorSwapper
    sw = AsyncPartitionedParameterSwapper(model_dtype=torch.float32, nvme_path=str(tmp_path), buffer_count=3, buffer_size=1 << 20,
                                          aio_config={"block_size": 4096})
    sw.min_aio_bytes = 0
    ps = []
    for i in range(2):
        t = torch.randn(5000 + i)
        t.ds_numel = t.numel()
        ps.append(types.SimpleNamespace(ds_id=i, ds_tensor=t))
    want = [p.ds_tensor.clone() for p in ps]
    sw.swap_out_and_release(ps)
    assert ps[0].ds_tensor.numel() == 1 and sw.available_swap_in_buffers() == 3
    sw.swap_in(ps, async_op=True)
    sw.synchronize_reads()
    for p, w in zip(ps, want):
        assert torch.equal(p.ds_tensor.data, w)
    # direct reads into a caller's buffer, updating the files from fp32 slices, a pool buffer for a new partition
    sw.remove_partition_and_release_buffers(ps)
    dest = torch.zeros(sw._aligned(5000))
    sw.swap_into_buffer(ps[0], dest)
    assert torch.equal(dest[:5000], want[0]) and sw.available_swap_in_buffers() == 3
    odd = torch.zeros(5003)                                  # not aligned: goes through a pool buffer and gives it back
    sw.swap_into_buffer(ps[1], odd)
    assert torch.equal(odd[:5001], want[1]) and sw.available_swap_in_buffers() == 3
    sw.reserve_partitioned_swap_space([5000, 5001])
    new = [w.double() * 2 for w in want]
    sw.swap_out_partitioned_params(ps, new)
    sw.swap_in(ps, async_op=False)
    for p, w in zip(ps, want):
        assert torch.equal(p.ds_tensor.data, w * 2)
    sw.remove_partition_and_release_buffers(ps)
    extra = types.SimpleNamespace(ds_id=7, ds_tensor=torch.zeros(1))
    buf = sw.get_buffer(extra, 1234)
    assert buf.numel() == 1234 and sw.available_swap_in_buffers() == 2
    ats = AsyncTensorSwapper(aio_handle(4096, 8, False, True, 1), 256)
    ats.add_buffers([torch.empty(1 << 16) for _ in range(2)])
    ts = [torch.randn(3000) for _ in range(40)]
    paths = [str(tmp_path / f"t{i}.swp") for i in range(40)]
    ats.swap_out_tensors(ts, paths)
    ats.release_buffers()
    assert torch.equal(torch.from_numpy(np.fromfile(paths[17], dtype=np.float32)[:3000]), ts[17])


def test_megatron_state_dict_split_then_merge(tmp_path):
    from hcache_deepspeed_b200.runtime.state_dict_factory import MegatronSDLoader
    full = {"attention.query_key_value.weight": torch.arange(24.).reshape(12, 2), "attention.dense.weight": torch.arange(16.).reshape(4, 4),
            "mlp.dense_h_to_4h.weight": torch.arange(32.).reshape(8, 4), "mlp.dense_4h_to_h.weight": torch.arange(32.).reshape(4, 8),
            "ln.weight": torch.ones(4)}
    torch.save({"module": full}, tmp_path / "mp0.pt")
    loader = MegatronSDLoader([str(tmp_path / "mp0.pt")], 0)
    parts = [loader.load(2, r)[1]["module"] for r in range(2)]
    assert parts[0]["attention.dense.weight"].shape == (4, 2) and parts[0]["attention.query_key_value.weight"].shape == (6, 2)
    for r in range(2):
        torch.save({"module": parts[r]}, tmp_path / f"s{r}.pt")
    merged = MegatronSDLoader([str(tmp_path / "s0.pt"), str(tmp_path / "s1.pt")], 0).load(1, 0)[1]["module"]
    for k in full:
        assert torch.equal(merged[k], full[k]), k


def test_sparse_attention_layouts_and_op():
    from hcache_deepspeed_b200.ops.sparse_attention import (BigBirdSparsityConfig, BSLongformerSparsityConfig, DenseSparsityConfig,
                                                            FixedSparsityConfig, LocalSlidingWindowSparsityConfig,
                                                            SparseSelfAttention, VariableSparsityConfig)
    fixed = FixedSparsityConfig(2, block=4, num_local_blocks=2, attention="unidirectional").make_layout(32)
    assert fixed.shape == (2, 8, 8) and torch.equal(fixed[0], torch.tril(fixed[0])) and int(fixed[0, 3, 2]) == 1
    for cfg in (VariableSparsityConfig(2, block=4, local_window_blocks=[2]), BigBirdSparsityConfig(2, block=4),
                BSLongformerSparsityConfig(2, block=4), LocalSlidingWindowSparsityConfig(2, block=4)):
        lay = cfg.make_layout(32)
        assert lay.shape == (2, 8, 8) and all(int(lay[0, i, i]) == 1 for i in range(8))
    with pytest.raises(ValueError):
        DenseSparsityConfig(2, block=4).make_layout(30)
    q, k, v = (torch.randn(1, 2, 16, 8) for _ in range(3))
    att = SparseSelfAttention(DenseSparsityConfig(2, block=4), max_seq_length=32)
    torch.testing.assert_close(att(q, k, v), torch.nn.functional.scaled_dot_product_attention(q, k, v))
    local = SparseSelfAttention(LocalSlidingWindowSparsityConfig(2, block=4, num_sliding_window_blocks=1), max_seq_length=32)
    out = local(q, k, v)     # block-diagonal, block-causal: the first block never sees later tokens
    want0 = torch.nn.functional.scaled_dot_product_attention(q[:, :, :4], k[:, :, :4], v[:, :, :4])
    torch.testing.assert_close(out[:, :, :4], want0)


def test_evoformer_attention_matches_manual():
    from hcache_deepspeed_b200.ops.deepspeed4science import DS4Sci_EvoformerAttention
    torch.manual_seed(0)
    Q, K, V = (torch.randn(1, 3, 6, 2, 4) for _ in range(3))
    b1, b2 = torch.randn(1, 3, 1, 1, 6), torch.randn(1, 1, 2, 6, 6)
    out = DS4Sci_EvoformerAttention(Q, K, V, [b1, b2])
    q, k, v = (t.permute(0, 1, 3, 2, 4) for t in (Q, K, V))
    s = q @ k.transpose(-1, -2) / 2.0 + b1 + b2
    want = (torch.softmax(s, -1) @ v).permute(0, 1, 3, 2, 4)
    torch.testing.assert_close(out, want, rtol=1e-5, atol=1e-5)


def test_accelerator_and_op_builders():
    from hcache_deepspeed_b200.accelerator import get_accelerator
    from hcache_deepspeed_b200.ops.op_builder import ALL_OPS, CPUAdamBuilder
    acc = get_accelerator()
    assert acc.communication_backend_name() in ("nccl", "gloo") and acc.is_available()
    assert len(ALL_OPS) >= 20 and CPUAdamBuilder().load() is not None and ALL_OPS["async_io"].is_compatible()


def test_curriculum_sampler_and_random_ltd():
    from hcache_deepspeed_b200.runtime.data_pipeline.data_sampler import DeepSpeedDataSampler
    from hcache_deepspeed_b200.runtime.data_pipeline.random_ltd import RandomLayerTokenDrop, RandomLTDScheduler
    cfg = {"data_sampling": {"curriculum_learning": {"enabled": True, "curriculum_metrics": {"len": {
        "min_difficulty": 2, "max_difficulty": 10, "schedule_type": "fixed_linear",
        "schedule_config": {"total_curriculum_step": 5, "difficulty_step": 1}}}}}}
    diff = np.arange(100) % 11
    s = DeepSpeedDataSampler(cfg, 100, 2, 0, 2, difficulties={"len": diff})
    it = iter(s)
    first = next(it)
    assert len(first) == 2 and all(diff[i] <= 2 for i in first)
    sd = s.state_dict()
    s.load_state_dict(sd)
    sched = RandomLTDScheduler({"total_layer_num": 4, "random_ltd_layer_num": 2, "global_batch_size": 2,
                                "random_ltd_schedule": {"min_value": 4, "max_value": 16, "schedule_config": {"require_steps": 10, "seq_per_step": 4}}})
    assert sched.update_seq(0) == 4 and sched.update_seq(5) == 8 and sched.update_seq(100) == 16
    sched.set_current_seq(4)
    layer = RandomLayerTokenDrop(torch.nn.Linear(8, 8), scheduler=sched).train()
    x = torch.randn(2, 16, 8)
    y = layer(x)
    assert y.shape == x.shape and int((y == x).all(-1).sum()) == 2 * 12      # 12 of 16 tokens per sample skipped the layer


def test_multinode_runner_commands():
    from hcache_deepspeed_b200.launcher.multinode_runner import RUNNERS
    args = types.SimpleNamespace(user_script="train.py", user_args=["--x", "1"], master_addr="10.0.0.1", master_port=29500,
                                 hostfile="/job/hostfile", launcher_args="", include="", exclude="", num_nodes=-1, num_gpus=-1,
                                 no_python=False, module=False, comment="")
    res = {"n1": [0, 1], "n2": [0, 1]}
    cmds = {}
    for name, cls in RUNNERS.items():
        r = cls(args, "QUJD")
        r.add_export("NCCL_DEBUG", "INFO")
        cmds[name] = r.get_cmd({}, res)
    assert cmds["pdsh"][:2] == ["pdsh", "-S"] and "n1,n2" in cmds["pdsh"] and "--node_rank=%n" in cmds["pdsh"]
    assert cmds["openmpi"][:3] == ["mpirun", "-n", "4"] and "NCCL_DEBUG=INFO" in cmds["openmpi"]
    i = cmds["mpich"].index("-genv")
    assert cmds["mpich"][i:i + 3] == ["-genv", "NCCL_DEBUG", "INFO"] and cmds["mpich"][3:5] == ["-ppn", "2"]
    assert cmds["impi"][1:3] == ["-hosts", "n1,n2"] and cmds["slurm"][0] == "srun" and "--export=ALL,NCCL_DEBUG=INFO" in cmds["slurm"]
    assert cmds["mvapich"][-3:] == ["train.py", "--x", "1"]
    with pytest.raises(ValueError):
        RUNNERS["mpich"](args, "x").get_cmd({}, {"a": [0], "b": [0, 1]})


def test_nvme_tune_sweep(tmp_path):
    """Sweep -> JSON-lines logs -> best "aio" block (reference: nvme/perf_run_sweep.py, perf_generate_param.py)."""
    from hcache_deepspeed_b200.nvme import tune
    assert tune.parse_size("8M") == 8 << 20 and tune.parse_size("400M") == 400 << 20 and tune.parse_size(4096) == 4096
    space = {"block_size": ["256K", "1M"], "queue_depth": [8], "io_parallel": [1, 2], "use_direct": [False]}
    assert len(tune.sweep_points(space)) == 4
    out = tune.sweep_main(["--nvme_dir", str(tmp_path / "d0"), "--io_size", "2M", "--log_dir", str(tmp_path / "logs")]
                          + ["--sweep_config", _write_json(tmp_path / "space.json", space)])
    aio = out["aio"]
    assert aio["block_size"] in (256 << 10, 1 << 20) and aio["intra_op_parallelism"] in (1, 2) and aio["queue_depth"] == 8
    assert out["read_GBps"] > 0 and out["write_GBps"] > 0
    assert len(os.listdir(tmp_path / "logs" / "read")) == 4 and len(os.listdir(tmp_path / "logs" / "write")) == 4
    assert not os.path.exists(tmp_path / "d0" / "hds_nvme_tune.bin")
    assert tune.generate_main(str(tmp_path / "logs"))["aio"] == aio


def _write_json(path, obj):
    import json
    with open(path, "w") as f:
        json.dump(obj, f)
    return str(path)


def test_comet_monitor_rate_limit():
    from hcache_deepspeed_b200.monitor import CometMonitor, EventsLogScheduler
    from hcache_deepspeed_b200.monitor.config import CometConfig
    s = EventsLogScheduler(100)
    assert [s.needs_logging("loss", x) for x in (0, 50, 100, 150, 230)] == [True, False, True, False, True]

    class _Exp:
        def __init__(self):
            self.logged = []

        def log_metric(self, name, value, step):
            self.logged.append((name, value, step))

    exp = _Exp()
    m = CometMonitor(CometConfig(enabled=True, samples_log_interval=10), experiment=exp)
    m.write_events([("lr", 0.1, 0), ("lr", 0.2, 5), ("loss", 3.0, 5), ("lr", 0.3, 10)])
    assert exp.logged == [("lr", 0.1, 0), ("loss", 3.0, 5), ("lr", 0.3, 10)]


def test_small_utils():
    from hcache_deepspeed_b200.utils import bwc, mixed_precision_linkage as mpl
    from hcache_deepspeed_b200.utils.torch import required_torch_version
    from hcache_deepspeed_b200.utils.types import GATED_ACTIVATION_TYPES, ActivationFuncType, NormType
    assert bwc.bwc_tensor_model_parallel_rank(None) == 0 and bwc.bwc_pipeline_parallel_world_size(None) == 1
    old = types.SimpleNamespace(get_model_parallel_rank=lambda: 3, get_model_parallel_world_size=lambda: 4,
                                get_model_parallel_group=lambda: "g", get_pipe_parallel_world_size=lambda: 2)
    assert bwc.bwc_tensor_model_parallel_rank(old) == 3 and bwc.bwc_tensor_model_parallel_world_size(old) == 4
    assert bwc.bwc_tensor_model_parallel_group(old) == "g" and bwc.bwc_pipeline_parallel_world_size(old) == 2
    assert required_torch_version(min_version="2.0") and not required_torch_version(max_version="1.13")
    assert ActivationFuncType.GATED_SILU in GATED_ACTIVATION_TYPES and NormType.RMSNorm == 3
    # lp params of 6, 10, 8 elements; this rank owns flat elements [4, 16): all of none, tail of p0, all of... etc.
    ps = [torch.arange(6.0).bfloat16(), torch.arange(10.0).bfloat16(), torch.arange(8.0).bfloat16()]
    flat = torch.zeros(12)
    mpl.link_hp_params(ps, flat, partition_start=4, partition_size=12)
    m0, m1, m2 = (p._hp_mapping for p in ps)
    assert (m0.lp_fragment_address.start, m0.lp_fragment_address.numel, m0.hp_fragment_address.start) == (4, 2, 0)
    assert (m1.lp_fragment_address.start, m1.lp_fragment_address.numel, m1.hp_fragment_address.start) == (0, 10, 2)
    assert m2 is None or m2.lp_fragment_address.numel == 0 or True
    assert ps[2]._hp_mapping is None  # [16, 24) lies outside the partition
    for m in (m0, m1):
        m.update_hp()
    assert flat.tolist() == [4, 5] + list(range(10))
    mpl.lazy_init_hp_params_optimizer_state(ps[:2], flat, {flat: {"exp_avg": torch.arange(12.0), "step": torch.tensor(1)}})
    assert m1.get_hp_fragment("exp_avg").tolist() == list(range(2, 12))


def test_diffusers_wrappers_cpu_passthrough():
    """DSUNet / DSVAE / DSClipEncoder forward to the wrapped module (graphs engage only for CUDA tensors)."""
    from hcache_deepspeed_b200.model_implementations.diffusers import DSUNet, DSVAE
    from hcache_deepspeed_b200.model_implementations.features import GraphedCallable
    from hcache_deepspeed_b200.model_implementations.transformers.clip_encoder import DSClipEncoder
    from hcache_deepspeed_b200.module_inject.replace_module import generic_injection

    class UNet(torch.nn.Module):
        in_channels = 4

        def __init__(self):
            super().__init__()
            self.conv = torch.nn.Conv2d(4, 4, 1)

        def forward(self, sample, t, ehs, return_dict=True):
            return self.conv(sample) + t + ehs.mean()

    class VAE(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = torch.nn.Linear(4, 4)

        def encode(self, x):
            return self.lin(x)

        def decode(self, z):
            return self.lin(z) * 2

        def forward(self, x):
            return self.decode(self.encode(x))

    pipe = types.SimpleNamespace(unet=UNet(), vae=VAE(), text_encoder=torch.nn.Linear(4, 4))
    generic_injection(pipe, dtype=torch.float32, enable_cuda_graph=True)
    assert isinstance(pipe.unet, DSUNet) and isinstance(pipe.vae, DSVAE) and isinstance(pipe.text_encoder, DSClipEncoder)
    x = torch.randn(1, 4, 2, 2)
    want = pipe.unet.unet(x, torch.tensor(1.0), torch.ones(3))
    torch.testing.assert_close(pipe.unet(x, torch.tensor(1.0), torch.ones(3)), want)
    assert pipe.unet.in_channels == 4 and not pipe.unet.cuda_graph_created
    assert not any(p.requires_grad for p in pipe.unet.parameters())
    z = torch.randn(2, 4)
    torch.testing.assert_close(pipe.vae.decode(z), pipe.vae.vae.decode(z))
    torch.testing.assert_close(pipe.vae(z), pipe.vae.vae(z))
    torch.testing.assert_close(pipe.text_encoder(z), pipe.text_encoder.enc(z))
    assert pipe.text_encoder.iter == 1
    calls = []
    g = GraphedCallable(lambda a, k=1: calls.append(k) or a * k)
    assert g(torch.ones(2), k=3).tolist() == [3, 3] and calls == [3]


def test_shim_packages_and_random_ltd_autograd():
    import numpy as np
    from hcache_deepspeed_b200.nebula import DeepSpeedNebulaConfig
    from hcache_deepspeed_b200.ops.aio import AsyncIOHandle  # noqa: F401
    from hcache_deepspeed_b200.ops.gds import gds_available
    from hcache_deepspeed_b200.ops.random_ltd import GatherTokens, ScatterTokens, bert_sample_tokens, gpt_sample_tokens
    from hcache_deepspeed_b200.pipe import LayerSpec, PipelineModule, TiedLayerSpec  # noqa: F401
    from hcache_deepspeed_b200.runtime.compression import CupyBackend
    assert isinstance(gds_available(), bool)
    n = DeepSpeedNebulaConfig({"nebula": {"enabled": True, "persistent_storage_path": "/x"}})
    assert n.enabled and n.persistent_storage_path == "/x" and n.persistent_time_interval == 100
    b = CupyBackend()
    bits = torch.rand(64) > 0.5
    chunks = b.compress_by_chunk(bits, 2)
    assert torch.cat(chunks).tolist() == np.packbits(bits.numpy()).tolist()
    assert b.unpackbits(torch.cat(chunks)).bool().tolist() == bits.tolist()

    torch.manual_seed(0)
    idx, mask = gpt_sample_tokens(5, 12, 3, layers=2, attn_mask=torch.ones(3, 1, 12, 12))
    assert idx.shape == (2, 3, 5) and idx.dtype == torch.int32 and mask.shape == (3, 1, 5, 5)
    assert (idx[..., 1:] > idx[..., :-1]).all()
    full_mask = torch.arange(3 * 12 * 12, dtype=torch.float32).reshape(3, 1, 12, 12)
    bidx, bmasks = bert_sample_tokens(4, 12, 3, layers=1, attn_mask=full_mask)
    r = bidx[0, 1].long()
    assert torch.equal(bmasks[0][1, 0], full_mask[1, 0][r][:, r])
    for batch_first in (True, False):
        x = torch.randn(3, 12, 4, requires_grad=True) if batch_first else torch.randn(12, 3, 4, requires_grad=True)
        _, part = GatherTokens.apply(x, idx[0], batch_first)
        assert part.shape == ((3, 5, 4) if batch_first else (5, 3, 4))
        out = ScatterTokens.apply(x, part * 2, idx[0], batch_first)
        out.sum().backward()
        kept = torch.zeros(3, 12)
        kept.scatter_(1, idx[0].long(), 1.0)
        want = (1 + kept)[:, :, None].expand(3, 12, 4)       # dropped tokens: identity; kept tokens: 2x path
        torch.testing.assert_close(x.grad if batch_first else x.grad.transpose(0, 1), want)


def test_elastic_agent_restarts_and_rescales(tmp_path):
    """A failing worker group is restarted (restart count exported); a membership change relaunches with the new world
    size without consuming a restart; restarts are bounded (ref: elasticity/elastic_agent.py:127-190)."""
    import sys
    from hcache_deepspeed_b200.elasticity.elastic_agent import DSElasticAgent, WorkerState
    marks = tmp_path / "marks"
    marks.mkdir()
    script = tmp_path / "worker.py"
    script.write_text(
        "import os, sys, time\n"
        "a = int(os.environ['TORCHELASTIC_RESTART_COUNT']); r = os.environ['RANK']; w = os.environ['WORLD_SIZE']\n"
        f"open(os.path.join({str(marks)!r}, f'a{{a}}_r{{r}}_w{{w}}'), 'w').close()\n"
        "mode = sys.argv[1]\n"
        "if mode == 'flaky' and a == 0 and r == '1': sys.exit(3)\n"
        "if mode == 'always': sys.exit(5)\n"
        "if mode == 'scale' and w == '2': time.sleep(30)\n"
        "time.sleep(0.2)\n")
    res = DSElasticAgent([sys.executable, str(script), "flaky"], 2, max_restarts=2, monitor_interval=0.05).run()
    assert res.state == WorkerState.SUCCEEDED and res.restarts == 1 and res.world_sizes == [2, 2]
    # (rank 0 of attempt 0 may be stopped before it gets to write its mark)
    assert {"a0_r1_w2", "a1_r0_w2", "a1_r1_w2"} <= set(os.listdir(marks))
    res = DSElasticAgent([sys.executable, str(script), "always"], 1, max_restarts=1, monitor_interval=0.05).run()
    assert res.state == WorkerState.FAILED and res.restarts == 1 and res.return_codes[0] == 5
    nodes = iter([1, 1, 1, 2])
    last = [1]

    def members():
        last[0] = next(nodes, last[0] if last[0] == 2 else 2)
        return last[0]

    res = DSElasticAgent([sys.executable, str(script), "scale"], 2, max_restarts=0, monitor_interval=0.05, membership_fn=members).run()
    assert res.state == WorkerState.SUCCEEDED and res.restarts == 0 and res.world_sizes == [2, 4]


def test_top_level_namespace_mirrors_deepspeed():
    import hcache_deepspeed_b200 as ds
    for name in ("zero", "moe", "pipe", "checkpoint", "module_inject", "ops", "checkpointing", "DeepSpeedEngine", "PipelineModule",
                 "InferenceEngine", "DeepSpeedInferenceConfig", "DeepSpeedTransformerInference", "OnDevice", "tp_model_init",
                 "init_inference", "initialize", "init_distributed", "add_config_arguments", "comm"):
        assert getattr(ds, name) is not None, name
    assert hasattr(ds.zero, "Init") and hasattr(ds.pipe, "PipelineModule") and not ds.is_compile_supported()
    with pytest.raises(AttributeError):
        ds.no_such_thing


def test_linear_init_context_and_domino_linears():
    from hcache_deepspeed_b200.linear import Init, LoRAConfig, LoRAOptimizedLinear
    from hcache_deepspeed_b200.runtime.domino import DominoAsyncColumnParallelLinear, RowParallelLinearNoComm
    with Init(lora_config=LoRAConfig(lora_r=4), dtype=torch.float32) as ctx:
        m = torch.nn.Sequential(torch.nn.Linear(16, 32, bias=False), torch.nn.ReLU(), torch.nn.Linear(32, 8))
    assert isinstance(m[0], LoRAOptimizedLinear) and type(m[2]) is torch.nn.Linear and torch.nn.Linear.__name__ == "Linear"
    ctx.finalize(m)
    assert sorted(n for n, p in m.named_parameters() if p.requires_grad) == ["0.lora_weight_1.weight", "0.lora_weight_2.weight"]
    m(torch.randn(2, 16)).sum().backward()
    # single-process Domino linears == plain linears; the handle dict stays empty without a tensor-parallel group
    torch.manual_seed(0)
    col, row = DominoAsyncColumnParallelLinear(8, 12, bias=True), RowParallelLinearNoComm(12, 8, bias=True)
    x = torch.randn(3, 8, requires_grad=True)
    handles = {}
    h, _ = col(x, handles, 0)
    y, b = row(torch.relu(h))
    (y + b).sum().backward()
    ref = torch.nn.functional.linear(torch.relu(torch.nn.functional.linear(x, col.weight, col.bias)), row.weight, row.bias)
    assert torch.allclose(y + b, ref, atol=1e-6) and handles == {} and x.grad is not None and col.weight.grad is not None


def test_module_quantize_and_launcher_helper(monkeypatch):
    from types import SimpleNamespace

    from hcache_deepspeed_b200.launcher.launcher_helper import build_command, env_mapping, parse_args
    from hcache_deepspeed_b200.module_inject.module_quantize import quantize_transformer_layer

    class Layer(torch.nn.Module):
        def __init__(self):
            super().__init__()
            lin = lambda: torch.nn.Linear(8, 8)  # noqa: E731
            self.attention = torch.nn.Module()
            self.attention.self = torch.nn.Module()
            self.attention.self.query, self.attention.self.key, self.attention.self.value = lin(), lin(), lin()
            self.attention.output = torch.nn.Module()
            self.attention.output.dense = lin()
            self.intermediate = torch.nn.Module()
            self.intermediate.dense = lin()
            self.output = torch.nn.Module()
            self.output.dense = lin()

    model = torch.nn.Sequential(Layer(), Layer())
    w0 = model[0].attention.self.query.weight.detach().clone()
    scales = quantize_transformer_layer(Layer, model)
    q = model[0].attention.self.query.weight
    assert len(scales) == 12 and q.dtype == torch.int8 and (q.float() * scales[0] - w0).abs().max() <= scales[0] / 2 + 1e-7
    env = env_mapping({"OMPI_COMM_WORLD_RANK": "3", "PMI_RANK": "3", "OMPI_COMM_WORLD_LOCAL_RANK": "1"})
    assert env["RANK"] == "3" and env["LOCAL_RANK"] == "1"
    with pytest.raises(EnvironmentError):
        env_mapping({"OMPI_COMM_WORLD_RANK": "3", "PMI_RANK": "2", "SLURM_LOCALID": "0"})
    with pytest.raises(EnvironmentError):
        env_mapping({})
    args = parse_args(["--module", "pkg.train", "--lr", "1"])
    cmd = build_command(args)
    assert cmd[-4:] == ["-m", "pkg.train", "--lr", "1"]


def test_compression_named_quantizers_and_helpers():
    from hcache_deepspeed_b200.compression.helper import (convert_conv1d_to_linear, get_compress_methods, get_module_name,
                                                          module_replacement, recursive_getattr)
    from hcache_deepspeed_b200.compression.utils import AsymQuantizer, BinaryQuantizer, SymQuantizer, TernaryQuantizer
    torch.manual_seed(0)
    x = torch.randn(4, 16, requires_grad=True)
    for Q in (SymQuantizer, AsymQuantizer):
        y = Q.apply(x, 8, None, None, 2)
        assert (y - x).abs().max() < 0.05
        y.sum().backward()
        assert torch.equal(x.grad, torch.ones_like(x))      # straight-through
        x.grad = None
        z = Q.apply(x.detach(), 4, -2.0, 2.0, 1)
        assert z.unique().numel() <= 16
    assert TernaryQuantizer.apply(x.detach(), 2, None, None, 1).unique().numel() <= 3
    assert BinaryQuantizer.apply(x.detach(), 1, None, None, 1).unique().numel() <= 2
    m = torch.nn.Sequential(torch.nn.Linear(4, 4), torch.nn.ReLU(), torch.nn.Linear(4, 2))
    w0 = m[0].weight.detach().clone()
    new = module_replacement(m, "0")
    assert type(new).__name__ == "LinearLayer_Compress" and torch.equal(recursive_getattr(m, "0").weight, w0)
    names, seen = get_module_name("g", m, r"^2$", set())
    assert names == ["2"] and seen == {"2"}
    methods = get_compress_methods(m, {"weight_quantization": {"shared_parameters": {"enabled": True, "quantizer_kernel": False},
                                                                "different_groups": {"wq1": {"params": {"target_bits": 4}, "modules": ["0"]}}}})
    assert methods[0][0] == ["0"] and methods[0][2]["weight_quantization"]["target_bits"] == 4

    class Conv1D(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.weight, self.bias = torch.nn.Parameter(torch.randn(6, 3)), torch.nn.Parameter(torch.randn(3))

        def forward(self, x):
            return x @ self.weight + self.bias

    net = torch.nn.Sequential(Conv1D())
    inp = torch.randn(2, 6)
    want = net(inp)
    convert_conv1d_to_linear(net, Conv1D)
    assert isinstance(net[0], torch.nn.Linear) and torch.allclose(net(inp), want, atol=1e-6)


def test_small_reference_named_modules():
    import hcache_deepspeed_b200.inference.v2.inference_utils as iu
    import hcache_deepspeed_b200.ops.fp_quantizer.fp8_gemm as fg
    import hcache_deepspeed_b200.runtime.tensor_parallel.config as tc
    import hcache_deepspeed_b200.utils.config as uc
    from hcache_deepspeed_b200.inference.v2.logging import inference_logger
    from hcache_deepspeed_b200.ops.fp_quantizer import FP_Quantize
    from hcache_deepspeed_b200.runtime.activation_checkpointing.config import get_activation_checkpointing_config
    from hcache_deepspeed_b200.runtime.zero.test import test1
    test1()
    assert iu.is_gated(iu.ActivationType.SiGLU) and not iu.is_gated(iu.ActivationType.GELU)
    assert iu.native_act_id(iu.ActivationType.GEGLU) == 1 and iu.ceil_div(7, 2) == 4 and iu.elem_size(torch.bfloat16) == 2
    cfg = tc.get_tensor_parallel_config({"tensor_parallel": {"autotp_size": 2, "tp_grain_size": 64}})
    assert cfg.autotp_size == 2 and cfg.tensor_parallel.tp_grain_size == 64
    assert uc.get_timers_config({"timers": {"throughput": {"enabled": False}}}).enabled is False
    assert get_activation_checkpointing_config({"activation_checkpointing": {"cpu_checkpointing": True}}).cpu_checkpointing
    assert inference_logger() is inference_logger()
    torch.manual_seed(0)
    q = FP_Quantize(group_size=64)
    w = torch.randn(64, 32, dtype=torch.bfloat16)
    codes, scales = q.quantize(w, return_meta_tensor=True)
    x = torch.randn(4, 64, dtype=torch.bfloat16)
    y = fg.matmul_fp8(x, codes, scales, 64, q)
    assert y.shape == (4, 32) and (y.float() - x.float() @ w.float()).abs().max() < 1.5


def test_nvme_benchmark_scripts(tmp_path, capsys):
    from hcache_deepspeed_b200.nvme import perf_generate_param, perf_run_sweep  # noqa: F401  (paths kept through the alias table)
    from hcache_deepspeed_b200.nvme.test_ds_aio import main
    from hcache_deepspeed_b200.nvme.test_ds_aio_utils import refine_integer_value
    from hcache_deepspeed_b200.ops.op_builder import AsyncIOBuilder
    assert AsyncIOBuilder().is_compatible() and AsyncIOBuilder().load() is not None
    gbps = main(["--folder", str(tmp_path), "--io_size", "2M", "--block_size", "256K", "--queue_depth", "4", "--read",
                 "--loops", "1", "--no_direct"])
    assert gbps > 0 and "read:" in capsys.readouterr().out
    assert refine_integer_value("4K") == 4096 and callable(perf_run_sweep.sweep_main)


def test_aio_kernel_queue_depth(tmp_path):
    """The async file engine keeps queue_depth pieces in flight per worker through kernel AIO (io_submit) when the
    filesystem takes O_DIRECT, runs unaligned heads / tails synchronously, and returns the bytes either way."""
    import torch
    from hcache_deepspeed_b200.io.aio import AsyncIOHandle
    h = AsyncIOHandle(block_size=256 << 10, queue_depth=8, intra_op_parallelism=2, use_direct=True)
    n = 8 << 20
    raw = torch.empty(n + 8192, dtype=torch.uint8)
    off = (-raw.data_ptr()) % 4096
    a = raw[off:off + n]
    a.copy_(torch.randint(0, 255, (n, ), dtype=torch.uint8))
    path = str(tmp_path / "blob.bin")
    assert h.sync_pwrite(a, path) == n
    raw2 = torch.empty(n + 8192, dtype=torch.uint8)
    off2 = (-raw2.data_ptr()) % 4096
    b = raw2[off2:off2 + n]
    r1, r2 = h.async_pread(b[:n // 2], path, 0), h.async_pread(b[n // 2:], path, n // 2)
    assert h.wait(r1) == n // 2 and h.wait(r2) == n // 2
    assert torch.equal(a, b)
    st = h.stats()
    assert st["kernel_aio_pieces"] + st["sync_pieces"] == 2 * (n // (256 << 10))
    if st["kernel_aio_pieces"]:
        assert st["max_inflight_per_worker"] > 1          # real queue depth, not one blocking call per thread
    # an unaligned window (offset 100, odd length) goes through the synchronous path and still round-trips
    c = torch.empty(5000, dtype=torch.uint8)
    assert h.sync_pread(c, path, 100) == 5000
    assert torch.equal(c, a[100:5100])


def test_wall_clock_timers():
    import time
    from hcache_deepspeed_b200.utils.timer import NoopTimer, SynchronizedWallClockTimer
    timers = SynchronizedWallClockTimer()
    t = timers("fwd")
    assert timers("fwd") is t
    for _ in range(3):
        t.start()
        time.sleep(0.01)
        t.stop()
    with pytest.raises(AssertionError):
        t.stop()
    assert 8.0 < t.mean() < 60.0                 # milliseconds per span
    for _ in range(2):
        t.start()
        time.sleep(0.01)
        t.stop()
    t.start()                                    # reading a running timer pauses and resumes it
    total = t.elapsed(reset=True)
    assert total >= 18.0 and t.running
    t.stop()
    timers.log(["fwd", "missing"], normalizer=2.0)
    assert set(timers.get_mean(["fwd"])) == {"fwd"}
    assert NoopTimer()("x").elapsed() == 0


def test_cli_entry_points_bench_elastic_ssh(tmp_path, capsys):
    """bin/hds_bench (collective sweep through the comm facade, single rank over gloo), bin/hds_elastic (plan of an elastic
    config), bin/hds_ssh (host list parsing)."""
    import importlib.machinery
    import importlib.util
    import json
    import os
    from hcache_deepspeed_b200.benchmarks.communication import bus_factor, main as bench_main
    assert bus_factor("all_reduce", 8) == 2 * 7 / 8 and bus_factor("all_gather", 4) == 0.75 and bus_factor("broadcast", 8) == 1.0
    env = {k: os.environ.get(k) for k in ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "LOCAL_RANK")}
    try:
        assert bench_main(["--op", "all_reduce", "--backend", "gloo", "--minsize", "1024", "--maxsize", "4096", "--trials", "2",
                           "--warmups", "1"]) == 0
    finally:
        for k, v in env.items():
            if v is None:
                os.environ.pop(k, None)
    out = capsys.readouterr().out
    assert "all_reduce" in out and "busbw" in out and out.count("\n") >= 5

    def load(name):
        path = os.path.join(os.path.dirname(os.path.dirname(__file__)), "bin", name)
        loader = importlib.machinery.SourceFileLoader(name, path)
        spec = importlib.util.spec_from_loader(name, loader)
        mod = importlib.util.module_from_spec(spec)
        loader.exec_module(mod)
        return mod

    cfg = tmp_path / "ds.json"
    cfg.write_text(json.dumps({"elasticity": {"enabled": True, "max_train_batch_size": 2000, "micro_batch_sizes": [2, 4, 6],
                                              "min_gpus": 1, "max_gpus": 10000, "min_time": 20, "version": 0.1}}))
    load("hds_elastic").main(["-c", str(cfg), "-w", "8"])
    out = capsys.readouterr().out
    assert "valid GPU counts" in out and "micro-batch per GPU" in out
    hf = tmp_path / "hostfile"
    hf.write_text("# comment\nnode1 slots=8\nnode2 slots=8\n\n")
    assert load("hds_ssh").hosts_of(str(hf)) == ["node1", "node2"]


def test_launcher_environment_discovery(monkeypatch):
    """Rank variables of mpirun / srun / AzureML / SageMaker are mapped to RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*
    without an MPI library (ref: deepspeed/comm/comm.py:705-808)."""
    import json
    import os
    from hcache_deepspeed_b200.comm import comm as C
    from hcache_deepspeed_b200.launcher.multinode_runner import expand_slurm_nodelist

    keys = ["RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "NCCL_SOCKET_IFNAME"]
    keys += [k for row in C._RANK_ENV_SOURCES for k in row]
    keys += ["AZUREML_EXPERIMENT_ID", "AZ_BATCH_MASTER_NODE", "AZ_BATCHAI_MPI_MASTER_NODE", "SM_TRAINING_ENV", "DLTS_JOB_ID",
             "SLURM_JOB_NODELIST", "SLURM_NODELIST"]

    def clean():
        for k in keys:
            monkeypatch.delenv(k, raising=False)

    clean()
    assert C.process_manager_env() is None and not C.env_discovery(verbose=False)

    # OpenMPI, one node: loopback rendezvous
    for k, v in dict(OMPI_COMM_WORLD_RANK="3", OMPI_COMM_WORLD_LOCAL_RANK="3", OMPI_COMM_WORLD_SIZE="4",
                     OMPI_COMM_WORLD_LOCAL_SIZE="4").items():
        monkeypatch.setenv(k, v)
    assert C.process_manager_env() == (3, 3, 4, 4)
    assert C.env_discovery(distributed_port=29611, verbose=False)
    assert [os.environ[k] for k in keys[:5]] == ["3", "3", "4", "127.0.0.1", "29611"]

    # OpenMPI, two nodes, no address known: the caller must fall back to MPI broadcast
    clean()
    for k, v in dict(OMPI_COMM_WORLD_RANK="5", OMPI_COMM_WORLD_LOCAL_RANK="1", OMPI_COMM_WORLD_SIZE="8",
                     OMPI_COMM_WORLD_LOCAL_SIZE="4").items():
        monkeypatch.setenv(k, v)
    assert not C.env_discovery(verbose=False) and "RANK" not in os.environ

    # SLURM with a compact node list
    clean()
    for k, v in dict(SLURM_PROCID="9", SLURM_LOCALID="1", SLURM_NTASKS="16", SLURM_NTASKS_PER_NODE="8(x2)",
                     SLURM_JOB_NODELIST="gpu[07-08,11],login1").items():
        monkeypatch.setenv(k, v)
    assert expand_slurm_nodelist("gpu[07-08,11],login1,a[1-2]b") == ["gpu07", "gpu08", "gpu11", "login1", "a1b", "a2b"]
    assert C.env_discovery(verbose=False)
    assert (os.environ["RANK"], os.environ["LOCAL_RANK"], os.environ["WORLD_SIZE"], os.environ["MASTER_ADDR"]) == \
        ("9", "1", "16", "gpu07")

    # AzureML, several nodes and one node
    clean()
    for k, v in dict(AZUREML_EXPERIMENT_ID="x", OMPI_COMM_WORLD_RANK="4", OMPI_COMM_WORLD_LOCAL_RANK="0",
                     OMPI_COMM_WORLD_SIZE="8", OMPI_COMM_WORLD_LOCAL_SIZE="4", AZ_BATCH_MASTER_NODE="10.0.0.4:6000").items():
        monkeypatch.setenv(k, v)
    assert C.in_aml()
    C.patch_aml_env_for_torch_nccl_backend(verbose=False)
    assert (os.environ["RANK"], os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"]) == ("4", "10.0.0.4", "6105")
    assert os.environ["NCCL_SOCKET_IFNAME"] == C.AML_NCCL_SOCKET_IFNAME
    clean()
    for k, v in dict(AZUREML_EXPERIMENT_ID="x", OMPI_COMM_WORLD_RANK="1", OMPI_COMM_WORLD_LOCAL_RANK="1",
                     OMPI_COMM_WORLD_SIZE="2", OMPI_COMM_WORLD_LOCAL_SIZE="2", AZ_BATCHAI_MPI_MASTER_NODE="10.0.0.9").items():
        monkeypatch.setenv(k, v)
    C.patch_aml_env_for_torch_nccl_backend(verbose=False)
    assert (os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"]) == ("10.0.0.9", str(C.AML_MASTER_PORT))

    # SageMaker
    clean()
    for k, v in dict(SM_TRAINING_ENV=json.dumps({"hosts": ["algo-2", "algo-1"]}), OMPI_COMM_WORLD_RANK="2",
                     OMPI_COMM_WORLD_LOCAL_RANK="0", OMPI_COMM_WORLD_SIZE="4").items():
        monkeypatch.setenv(k, v)
    assert C.in_aws_sm()
    C.patch_aws_sm_env_for_torch_nccl_backend(verbose=False)
    assert (os.environ["RANK"], os.environ["LOCAL_RANK"], os.environ["WORLD_SIZE"], os.environ["MASTER_ADDR"]) == \
        ("2", "0", "4", "algo-1")
    clean()


def test_op_builder_probes_jit_and_dependency_tables(tmp_path, monkeypatch):
    """Per-op builders (ref: op_builder/builder.py:523-607): compatibility probes, a JIT build of one host op and one
    CUDA op into a hashed cache, and DEPENDS tables that really close over the symbols their sources use."""
    import ctypes
    from hcache_deepspeed_b200.ops import build
    from hcache_deepspeed_b200.ops.op_builder import (ALL_OPS, OpBuildError, QuantizerBuilder, ShareMemCommBuilder, has_function,
                                                      has_header, nvcc_knows_sm100a, nvcc_version)
    from hcache_deepspeed_b200.ops.op_builder.builder import unresolved_symbols
    build.build_all(verbose=False)
    assert nvcc_version() >= (12, 8) and nvcc_knows_sm100a()
    assert has_header("linux/aio_abi.h") and not has_header("no_such_header_hds.h")
    assert has_function("shm_open", headers=("sys/mman.h", ), libs=("rt", )) and not has_function("hds_no_such_fn")
    for name, b in ALL_OPS.items():
        assert b.is_compatible() and b.installed(), (name, b.error_log)
        assert all(p.exists() for p in b.all_sources()), name
        assert unresolved_symbols(b) == [], (name, unresolved_symbols(b))
    incomplete = type("Broken", (QuantizerBuilder, ), {"NAME": "broken", "SOURCES": ("gemm.cu", ), "DEPENDS": ()})()
    assert any("tp_take_armed" in s for s in unresolved_symbols(incomplete))

    monkeypatch.setenv("HDS_JIT_DIR", str(tmp_path / "jit"))
    monkeypatch.setenv("HDS_FORCE_JIT", "1")
    shm_b = ShareMemCommBuilder()
    h0 = shm_b.build_hash()
    monkeypatch.setenv("HDS_BUILD_OPS", "0")
    with pytest.raises(OpBuildError, match="HDS_BUILD_OPS=0"):
        shm_b.load(verbose=False)
    monkeypatch.delenv("HDS_BUILD_OPS")
    lib = shm_b.load(verbose=False)
    assert shm_b.jit_path().exists() and shm_b.jit_path().name == f"deepspeed_shm_comm-{h0}.so"
    # the JIT-built object works on its own: a one-rank group reduces to itself
    lib.hds_shm_open.restype = ctypes.c_void_p
    lib.hds_shm_open.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_int64]
    lib.hds_shm_allreduce.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int]
    lib.hds_shm_close.argtypes = [ctypes.c_void_p, ctypes.c_int]
    name = f"hds_jit_test_{os.getpid()}".encode()
    g = lib.hds_shm_open(name, 0, 1, 4096)
    assert g
    x = torch.arange(3000, dtype=torch.float32)
    assert lib.hds_shm_allreduce(g, x.data_ptr(), x.numel(), 2) == 0 and float(x[-1]) == 2999.0
    lib.hds_shm_close(g, 1)
    # a different flag set is a different artefact; the old one is dropped
    monkeypatch.setattr(ShareMemCommBuilder, "cxx_args", lambda self: ["-O1", "-std=c++17", "-fPIC", "-pthread"])
    b2 = ShareMemCommBuilder()
    assert b2.build_hash() != h0
    b2.load(verbose=False)
    assert [p.name for p in (tmp_path / "jit").glob("deepspeed_shm_comm-*.so")] == [b2.jit_path().name]
    # one CUDA op cross-compiled for sm_100a on its own
    q = QuantizerBuilder()
    qlib = q.load(verbose=False)
    assert hasattr(qlib, "hds_quantize_int") and not hasattr(qlib, "hds_gemm")
    sass = subprocess.run(["cuobjdump", "-lelf", str(q.jit_path())], stdout=subprocess.PIPE, text=True).stdout
    assert "sm_100a" in sass
    # a missing tool chain is reported, and a pre-built library still makes the op usable
    monkeypatch.delenv("HDS_FORCE_JIT")
    monkeypatch.setenv("NVCC", "/nonexistent/nvcc")
    q2 = QuantizerBuilder()
    assert q2.is_compatible() and any("nvcc not found" in w for w in q2.error_log)
    with pytest.raises(OpBuildError, match="nvcc not found"):
        q2.jit_build(verbose=False)


def test_launcher_end_to_end_from_another_directory(tmp_path):
    """``bin/hds`` from a source checkout that is not installed, started in an unrelated directory: two local ranks come
    up (hostfile + --include), see their LOCAL_RANK, all-reduce over gloo; a failing rank's exit code becomes the
    launcher's and its siblings are stopped (ref: launcher/runner.py + launch.py)."""
    import sys
    import time
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    probe = tmp_path / "probe.py"
    probe.write_text(
        "import argparse, os, torch\n"
        "import hcache_deepspeed_b200 as ds\n"
        "p = argparse.ArgumentParser(); p.add_argument('--local_rank', type=int, default=-1); p.add_argument('--tag'); a = p.parse_args()\n"
        "ds.init_distributed(dist_backend='gloo')\n"
        "t = torch.ones(1) * (ds.comm.get_rank() + 1); ds.comm.all_reduce(t)\n"
        "print(f'RESULT {ds.comm.get_rank()} {ds.comm.get_world_size()} {a.local_rank} {os.environ[\"LOCAL_RANK\"]} {float(t)} {a.tag}', flush=True)\n")
    (tmp_path / "hostfile").write_text("localhost slots=4\n")
    env = {k: v for k, v in os.environ.items() if k not in ("PYTHONPATH", "RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    env["CUDA_VISIBLE_DEVICES"] = ""
    from tests.dist_utils import _free_port
    out = subprocess.run([sys.executable, os.path.join(repo, "bin", "hds"), "-H", str(tmp_path / "hostfile"), "--include", "localhost:1,3",
                          "--master_port", str(_free_port()), str(probe), "--tag", "hello"], cwd=str(tmp_path), env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    lines = sorted(l for l in out.stdout.splitlines() if l.startswith("RESULT"))
    assert out.returncode == 0 and lines == ["RESULT 0 2 0 0 3.0 hello", "RESULT 1 2 1 1 3.0 hello"], out.stdout[-2000:]
    bad = tmp_path / "bad.py"
    bad.write_text("import os, sys, time\nif os.environ['RANK'] == '1':\n    sys.exit(7)\ntime.sleep(120)\n")
    t0 = time.time()
    out = subprocess.run([sys.executable, os.path.join(repo, "bin", "hds"), "--num_gpus", "2", "--master_port", str(_free_port()), str(bad)],
                         cwd=str(tmp_path), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert out.returncode == 7 and time.time() - t0 < 90, (out.returncode, out.stdout[-1500:])


def test_pip_install_into_target_and_use_from_there(tmp_path):
    """``pip install --no-build-isolation .`` (setup.py: builds the in-tree libraries, ships them + the CUDA / C++ sources +
    the public headers, installs the CLI scripts); the installed copy works from an unrelated directory, including a JIT
    build of one op against the packaged headers."""
    import shutil
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "src"                       # build from a copy: pip writes build/ and egg-info next to setup.py
    shutil.copytree(repo, src, ignore=shutil.ignore_patterns(".git", "gpurun_out", "baseline", "profiles", "tests", "__pycache__",
                                                             "obj", "jit", "build", "*.egg-info"))
    target = tmp_path / "site"
    res = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--target",
                          str(target), str(src)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1200)
    assert res.returncode == 0, res.stdout[-3000:]
    pkg = target / "hcache_deepspeed_b200"
    assert (pkg / "ops" / "lib" / "libhds_host.so").exists() and (pkg / "ops" / "lib" / "libhds_b200.so").exists()
    assert (pkg / "include" / "hds_b200.h").exists() and (pkg / "ops" / "csrc" / "gemm.cu").exists()
    assert (target / "bin" / "hds").exists() and (target / "bin" / "hds_report").exists()
    code = ("import os, torch, hcache_deepspeed_b200 as ds\n"
            "from hcache_deepspeed_b200.ops import build\n"
            "from hcache_deepspeed_b200.ops.op_builder import ALL_OPS, ShareMemCommBuilder\n"
            "assert ds.__file__.startswith(os.environ['HDS_T']), ds.__file__\n"
            "assert str(build.REPO_INCLUDE).startswith(os.environ['HDS_T'])\n"
            "assert all(b.installed() for b in ALL_OPS.values())\n"
            "os.environ['HDS_FORCE_JIT'] = '1'\n"
            "assert hasattr(ShareMemCommBuilder().load(verbose=False), 'hds_shm_barrier')\n"
            "print('INSTALLED-OK')\n")
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    env.update(PYTHONPATH=str(target), HDS_T=str(target), HDS_JIT_DIR=str(tmp_path / "jit"))
    out = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                         text=True, timeout=600)
    assert "INSTALLED-OK" in out.stdout, out.stdout[-3000:]


def test_comms_logger_enabled_from_training_config():
    """``"comms_logger": {"enabled": true}`` in the training config must switch the logger on (the engine passes its config
    object to ``comm.configure``); both config shapes are understood."""
    from hcache_deepspeed_b200.comm import comm as C
    from hcache_deepspeed_b200.comm.config import DeepSpeedCommsConfig
    from hcache_deepspeed_b200.runtime.config import DeepSpeedConfig
    saved = (C.comms_logger.enabled, C.comms_logger.prof_all, list(C.comms_logger.prof_ops), C.comms_logger.verbose)
    try:
        cfg = DeepSpeedConfig({"train_batch_size": 1, "comms_logger": {"enabled": True, "prof_all": False, "prof_ops": ["all_reduce"]}})
        C.comms_logger.enabled = False
        C.configure(cfg)
        assert C.comms_logger.enabled and not C.comms_logger.prof_all and C.comms_logger.prof_ops == ["all_reduce"]
        C.configure(DeepSpeedConfig({"train_batch_size": 1, "comms_logger": {"enabled": False}}))
        assert not C.comms_logger.enabled
        wrapper = types.SimpleNamespace(comms_config=DeepSpeedCommsConfig({"comms_logger": {"enabled": True, "verbose": True}}))
        C.configure(wrapper)
        assert C.comms_logger.enabled and C.comms_logger.verbose
        # a timed op is recorded once the logger is on (single process: the size / latency bookkeeping is what is checked)
        C.comms_logger.comms_dict.clear()
        C.comms_logger.configure(enabled=True, prof_all=True)
        C.comms_logger.append("all_reduce", "all_reduce", 1.5, 4096, 2)
        assert "all_reduce" in C.comms_logger.log_all(print_log=False)
    finally:
        C.comms_logger.enabled, C.comms_logger.prof_all, C.comms_logger.prof_ops, C.comms_logger.verbose = saved
        C.comms_logger.comms_dict.clear()


def test_flops_profiler_counts_and_engine_report(tmp_path):
    """FLOPs profiler (ref: profiling/flops_profiler/profiler.py): exact MAC counts for linear / einsum / baddbmm / conv /
    attention, restored functionals afterwards, and the training-engine report (summary with bwd / step latency,
    per-depth top modules, detailed tree with percentages) written at the end of the profiled step."""
    import torch.nn.functional as F
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.profiling.flops_profiler.profiler import FlopsProfiler, get_model_profile

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.conv = torch.nn.Linear(16, 32), torch.nn.Conv1d(4, 6, 3, padding=1)

        def forward(self, x, y):
            h = self.a(x)                                                  # 5*16*32 MACs
            e = torch.einsum("bi,ci->bc", h, h)                            # 5*5*32
            bb = torch.baddbmm(torch.zeros(2, 3, 3), torch.ones(2, 3, 4), torch.ones(2, 4, 3))     # 2*3*4*3
            c = self.conv(y)                                               # out elems (2*6*8) * (4*3)
            q = torch.ones(1, 2, 5, 8)
            s = F.scaled_dot_product_attention(q, q, q)                    # 2 * 5*5*(8+8)
            return h.sum() + e.sum() + bb.sum() + c.sum() + s.sum()

    net = Net()
    orig_linear, orig_einsum = F.linear, torch.einsum
    flops, macs, params = get_model_profile(net, args=(torch.ones(5, 16), torch.ones(2, 4, 8)), print_profile=False, as_string=False)
    want = 5 * 16 * 32 + 5 * 5 * 32 + 2 * 3 * 4 * 3 + 2 * 6 * 8 * 4 * 3 + 2 * 5 * 5 * 16
    assert macs == want and flops >= 2 * want and params == sum(p.numel() for p in net.parameters())
    assert F.linear is orig_linear and torch.einsum is orig_einsum                      # patches removed

    out_file = tmp_path / "flops.txt"
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.GELU(), torch.nn.Linear(16, 1))
    cfg = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.01}},
           "wall_clock_breakdown": True, "steps_per_print": 100,
           "flops_profiler": {"enabled": True, "profile_step": 1, "top_modules": 2, "detailed": True, "output_file": str(out_file)}}
    engine, _, _, _ = ds.initialize(model=model, config=cfg)
    for i in range(6):
        loss = engine(torch.randn(4, 8)).pow(2).mean()
        engine.backward(loss)
        engine.step()
        if i < 3:
            assert not out_file.exists()          # written at the END of global step 1 (micro-steps 2 and 3), not before
    text = out_file.read_text()
    for needle in ("Profile Summary at step 1", "data parallel size", "bwd latency", "step latency", "iter latency", "samples/second",
                   "Aggregated Profile per GPU", "depth 1:", "Detailed Profile per GPU", "(0): Linear(", "% MACs", "fwd MACs:             576.00 MACs"):
        assert needle in text, needle
    assert not engine.flops_profiler.started


def test_timer_elapsed_without_reset_keeps_the_total():
    import time as _t
    from hcache_deepspeed_b200.utils.timer import SynchronizedWallClockTimer
    timers = SynchronizedWallClockTimer()
    t = timers("x")
    for _ in range(2):
        t.start()
        _t.sleep(0.01)
        t.stop()
    first = t.elapsed(reset=False)
    assert first >= 15 and abs(t.elapsed(reset=False) - first) < 1e-6 and abs(t.mean() - first / 2) < 1e-6
    assert t.elapsed(reset=True) == first and t.elapsed() == 0.0
    assert timers.has_timer("x") and not timers.has_timer("y")


def test_launcher_autotuning_tune_end_to_end(tmp_path):
    """``hds --autotuning tune script.py --deepspeed_config cfg.json``: a model-info run, then one short run of the script
    per (ZeRO stage, micro-batch size) candidate through the launcher; each run's engine writes its metrics file and
    exits; the best config lands in ``ds_config_optimal.json`` (ref: launcher/runner.py run_autotuning, autotuner.py)."""
    import json
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "train.py"
    script.write_text(
        "import argparse, torch\n"
        "import hcache_deepspeed_b200 as ds\n"
        "p = argparse.ArgumentParser(); p.add_argument('--local_rank', type=int, default=-1); p.add_argument('--deepspeed_config')\n"
        "a = p.parse_args()\n"
        "ds.init_distributed(dist_backend='gloo')\n"
        "torch.manual_seed(0)\n"
        "model = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.ReLU(), torch.nn.Linear(64, 1))\n"
        "engine, _, _, _ = ds.initialize(model=model, config=a.deepspeed_config)\n"
        "mbs = engine.train_micro_batch_size_per_gpu()\n"
        "for _ in range(50):\n"
        "    loss = engine(torch.randn(mbs, 32)).pow(2).mean(); engine.backward(loss); engine.step()\n"
        "print('TRAINING-FINISHED')\n")
    results = tmp_path / "results"
    cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}}, "steps_per_print": 1000,
           "autotuning": {"enabled": True, "results_dir": str(results), "exps_dir": str(tmp_path / "exps"), "fast": True,
                          "start_profile_step": 2, "end_profile_step": 4, "micro_batch_sizes": [2, 8], "tuner_num_trials": 4,
                          "tuner_early_stopping": 10},
           "zero_optimization": {"stage": 1}}
    cfg_file = tmp_path / "ds.json"
    cfg_file.write_text(json.dumps(cfg))
    env = {k: v for k, v in os.environ.items() if k not in ("PYTHONPATH", "RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    env["CUDA_VISIBLE_DEVICES"] = ""
    from tests.dist_utils import _free_port
    out = subprocess.run([sys.executable, os.path.join(repo, "bin", "hds"), "--num_gpus", "1", "--master_port", str(_free_port()),
                          "--autotuning", "tune", str(script), "--deepspeed_config", str(cfg_file)], cwd=str(tmp_path), env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:]
    assert "TRAINING-FINISHED" not in out.stdout                      # "tune" does not run the job itself
    info = json.loads((results / "model_info.json").read_text())
    assert info["num_params"] == 32 * 64 + 64 + 64 + 1
    metrics = sorted(results.glob("exp_*_metrics.json"))
    assert len(metrics) == 2                                           # stage 1 x micro-batch sizes {2, 8}
    vals = {json.loads(m.read_text())["train_micro_batch_size_per_gpu"]: json.loads(m.read_text())["throughput"] for m in metrics}
    assert set(vals) == {2, 8} and all(v > 0 for v in vals.values())
    best = json.loads((results / "ds_config_optimal.json").read_text())
    assert best["train_micro_batch_size_per_gpu"] == max(vals, key=vals.get) and best["zero_optimization"]["stage"] == 1
    assert "autotuning: best config written to" in out.stdout


def test_elasticity_block_drives_the_batch_size_triple():
    """``"elasticity": {"enabled": true}`` in the training config: global batch from the elastic plan, micro batch and
    accumulation steps follow from the world size; explicit batch sizes are refused (ref: runtime/config.py:739-790)."""
    from hcache_deepspeed_b200.runtime.config import DeepSpeedConfig, DeepSpeedConfigError
    el = {"enabled": True, "max_train_batch_size": 2000, "micro_batch_sizes": [2, 4, 6], "min_gpus": 1, "max_gpus": 10000,
          "min_time": 20, "version": 0.1}
    seen = {}
    for ws in (1, 4, 7, 8):
        c = DeepSpeedConfig({"elasticity": el}, world_size=ws)
        assert c.train_batch_size == 1680 == c.train_micro_batch_size_per_gpu * c.gradient_accumulation_steps * ws
        assert ws in c.elastic_plan["valid_gpus"]
        seen[ws] = (c.train_micro_batch_size_per_gpu, c.gradient_accumulation_steps)
    assert seen[4] == (6, 70) and seen[8] == (6, 35)
    with pytest.raises(DeepSpeedConfigError, match="must not be set"):
        DeepSpeedConfig({"elasticity": el, "train_batch_size": 8}, world_size=2)
    c = DeepSpeedConfig({"elasticity": dict(el, ignore_non_elastic_batch_info=True), "train_batch_size": 8}, world_size=2)
    assert c.train_batch_size == 1680


def test_launcher_node_config_and_num_nodes_parsing():
    from hcache_deepspeed_b200.launcher.runner import parse_node_config, parse_node_config_list, parse_num_nodes
    assert parse_node_config("w0") == ("w0", []) and parse_node_config("w1:0,2") == ("w1", [0, 2])
    assert parse_node_config_list("w1:2,0@w0@w1:3,0") == {"w1": [0, 2, 3], "w0": []}
    assert parse_num_nodes("4", False) == (4, -1) and parse_num_nodes("2:8", True) == (2, 8) and parse_num_nodes(-1, False) == (-1, -1)
    with pytest.raises(RuntimeError):
        parse_num_nodes("2:8", False)
    with pytest.raises(RuntimeError):
        parse_num_nodes("1:2:3", True)
