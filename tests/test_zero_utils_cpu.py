"""ZeRO support modules: contiguous allocator with compaction, memory-lean linear, offload-state helpers, utils."""
from collections import namedtuple

import pytest
import torch

from hcache_deepspeed_b200.runtime.zero.contiguous_memory_allocator import ContiguousMemoryAllocator
from hcache_deepspeed_b200.runtime.zero.linear import LinearModuleForZeroStage3, zero3_linear_wrap
from hcache_deepspeed_b200.runtime.zero.offload_states import (OffloadStateTypeEnum, get_state_devices, offload_adam_states,
                                                              reload_adam_states)
from hcache_deepspeed_b200.runtime.zero.partitioned_param_profiler import PartitionedParameterProfiler
from hcache_deepspeed_b200.runtime.zero.utils import (apply_to_tensors_only, assert_ints_same_as_other_ranks,
                                                     get_mapping_to_flat_buffer, is_zero_supported_optimizer,
                                                     isinstance_namedtuple)


def test_contiguous_allocator_defragments_and_keeps_handles():
    mem = ContiguousMemoryAllocator(1024, torch.float32, "cpu")
    a = mem.allocate_tensor(64)
    b = mem.allocate_tensor(64)
    c = mem.allocate_tensor(128)
    d = mem.allocate_tensor(256)
    e = mem.allocate_tensor(512)
    assert mem.total_free == 0 and mem.max_allocated == 1024
    for i, t in enumerate((a, b, c, d, e)):
        t.fill_(float(i + 1))
    p = torch.nn.Parameter(torch.empty(0))
    mem.assign_to_param(d, p, 200, (10, 20))
    assert p.shape == (10, 20) and float(p.sum()) == 4.0 * 200
    mem.release_tensor(a)
    mem.release_tensor(c)
    assert mem.total_free == 192 and mem.largest_contiguous == 128
    assert mem.contiguous_sizes == {0: 64, 128: 128}
    # 192 free in total but not in one piece: compaction has to move b, d and e
    f = mem.allocate_tensor(192)
    f.fill_(9.0)
    assert mem.total_free == 0
    assert torch.all(b == 2.0) and torch.all(d == 4.0) and torch.all(e == 5.0)
    assert float(p.sum()) == 4.0 * 200
    assert p.data_ptr() == d.data_ptr() and b.data_ptr() == mem.buffer.data_ptr()
    # writes through the handle still land in the parameter after the move
    d.fill_(7.0)
    assert float(p.sum()) == 7.0 * 200
    with pytest.raises(AssertionError):
        mem.allocate_tensor(1)
    mem.release_tensor(e)
    assert mem.largest_contiguous == 512 and "x" in mem.print_allocation(32)


def test_zero3_linear_matches_nn_linear():
    torch.manual_seed(0)
    ref = torch.nn.Linear(24, 12)
    mod = LinearModuleForZeroStage3(24, 12)
    mod.load_state_dict(ref.state_dict())
    x1 = torch.randn(3, 5, 24, requires_grad=True)
    x2 = x1.detach().clone().requires_grad_(True)
    ref(x1).square().sum().backward()
    mod(x2).square().sum().backward()
    assert torch.allclose(x1.grad, x2.grad, atol=1e-5)
    assert torch.allclose(ref.weight.grad, mod.weight.grad, atol=1e-4)
    assert torch.allclose(ref.bias.grad, mod.bias.grad, atol=1e-4)
    y = zero3_linear_wrap(x1.detach(), ref.weight.detach())
    assert torch.allclose(y, x1.detach() @ ref.weight.detach().t())


def test_offload_and_reload_adam_states():
    lin = torch.nn.Linear(8, 8)
    opt = torch.optim.Adam(lin.parameters(), lr=1e-2)
    lin(torch.randn(4, 8)).sum().backward()
    opt.step()
    before = {id(p): opt.state[p]["exp_avg"].clone() for p in lin.parameters()}
    offload_adam_states(opt, "cpu")
    assert get_state_devices(opt, OffloadStateTypeEnum.optim_states) == {torch.device("cpu")}
    reload_adam_states(opt, "cpu")
    for p in lin.parameters():
        assert torch.equal(opt.state[p]["exp_avg"], before[id(p)])
    opt.zero_grad()
    lin(torch.randn(4, 8)).sum().backward()
    opt.step()       # the states are still usable


def test_zero_utils_tree_map_and_flat_map():
    Pair = namedtuple("Pair", "a b")
    tree = {"x": [torch.ones(2), 3], "y": Pair(torch.zeros(1), "s")}
    out = apply_to_tensors_only(lambda t: t + 1, tree)
    assert torch.equal(out["x"][0], torch.full((2,), 2.0)) and out["x"][1] == 3
    assert isinstance_namedtuple(out["y"]) and float(out["y"].a) == 1.0 and out["y"].b == "s"
    m = get_mapping_to_flat_buffer([torch.empty(3), torch.empty(2, 2)])
    assert [(o, n) for _, o, n in m] == [(0, 3), (3, 4)]
    assert is_zero_supported_optimizer(torch.optim.AdamW([torch.nn.Parameter(torch.ones(1))]))
    assert not is_zero_supported_optimizer(object())
    assert_ints_same_as_other_ranks([1, 2, 3])      # single process: a no-op


def test_partitioned_param_profiler_counts():
    prof = PartitionedParameterProfiler()
    for n in (10, 20):
        prof.start_event("fetch")
        prof.stop_event("fetch", n)
    c = prof.event_counters["fetch"]
    assert (c.count, c.num_elem) == (2, 30) and prof.elapsed["fetch"] >= 0
    prof.reset_events()
    assert prof.event_counters["fetch"].count == 0


def test_runtime_name_parity_modules():
    import json

    from hcache_deepspeed_b200.runtime import base_optimizer, compiler, config_utils, constants
    from hcache_deepspeed_b200.runtime.bf16_optimizer import BF16_Optimizer
    from hcache_deepspeed_b200.runtime.zero.stage3 import DeepSpeedZeroOptimizer_Stage3
    assert issubclass(BF16_Optimizer, base_optimizer.ZeROOptimizer)
    assert issubclass(DeepSpeedZeroOptimizer_Stage3, base_optimizer.DeepSpeedOptimizer)
    assert hasattr(BF16_Optimizer, "load_hp_checkpoint_state_from_checkpoint_dir")
    assert constants.TRAIN_BATCH_SIZE == "train_batch_size" and constants.GRADIENT_CLIPPING_DEFAULT == 0.0
    assert constants.defaults()["steps_per_print"] == 10
    text = json.dumps({"bucket": 5e8, "small": 12, "flag": True, "l": [1, 2]}, cls=config_utils.ScientificNotationEncoder)
    assert "5.000000e+08" in text and '"small": 12' in text and "true" in text
    assert config_utils.get_scalar_param({"a": 1}, "b", 7) == 7
    f = compiler.disable(lambda x: x + 1)
    assert f(1) == 2 and not compiler.is_compiling()


def _fake_param(n, persist=False):
    p = torch.nn.Parameter(torch.empty(0))
    p.ds_numel, p.ds_persist = n, persist
    return p


def test_schedule_passes_on_a_synthetic_trace():
    from hcache_deepspeed_b200.compile.list_schedule import peak_live_numel, schedule_gathers
    from hcache_deepspeed_b200.compile.passes.prefetch import schedule_prefetch
    from hcache_deepspeed_b200.compile.passes.selective_gather import selective_gather
    from hcache_deepspeed_b200.compile.profilers import fit_alpha_beta
    from hcache_deepspeed_b200.compile.trace import ModuleTrace, TraceEvent
    a, b, c, d = _fake_param(100), _fake_param(4000), _fake_param(50, persist=True), _fake_param(1000)
    order = [([a], "fwd"), ([b], "fwd"), ([c, d], "fwd"), ([c, d], "bwd"), ([b], "bwd"), ([a], "bwd")]
    plan = [[id(p) for p in step] for step in schedule_prefetch(order, bucket_numel=4500)]
    assert plan[0] == [id(b)]                  # d does not fit into what is left of the bucket after b
    assert plan[1] == [id(d), id(a)]           # c is persistent: never prefetched; (c, d) of bwd is the same d
    assert plan[2] == [id(b), id(a)] and plan[4] == [id(a)] and plan[-1] == []
    assert schedule_prefetch(order, 0) == [[] for _ in order]
    trace = ModuleTrace([TraceEvent(i, f"m{i}", ph, ps) for i, (ps, ph) in enumerate(order)])
    assert trace.uses()[id(a)] == [0, 5]
    # benefit per byte: the small parameter first; the budget stops before the big one
    chosen = selective_gather(trace, budget_bytes=3000, world_size=2)
    assert chosen[0] is a and any(x is d for x in chosen) and not any(x is b for x in chosen)
    assert selective_gather(trace, 10 ** 9, world_size=1) == []
    launch = schedule_gathers(trace, max_live_numel=4200)
    assert any(x is b for x in launch[0]) and not any(x is c for v in launch.values() for x in v)
    assert peak_live_numel(trace, launch) <= 4200 + 4000
    fit = fit_alpha_beta([(1000, 1e-5 + 2000 / 1e9), (100000, 1e-5 + 200000 / 1e9), (10 ** 6, 1e-5 + 2e6 / 1e9)])
    assert fit["alpha_s"] == pytest.approx(1e-5, rel=1e-3) and fit["bandwidth_Bps"] == pytest.approx(1e9, rel=1e-3)


def test_adam_state_mover_wraps_step():
    from hcache_deepspeed_b200.compile.passes.offload_adam_states import move_opt_states
    lin = torch.nn.Linear(4, 4)
    ref = torch.nn.Linear(4, 4)
    ref.load_state_dict(lin.state_dict())
    opt, opt_ref = torch.optim.Adam(lin.parameters(), lr=1e-2), torch.optim.Adam(ref.parameters(), lr=1e-2)
    mover = move_opt_states(opt, "cpu", double_buffer=False)
    for i in range(3):
        x = torch.randn(2, 4)
        for m, o in ((lin, opt), (ref, opt_ref)):
            o.zero_grad()
            m(x).sum().backward()
            o.step()
        assert mover.offloaded
    assert torch.allclose(lin.weight, ref.weight) and mover.bytes_moved > 0


@pytest.mark.parametrize("pipelined", [False, True])
def test_optimizer_swapper_roundtrip(tmp_path, pipelined):
    from hcache_deepspeed_b200.runtime.swap_tensor import PartitionedOptimizerSwapper, PipelinedOptimizerSwapper
    torch.manual_seed(0)
    masters = [torch.randn(n, requires_grad=True) for n in (3000, 1000, 5000)]
    twins = [m.detach().clone().requires_grad_(True) for m in masters]
    opt = torch.optim.Adam([{"params": [m]} for m in masters], lr=1e-2)
    ref = torch.optim.Adam([{"params": [m]} for m in twins], lr=1e-2)
    cls = PipelinedOptimizerSwapper if pipelined else PartitionedOptimizerSwapper
    sw = cls(None, {"block_size": 4096}, str(tmp_path), opt, 5000, "cpu", torch.float32, None, buffer_count=4)
    sw.initialize_parameters(masters, masters)
    assert all(m.numel() == 0 for m in masters)
    groups = opt.param_groups
    for it in range(3):
        grads = [torch.randn(n) for n in (3000, 1000, 5000)]
        for i, m in enumerate(masters):
            sw.swap_in_optimizer_state(m, async_parameter=masters[i + 1] if i + 1 < len(masters) else None)
            assert m.numel() == grads[i].numel()
            m.grad = grads[i].clone()
            opt.param_groups = [groups[i]]
            opt.step()
            m.grad = None
            sw.swap_out_optimizer_state(m, async_swap=True)
            assert m.numel() == 0 and opt.state[m]["exp_avg"].numel() == 0
        opt.param_groups = groups
        for t, g in zip(twins, grads):
            t.grad = g.clone()
        ref.step()
    sw.swap_in_all(masters)
    for m, t in zip(masters, twins):
        torch.testing.assert_close(m.detach(), t.detach())
        torch.testing.assert_close(opt.state[m]["exp_avg_sq"], ref.state[t]["exp_avg_sq"])
    assert sw.stats["bytes_read"] > 0 and sw.swappable_tensor(numel=10 ** 7) and not sw.swappable_tensor(numel=10)
    sw.swap_out_all(masters)
    sw.purge_state()
    assert len(sw._free) == 4


@pytest.mark.parametrize("max_live,expect_prefetch", [(10 ** 9, True), (300, False)])
def test_zero3_trace_prefetch_and_live_cap(max_live, expect_prefetch):
    """Second iteration on: the next modules' gathers are launched ahead — unless ``max_live_parameters`` forbids it."""
    from hcache_deepspeed_b200.runtime.zero.stage3 import DeepSpeedZeRoOffload
    torch.manual_seed(0)
    ref = torch.nn.Sequential(torch.nn.Linear(16, 16), torch.nn.Tanh(), torch.nn.Linear(16, 16), torch.nn.Tanh(), torch.nn.Linear(16, 4))
    net = torch.nn.Sequential(torch.nn.Linear(16, 16), torch.nn.Tanh(), torch.nn.Linear(16, 16), torch.nn.Tanh(), torch.nn.Linear(16, 4))
    net.load_state_dict(ref.state_dict())
    off = DeepSpeedZeRoOffload(net, prefetch_bucket_size=10 ** 6, max_live_parameters=max_live, param_persistence_threshold=0)
    grads = {}

    def take(ps):          # what the optimizer's reduce does: consume the gradient
        for p in ps:
            grads[id(p)] = p.grad.clone()
            p.grad = None

    off.reduce_fn = take
    x = torch.randn(3, 16)
    for it in range(3):
        out = net(x)
        out.sum().backward()
        assert all(p.numel() == 0 for p in net.parameters())          # everything released again
    assert off.trace_complete and len(off.trace) == 6 and [ph for _, ph in off.trace] == ["fwd"] * 3 + ["bwd"] * 3
    assert (off.stats["prefetched"] > 0) == expect_prefetch and off.stats["trace_invalidations"] == 0
    torch.testing.assert_close(out, ref(x))
    ref(x).sum().backward()
    for p, q in zip(net.parameters(), ref.parameters()):
        torch.testing.assert_close(grads[id(p)], q.grad)
    # a different module order drops the trace instead of prefetching the wrong parameters
    net[4](torch.tanh(net[0](x)))
    assert off.stats["trace_invalidations"] == 1 and not off.trace_complete
