"""MoE (gating + expert parallelism) and sequence parallelism (Ulysses, FPDT) on CPU ranks (gloo).
Mirrors tests/unit/moe/test_moe.py and tests/unit/sequence_parallelism/test_ulysses.py of the reference."""
import math

import pytest
import torch

from tests.dist_utils import run_distributed
from tests.parallel_helper import fpdt_worker, moe_worker, ulysses_worker

pytestmark = pytest.mark.timeout(600)


def test_gating_functions():
    from hcache_deepspeed_b200.moe.sharded_moe import top1gating, top2gating, topkgating
    torch.manual_seed(0)
    S, E = 64, 4
    logits = torch.randn(S, E)
    l, combine, dispatch, counts = top1gating(logits, capacity_factor=1.0, min_capacity=4, use_rts=False)
    C = combine.shape[2]
    assert C == 16 and dispatch.shape == (S, E, C) and int(counts.sum()) == S
    assert int(dispatch.sum(dim=(0, 2)).max()) <= C          # capacity respected
    assert int(dispatch.sum(dim=(1, 2)).max()) <= 1          # a token goes to at most one slot
    assert float(l) > 0
    l, combine, dispatch, counts = top2gating(logits, capacity_factor=2.0, min_capacity=4, top2_2nd_expert_sampling=False)
    per_tok = dispatch.sum(dim=(1, 2))
    assert int(per_tok.max()) <= 2 and int(counts.sum()) == 2 * S
    kept = per_tok == 2
    torch.testing.assert_close(combine.sum(dim=(1, 2))[kept], torch.ones(int(kept.sum())))   # weights renormalised
    l, combine, dispatch, counts = topkgating(logits, 3, capacity_factor=4.0, min_capacity=4)
    assert int(dispatch.sum(dim=(1, 2)).max()) == 3 and int(dispatch.sum(dim=(1, 2)).min()) == 3
    # each (expert, slot) holds at most one token
    assert int(dispatch.sum(dim=0).max()) <= 1


def test_moe_param_groups():
    from hcache_deepspeed_b200.moe.layer import MoE
    from hcache_deepspeed_b200.moe.utils import (has_moe_layers, is_moe_param,
                                                 split_params_into_different_moe_groups_for_optimizer)
    m = torch.nn.Sequential(torch.nn.Linear(8, 8), MoE(8, torch.nn.Linear(8, 8), num_experts=2, ep_size=1))
    assert has_moe_layers(m) == (True, 2)
    groups_ = split_params_into_different_moe_groups_for_optimizer({"params": list(m.parameters()), "lr": 0.1})
    moe_g = [g for g in groups_ if g.get("moe")]
    assert len(moe_g) == 1 and moe_g[0]["name"] == "ep_size_1" and all(is_moe_param(p) for p in moe_g[0]["params"])
    assert not any(is_moe_param(p) for p in groups_[0]["params"]) and groups_[0]["lr"] == 0.1


@pytest.mark.parametrize("world,ep,k", [(2, 2, 1), (4, 2, 2)])
def test_expert_parallel_matches_local(world, ep, k):
    for r in run_distributed(moe_worker, world, ep, k):
        assert r["err"] < 1e-5 and r["gerr"] < 1e-5 and r["is_moe_param"]
        assert r["laux"] == pytest.approx(r["laux_ref"], rel=1e-5)


@pytest.mark.parametrize("world,heads", [(2, 4), (4, 4), (2, 5)])
def test_ulysses_matches_full_attention(world, heads):
    for r in run_distributed(ulysses_worker, world, heads):
        assert r["err"] < 1e-5 and r["gerr"] < 1e-5


@pytest.mark.parametrize("world,chunks,offload", [(2, 2, False), (2, 3, True)])
def test_fpdt_chunked_attention(world, chunks, offload):
    for r in run_distributed(fpdt_worker, world, chunks, offload):
        assert r["err"] < 1e-5


def test_update_out_and_lse_merges_blocks():
    from hcache_deepspeed_b200.sequence.fpdt_layer import _attn_with_lse, update_out_and_lse
    torch.manual_seed(0)
    q, k, v = (torch.randn(1, 6, 2, 8) for _ in range(3))
    full, _ = _attn_with_lse(q, k, v, None, 0.3)
    o, l = None, None
    for ks, vs in zip(k.chunk(3, dim=1), v.chunk(3, dim=1)):
        bo, bl = _attn_with_lse(q, ks, vs, None, 0.3)
        o, l = update_out_and_lse(o, l, bo, bl)
    torch.testing.assert_close(o, full.float(), rtol=1e-5, atol=1e-6)


def test_sequence_parallel_cross_entropy_single_rank():
    from hcache_deepspeed_b200.sequence.cross_entropy import vocab_sequence_parallel_cross_entropy
    torch.manual_seed(0)
    logits = torch.randn(5, 2, 11, requires_grad=True)
    tgt = torch.randint(0, 11, (5, 2))
    loss = vocab_sequence_parallel_cross_entropy(logits, tgt, None)
    want = torch.nn.functional.cross_entropy(logits.reshape(-1, 11), tgt.reshape(-1), reduction="none").reshape(5, 2)
    torch.testing.assert_close(loss, want, rtol=1e-5, atol=1e-6)
    loss.sum().backward()
    l2 = logits.detach().clone().requires_grad_(True)
    torch.nn.functional.cross_entropy(l2.reshape(-1, 11), tgt.reshape(-1), reduction="sum").backward()
    torch.testing.assert_close(logits.grad, l2.grad, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("world", [2, 3])
def test_shm_allreduce_matches_gloo(world):
    import os
    from tests.parallel_helper import shm_allreduce_worker
    res = run_distributed(shm_allreduce_worker, world)
    for r in res:
        assert r["torch.float32_7"] < 1e-6 and r["torch.float32_5000"] < 1e-5
        assert r["torch.bfloat16_3001"] < 0.07        # one bf16 rounding of an fp32 sum vs gloo's pairwise bf16 adds
        assert r["int_fallback"] == world
    assert len({r["name"] for r in res}) == 1 and not os.path.exists("/dev/shm/" + res[0]["name"])


def _torch_backend_worker(rank, world):
    import torch
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.comm import comm as c
    from hcache_deepspeed_b200.comm.torch import TorchBackend
    from hcache_deepspeed_b200.comm.utils import (get_local_rank_from_launcher, get_msg_size_from_args,
                                                  get_world_rank_from_launcher, get_world_size_from_launcher)
    ds.init_distributed(dist_backend="gloo")
    b = TorchBackend("gloo")
    t = torch.full((4,), float(rank + 1))
    b.all_reduce(t)
    out = torch.empty(2 * world)
    b.all_gather_into_tensor(out, torch.full((2,), float(rank)))
    assert b.is_initialized() and b.get_backend() == "gloo" and b.has_all_gather_into_tensor() and b.has_reduce_scatter_tensor()
    assert get_world_rank_from_launcher() == rank and get_world_size_from_launcher() == world and get_local_rank_from_launcher() == rank
    return {"sum": t.tolist(), "gathered": out.tolist(), "size": b.get_world_size(),
            "msg": get_msg_size_from_args(c.all_reduce, t)}


def test_torch_backend_object_2ranks():
    res = run_distributed(_torch_backend_worker, 2)
    for r in res:
        assert r["sum"] == [3.0] * 4 and r["gathered"] == [0.0, 0.0, 1.0, 1.0] and r["size"] == 2 and r["msg"] == 16


@pytest.mark.parametrize("world,simd", [(2, "on"), (3, "on"), (4, "off")])
def test_ccl_backend_host_collectives(world, simd, monkeypatch):
    """The host back end (``dist_backend="ccl"``): all-reduce sum/max/min over five dtypes, broadcast, all-gather,
    reduce-scatter and barrier through the C++ shared-memory engine (both all-reduce shapes, AVX-512 and portable
    kernels) against gloo; uncovered calls fall back to torch.distributed."""
    import os
    from tests.parallel_helper import ccl_backend_worker
    monkeypatch.setenv("HDS_SHM_SIMD", simd)
    res = run_distributed(ccl_backend_worker, world)
    for r in res:
        calls = r["calls"]
        assert calls["all_reduce.native"] >= 60 and calls["all_reduce.torch"] == 3, calls
        assert calls["broadcast.native"] == 2 and calls["reduce_scatter_tensor.native"] == 4 and calls["barrier.native"] == 1, calls
        if simd == "off":
            assert r["avx512"] == 0
    assert not os.path.exists("/dev/shm/" + res[0]["name"])


def test_ulysses_training_through_the_engine_matches_full_sequence():
    from tests.parallel_helper import ulysses_engine_worker
    for r in run_distributed(ulysses_engine_worker, 2):
        assert r["sp"] == 2 and r["got"] == pytest.approx(r["want"], rel=1e-5, abs=1e-5), r


def test_moe_wall_clock_breakdown_counters_and_engine_report():
    """``wall_clock_breakdown`` switches on the MoE layer's gate / all-to-all / total timers, and
    ``engine.print_forward_breakdown`` sums them over the model's MoE layers (ref: sharded_moe.py:590-650, engine.py:2050)."""
    import hcache_deepspeed_b200 as ds
    from hcache_deepspeed_b200.moe.layer import MoE

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a = MoE(8, torch.nn.Linear(8, 8), num_experts=2, ep_size=1, min_capacity=4)
            self.b = MoE(8, torch.nn.Linear(8, 8), num_experts=2, ep_size=1, min_capacity=4)

        def forward(self, x):
            return self.b(self.a(x)[0])[0].pow(2).mean()

    engine, _, _, _ = ds.initialize(model=Net(), config={"train_micro_batch_size_per_gpu": 2, "wall_clock_breakdown": True,
                                                         "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "steps_per_print": 1000})
    layers = [engine.module.a.deepspeed_moe, engine.module.b.deepspeed_moe]
    assert all(l.wall_clock_breakdown for l in layers) and all(l.time_moe == 0.0 for l in layers)
    engine(torch.randn(2, 4, 8))
    assert all(l.time_moe > 0.0 and 0.0 < l.gate_time <= l.time_moe for l in layers)
    rep = engine.print_forward_breakdown(fwd_time=12.5)
    assert rep["fwd"] == 12.5 and rep["fwd_moe"] == pytest.approx(sum(l.time_moe for l in layers))
    assert rep["top_k"] == pytest.approx(sum(l.gate_time for l in layers)) and rep["1st_a2a"] >= 0.0


def test_comm_off_switches_2ranks():
    from tests.parallel_helper import comm_off_worker
    for rank, r in enumerate(run_distributed(comm_off_worker, 2)):
        assert r["noop"] == "Noop" and r["wait"] is None and r["skipped"] == [rank + 1.0] * 2 and r["reduced"] == [3.0, 3.0], r
        assert r["flags"] == (True, True) and r["gather_noop"] == "Noop" and r["probes"][0] is True, r


def test_model_parallel_compress_layers_2ranks():
    from tests.parallel_helper import mp_compress_worker
    for rank, r in enumerate(run_distributed(mp_compress_worker, 2)):
        assert r["fwd"] < 1e-5 and r["grad"] < 1e-4 and r["gathered_shape"] == (3, 8) and r["bias_back"], r
        assert r["scatter"] == [[4.0 * rank + i for i in range(4)]] and r["gather"] == [[float(i) for i in range(8)]], r
