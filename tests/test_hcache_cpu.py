"""HCache subsystem on CPU: planner properties, on-disk format, DRAM->NVMe tiering through the native aio engine,
mixed restore plans (recompute prefix + hidden states + KV pages) == live KV, multi-turn session manager."""
import os

import pytest
import torch

from hcache_deepspeed_b200.hcache import (HCacheSessionManager, HiddenStateStore, LinkModel, SessionState, plan_restore,
                                          storage_bytes)
from hcache_deepspeed_b200.hcache.format import StateHeader, plan_layout
from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch
from hcache_deepspeed_b200.io import aio_handle
from hcache_deepspeed_b200.models.arch import named_arch


def _engine(arch="tiny-llama-test", **hc):
    cfg = RaggedInferenceEngineConfig(state_manager=dict(max_tracked_sequences=16, max_ragged_batch_size=256,
                                                         max_ragged_sequence_count=8, max_context=512,
                                                         memory_config=dict(mode="allocate", size=32 * 1024 * 1024)),
                                      hcache=hc or {"enabled": True})
    return build_engine_from_arch(arch, cfg, seed=1, dtype=torch.float32, device=torch.device("cpu"))


def test_planner_mha_vs_gqa_and_balance():
    pcie = LinkModel()
    p7 = plan_restore(named_arch("llama-2-7b"), 4096, pcie, allow_recompute=False)
    assert p7.kinds == ["hidden"] * 32  # MHA: hidden states are half the bytes of KV and the link is the bottleneck
    p70 = plan_restore(named_arch("llama-2-70b"), 4096, pcie, allow_recompute=False)
    assert p70.kinds == ["kv"] * 80  # GQA 8 KV heads: KV is 4x smaller than the hidden state
    # with recompute allowed the idle GPU takes a prefix, and io/compute end up balanced
    p = plan_restore(named_arch("llama-2-7b"), 4096, pcie)
    r = len(p.layers_of("none"))
    assert 0 < r < 16 and p.kinds[:r] == ["none"] * r and abs(p.io_s - p.compute_s) / p.est_seconds < 0.15
    assert p.est_seconds < p7.est_seconds
    # a fast link makes compute the bottleneck: some layers move to KV offload (no projection needed)
    fast = plan_restore(named_arch("llama-2-7b"), 4096, LinkModel(io_bytes_per_s=400e9), allow_recompute=False)
    assert 0 < len(fast.layers_of("kv")) < 32 and abs(fast.io_s - fast.compute_s) / fast.est_seconds < 0.2
    assert storage_bytes(named_arch("llama-2-7b"), 4096, p7) == 32 * 4096 * 4096 * 2
    # TP: each rank moves 1/tp of the bytes
    p8 = plan_restore(named_arch("llama-2-7b"), 4096, LinkModel(tp_size=8), allow_recompute=False)
    assert p8.io_s < p7.io_s / 6


def test_format_roundtrip():
    hdr = plan_layout(torch.bfloat16, 100, 64, ["none", "hidden", "kv", "hidden"], kv_shape=[2, 2, 4, 64, 16], kv_heads=4,
                      head_dim=16, block_size=64, tokens_nbytes=400)
    back = StateHeader.from_bytes(hdr.to_bytes())
    assert [s.kind for s in back.layers] == ["none", "hidden", "kv", "hidden"]
    offs = [s.offset for s in back.layers if s.kind != "none"]
    assert all(o % 4096 == 0 for o in offs) and offs == sorted(offs)
    assert back.layers[1].nbytes == 100 * 64 * 2 and back.torch_dtype == torch.bfloat16
    with pytest.raises(ValueError):
        StateHeader.from_bytes(b"garbage!" + b"\0" * 32)


@pytest.mark.parametrize("use_direct", [False, True])
def test_aio_handle(tmp_path, use_direct):
    h = aio_handle(block_size=1 << 16, queue_depth=8, intra_op_parallelism=4, use_direct=use_direct)
    x = torch.randn(1 << 18)
    p = str(tmp_path / "blob.bin")
    assert h.sync_pwrite(x, p) == x.numel() * 4
    y = torch.empty_like(x)
    r1 = h.async_pread(y[:1 << 17], p, 0)
    r2 = h.async_pread(y[1 << 17:], p, (1 << 17) * 4)
    assert h.wait(r1) == (1 << 17) * 4 and h.wait(r2) == (1 << 17) * 4
    assert torch.equal(x, y)
    z = torch.empty(777)
    h.sync_pread(z, p, file_offset=4 * 1001)  # unaligned offset / length
    assert torch.equal(z, x[1001:1778])
    with pytest.raises(OSError):
        h.sync_pread(z, str(tmp_path / "missing.bin"))
    assert h.get_block_size() == 1 << 16 and h.get_intra_op_parallelism() == 4 and h.pending() == 0


def test_store_spills_to_nvme_and_loads_back(tmp_path):
    L, T, H = 3, 50, 32
    def mk(seed):
        g = torch.Generator().manual_seed(seed)
        return SessionState(tokens=torch.randint(0, 100, (T, ), generator=g).to(torch.int32), kinds=["hidden", "kv", "hidden"],
                            hidden=torch.randn(2, T, H, generator=g).to(torch.bfloat16),
                            kv=torch.randn(2, 1, 2, 2, 16, 8, generator=g).to(torch.bfloat16), meta={"seed": seed})
    one = mk(0).nbytes()
    store = HiddenStateStore(dram_bytes=int(2.5 * one), nvme_dir=str(tmp_path))
    states = {u: mk(u) for u in range(5)}
    for u, st in states.items():
        store.save(u, st)
    assert store.tier_of(4) == "dram" and store.tier_of(0) == "nvme" and store.stats["spills"] >= 3
    assert os.path.exists(tmp_path / "session_0.hcache")
    store.prefetch(1)
    for u in (0, 1, 4):
        got = store.load(u)
        assert torch.equal(got.tokens, states[u].tokens) and got.kinds == states[u].kinds
        want_kv = states[u].kv
        if got.kv_layer_major != states[u].kv_layer_major:   # states read back from NVMe come layer-major
            want_kv = want_kv.transpose(0, 1)
        assert torch.equal(got.hidden, states[u].hidden) and torch.equal(got.kv, want_kv)
        assert got.meta["seed"] == u
    assert store.stats["nvme_loads"] >= 2 and store.dram_bytes_used <= int(2.5 * one) + one
    store.drop(0)
    assert 0 not in store


@pytest.mark.parametrize("arch", ["tiny-llama-test", "tiny-opt-test"])
def test_mixed_plan_restore_equals_live(arch):
    eng = _engine(arch)
    spec = eng.model().spec
    g = torch.Generator().manual_seed(0)
    toks = torch.randint(0, spec.vocab, (70, ), generator=g)
    _, lat = eng.put([1], [toks])
    seq = eng.state_manager.get_sequence(1)
    # layer 0 recomputed, layer 1 from ... with only two layers: try every split of {recompute, hidden, kv}
    for kinds in (["none", "hidden"], ["hidden", "kv"], ["none", "kv"], ["kv", "hidden"], ["none", "none"]):
        hl = [i for i, k in enumerate(kinds) if k == "hidden"]
        kl = [i for i, k in enumerate(kinds) if k == "kv"]
        rl = [i for i, k in enumerate(kinds) if k == "none"]
        if rl != list(range(len(rl))):
            continue
        kv = eng.state_manager.kv_cache.offload_layers(seq.all_block_ids, kl) if kl else None
        uid = 100 + len(hl) * 10 + len(kl) + (50 if kinds[0] == "kv" else 0)
        major = bool(kl) and kinds[-1] == "kv"    # also exercise the layer-major (per-layer streaming) KV layout
        eng.restore_session(uid, toks, hidden=lat[0][hl] if hl else None, hidden_layers=hl,
                            kv=kv.transpose(0, 1).contiguous() if major else kv, kv_layers=kl, kv_layer_major=major,
                            recompute_layers=rl)
        probe = torch.randint(0, spec.vocab, (3, ), generator=g)
        la, _ = eng.put([1], [probe], capture_latents=False)
        lb, _ = eng.put([uid], [probe], capture_latents=False)
        torch.testing.assert_close(lb, la, rtol=1e-5, atol=1e-5)
        eng.flush(uid)
        # roll the live session back by re-creating it (its KV grew by the probe)
        eng.flush(1)
        _, lat = eng.put([1], [toks])
        seq = eng.state_manager.get_sequence(1)
    with pytest.raises(ValueError):
        eng.restore_session(999, toks, hidden=lat[0][[0]], hidden_layers=[0], recompute_layers=[1])


@pytest.mark.parametrize("kv_plan", [False, True])
def test_session_manager_multi_turn(tmp_path, kv_plan):
    eng = _engine()
    spec = eng.model().spec
    # kv_plan: a very fast link makes the projections the bottleneck, so the planner keeps layers as KV pages; a tiny DRAM
    # budget sends every saved state through the NVMe tier (layer-major KV segments) before it is restored
    store = HiddenStateStore(dram_bytes=(1 << 12) if kv_plan else (1 << 30), nvme_dir=str(tmp_path))
    mgr = HCacheSessionManager(eng, store, link=LinkModel(io_bytes_per_s=1e15) if kv_plan else LinkModel(),
                               allow_recompute=False, allow_kv=kv_plan)
    ref_eng = _engine()  # never flushed: the "recompute everything" truth
    for (n, t) in ref_eng.model().w.named_tensors().items():
        t.copy_(eng.model().w.named_tensors()[n])
    g = torch.Generator().manual_seed(5)
    history = []
    for turn in range(3):
        prompt = torch.randint(0, spec.vocab, (20 + turn, ), generator=g).to(torch.int32)
        logits = mgr.prefill(7, prompt)
        ref_logits, _ = ref_eng.put([7], [prompt], capture_latents=False)
        torch.testing.assert_close(logits, ref_logits, rtol=1e-5, atol=1e-5)
        gen = mgr.decode([7], 4, logits)[0]
        cur = ref_logits
        for i, tok in enumerate(gen):  # feed the same tokens to the reference engine
            assert int(cur.argmax(-1)) == tok
            if i < len(gen) - 1:
                cur, _ = ref_eng.put([7], [torch.tensor([tok], dtype=torch.int32)], capture_latents=False)
        history += prompt.tolist() + gen[:-1]
        st = mgr.end_turn(7)
        assert not mgr.is_resident(7) and eng.state_manager.get_sequence(7) is None
        assert st.n_tokens == len(history)
        if kv_plan:
            assert st.kv is not None and st.kv_layer_major and st.kv.shape[0] == len(st.layers_of("kv")) > 0
        else:
            assert st.hidden.shape[1] == len(history)
    assert store.stats["saves"] == 3 and store.stats["dram_hits"] + store.stats["nvme_loads"] == 2


def test_fp8_engine_mode_cpu():
    """quantization_mode='wf8af8': layer weights become E4M3 + per-channel scales; logits track the 16/32-bit engine,
    restore == live KV still holds (both paths project K/V through the same fp8 weights)."""
    from hcache_deepspeed_b200.ops.fp8 import Fp8Weight

    def build(mode):
        cfg = RaggedInferenceEngineConfig(state_manager=dict(max_tracked_sequences=16, max_ragged_batch_size=256,
                                                             max_ragged_sequence_count=8, max_context=512,
                                                             memory_config=dict(mode="allocate", size=32 * 1024 * 1024)),
                                          hcache={"enabled": True}, quantization={"quantization_mode": mode})
        return build_engine_from_arch("tiny-llama-test", cfg, seed=1, dtype=torch.float32, device=torch.device("cpu"))

    ref_eng, eng = build(None), build("wf8af8")
    lw = eng.model().w.layers[0]
    assert isinstance(lw.w_qkv, Fp8Weight) and isinstance(lw.w_down, Fp8Weight) and not isinstance(eng.model().w.lm_head, Fp8Weight)
    torch.testing.assert_close(lw.w_qkv.dequantize(), ref_eng.model().w.layers[0].w_qkv.float(), rtol=0.07, atol=1e-3)
    g = torch.Generator().manual_seed(0)
    prompt = torch.randint(0, 100, (40, ), generator=g)
    want, _ = ref_eng.put([1], [prompt])
    got, lat = eng.put([1], [prompt])
    cos = torch.nn.functional.cosine_similarity(got.float().flatten(), want.float().flatten(), dim=0)
    assert cos > 0.99, float(cos)
    # restore a second session from the captured hidden states: next-token logits equal the live session's
    eng.restore_kv([2], [prompt], [lat[0]])
    nxt = torch.tensor([7])
    a, _ = eng.put([1], [nxt])
    b, _ = eng.put([2], [nxt])
    torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-4)
    with pytest.raises(ValueError, match="wf8af8"):
        build("w4a16")


def test_fp6_engine_mode_cpu():
    """quantization_mode='wf6af16': E3M2 weights packed 6 bits each (input dims that are multiples of 128; others stay
    E4M3), per-channel scales; pack/unpack round-trips; logits track the unquantised engine."""
    from hcache_deepspeed_b200.models.arch import ArchSpec
    from hcache_deepspeed_b200.ops.fp8 import Fp8Weight, pack_6bit, quantize_rows_e3m2, unpack_6bit
    spec = ArchSpec(model_type="llama", n_layers=2, hidden=128, n_heads=4, n_kv_heads=2, head_dim=32, intermediate=256,
                    vocab=200, max_positions=256)

    def build(mode):
        cfg = RaggedInferenceEngineConfig(state_manager=dict(max_tracked_sequences=16, max_ragged_batch_size=128,
                                                             max_ragged_sequence_count=8, max_context=256,
                                                             memory_config=dict(mode="allocate", size=16 * 1024 * 1024)),
                                          hcache={"enabled": True}, quantization={"quantization_mode": mode})
        return build_engine_from_arch(spec, cfg, seed=1, dtype=torch.float32, device=torch.device("cpu"))

    ref_eng, eng = build(None), build("wf6af16")
    lw, lw_ref = eng.model().w.layers[0], ref_eng.model().w.layers[0]
    assert isinstance(lw.w_qkv, Fp8Weight) and lw.w_qkv.fmt == "e3m2" and lw.w_down.fmt == "e3m2"
    assert lw.w_qkv.q.shape == (lw_ref.w_qkv.shape[0], 128 * 3 // 4) and tuple(lw.w_qkv.shape) == tuple(lw_ref.w_qkv.shape)
    err = (lw.w_qkv.dequantize() - lw_ref.w_qkv.float()).abs().max() / lw_ref.w_qkv.float().abs().max()
    assert err < 0.07, float(err)        # 2 mantissa bits: at most 1/16 of the value, relative to the row maximum
    code, _ = quantize_rows_e3m2(torch.randn(5, 256))
    assert torch.equal(unpack_6bit(pack_6bit(code)), code) and int(code.max()) < 64
    g = torch.Generator().manual_seed(0)
    prompt = torch.randint(0, 200, (40, ), generator=g)
    want, _ = ref_eng.put([1], [prompt])
    got, lat = eng.put([1], [prompt])
    cos = torch.nn.functional.cosine_similarity(got.float().flatten(), want.float().flatten(), dim=0)
    assert cos > 0.97, float(cos)
    eng.restore_kv([2], [prompt], [lat[0]])
    a, _ = eng.put([1], [torch.tensor([7])])
    b, _ = eng.put([2], [torch.tensor([7])])
    torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-4)


def test_restore_layer_ready_hook_streams_layers_in_order():
    """``layer_ready(i)`` is called once per saved layer, in order, before that layer is consumed — the hook that lets
    storage reads complete layer by layer while earlier layers are already being restored."""
    cfg = RaggedInferenceEngineConfig(state_manager=dict(max_tracked_sequences=8, max_ragged_batch_size=128,
                                                         max_ragged_sequence_count=4, max_context=256,
                                                         memory_config=dict(mode="allocate", size=16 * 1024 * 1024)),
                                      hcache={"enabled": True})
    eng = build_engine_from_arch("tiny-llama-test", cfg, seed=2, dtype=torch.float32, device=torch.device("cpu"))
    prompt = torch.randint(0, 100, (33, ))
    _, lat = eng.put([1], [prompt])
    staged = torch.zeros_like(lat[0])          # "storage": a layer's bytes appear only when the hook is asked for them
    seen = []

    def ready(i):
        seen.append(i)
        staged[i].copy_(lat[0][i])

    eng.restore_kv([2], [prompt], [staged], layer_ready=ready)
    assert seen == list(range(lat[0].shape[0]))
    a, _ = eng.put([1], [torch.tensor([5])])
    b, _ = eng.put([2], [torch.tensor([5])])
    torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-4)


def test_many_sessions_interleaved_turns_through_dram_and_nvme(tmp_path):
    """Five conversations, three turns each in random order, a DRAM tier that holds about two saved states (the rest spill
    to the NVMe tier and come back): every turn's first logits and greedy continuation equal those of an engine that never
    dropped any KV cache."""
    import random
    eng, ref_eng = _engine(), _engine()
    spec = eng.model().spec
    for (n, t) in ref_eng.model().w.named_tensors().items():
        t.copy_(eng.model().w.named_tensors()[n])
    per_state = spec.n_layers * 40 * spec.hidden * 4          # fp32 hidden states of ~40 tokens
    store = HiddenStateStore(dram_bytes=2 * per_state, nvme_dir=str(tmp_path))
    mgr = HCacheSessionManager(eng, store, allow_recompute=False, allow_kv=False)
    rng = random.Random(1)
    g = torch.Generator().manual_seed(9)
    order = [u for u in range(5) for _ in range(3)]
    rng.shuffle(order)
    lengths = {u: 0 for u in range(5)}
    for uid in order:
        prompt = torch.randint(0, spec.vocab, (rng.randint(5, 18), ), generator=g).to(torch.int32)
        logits = mgr.prefill(uid, prompt)
        ref_logits, _ = ref_eng.put([uid], [prompt], capture_latents=False)
        torch.testing.assert_close(logits, ref_logits, rtol=2e-5, atol=2e-5)
        gen = mgr.decode([uid], 3, logits)[0]
        cur = ref_logits
        for i, tok in enumerate(gen):
            assert int(cur.argmax(-1)) == tok, (uid, i)
            if i < len(gen) - 1:
                cur, _ = ref_eng.put([uid], [torch.tensor([tok], dtype=torch.int32)], capture_latents=False)
        lengths[uid] += int(prompt.numel()) + len(gen) - 1
        st = mgr.end_turn(uid)
        assert st.n_tokens == lengths[uid] and not mgr.is_resident(uid)
    assert store.stats["saves"] == 15 and store.stats["spills"] > 0 and store.stats["nvme_loads"] > 0 and store.stats["dram_hits"] > 0
    assert int(eng.free_blocks.min()) == int(_engine().free_blocks.min())        # every KV block went back to the pool
