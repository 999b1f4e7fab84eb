"""Host-side ragged structures (ref: tests/unit/inference/v2/ragged/test_{blocked_allocator,manager_configs,ragged_wrapper}.py)."""
import random

import pytest
import torch
from pydantic import ValidationError

from hcache_deepspeed_b200.inference.v2.ragged import (BlockedAllocator, DSSequenceDescriptor, DSStateManager,
                                                         DSStateManagerConfig, KVCacheConfig, RaggedBatchWrapper)


def test_allocator_alloc_free_invariants():
    a = BlockedAllocator(64)
    assert a.free_blocks == 64
    x = a.allocate(16)
    assert len(set(x.tolist())) == 16 and a.free_blocks == 48
    with pytest.raises(ValueError):
        a.allocate(49)
    assert a.free_blocks == 48  # failed allocation takes nothing
    a.free(x[:5])
    assert a.free_blocks == 53
    with pytest.raises(ValueError):
        a.free(x[:1])  # double free
    with pytest.raises(ValueError):
        a.free([64])  # out of range
    with pytest.raises(ValueError):
        a.free([int(x[6]), int(x[6])])  # duplicate inside one call
    assert a.free_blocks == 53
    with pytest.raises(ValueError):
        BlockedAllocator(0)


def test_allocator_random_churn():
    random.seed(0)
    a = BlockedAllocator(128)
    held = []
    for _ in range(500):
        if held and (random.random() < 0.5 or a.free_blocks == 0):
            k = random.randrange(len(held))
            a.free(held.pop(k))
        else:
            n = random.randint(1, min(8, a.free_blocks))
            held.append(a.allocate(n))
        flat = [b for h in held for b in h.tolist()]
        assert len(flat) == len(set(flat)) and a.free_blocks == 128 - len(flat)


def test_manager_config_validation():
    DSStateManagerConfig(max_tracked_sequences=8, max_ragged_sequence_count=8, max_ragged_batch_size=8)
    with pytest.raises(ValidationError):
        DSStateManagerConfig(max_tracked_sequences=4, max_ragged_sequence_count=8)
    with pytest.raises(ValidationError):
        DSStateManagerConfig(max_ragged_sequence_count=64, max_ragged_batch_size=32)
    with pytest.raises(ValidationError):
        DSStateManagerConfig(max_context=-1)
    with pytest.raises(ValidationError):
        DSStateManagerConfig(not_a_field=1)


def _mgr(block=16, max_blocks=16):
    cfg = DSStateManagerConfig(max_tracked_sequences=4, max_ragged_batch_size=64, max_ragged_sequence_count=4, max_context=block * max_blocks,
                               memory_config={"mode": "allocate", "size": 1 << 20}, offload=True)
    kv = KVCacheConfig(block_size=block, cache_shape=(2, 2, 8), cache_dtype="fp32", max_blocks_per_allocation_group=max_blocks)
    return cfg, DSStateManager(cfg, (kv, ), device=torch.device("cpu"))


def test_batch_wrapper_metadata_and_limits():
    cfg, m = _mgr()
    b = RaggedBatchWrapper(cfg, block_size=16, max_blocks_per_seq=16, group=2, device=torch.device("cpu"))
    s1, s2, s3 = (m.get_or_create_sequence(u) for u in (1, 2, 3))
    s1.extend_kv_cache(m.allocate_blocks(3))
    s1._seen_tokens = 30  # mid-block continuation
    s1.pre_forward(4)
    s2.extend_kv_cache(m.allocate_blocks(2))
    s2.pre_forward(20)
    s3.extend_kv_cache(m.allocate_blocks(1))
    s3.pre_forward(1)
    b.insert_sequence(s1, torch.tensor([7, 8, 9, 10]))
    b.insert_sequence(s2, torch.arange(20))
    b.insert_sequence(s3, torch.tensor([5]))
    b.finalize()
    assert b.current_tokens == 25 and b.current_sequences == 3
    assert b.tok_pos().tolist()[:4] == [30, 31, 32, 33]
    blk = s1.all_block_ids.tolist()
    assert b.tok_slot().tolist()[:4] == [blk[1] * 16 + 14, blk[1] * 16 + 15, blk[2] * 16 + 0, blk[2] * 16 + 1]
    assert b.last_rows().tolist() == [3, 23, 24]
    assert b.tokens_to_seq().tolist() == [0] * 4 + [1] * 20 + [2]
    # group=2 -> at most 4 tokens per split-KV item; the 20-token sequence becomes one tensor-core item
    assert b.items_small().tolist() == [[0, 0, 4, 30], [2, 24, 1, 0]]
    assert b.items_large().tolist() == [[1, 4, 20, 0]]
    assert b.rows_per_small_item == 8 and b.max_kv_len_small == 34 and b.max_kv_len_large == 20
    assert b.block_tables()[0, :3].tolist() == blk
    assert b.inflight_seq_descriptors_shadow == [(0, 4, 30), (4, 20, 0), (24, 1, 0)]
    # limits
    b.clear()
    s4 = m.get_or_create_sequence(4)
    s4.extend_kv_cache(m.allocate_blocks(5))
    with pytest.raises(RuntimeError):
        b.insert_sequence(s4, torch.arange(65))
    # missing blocks are caught by the native builder
    b.clear()
    s4.pre_forward(50)
    s4._n_blocks = 1
    b.insert_sequence(s4, torch.arange(50), do_checks=False)
    with pytest.raises(RuntimeError):
        b.finalize()


def test_state_manager_tracking_flush_and_offload():
    cfg, m = _mgr()
    free0 = int(m.free_blocks[0])
    for u in range(4):
        m.get_or_create_sequence(u).extend_kv_cache(m.allocate_blocks(2))
    with pytest.raises(RuntimeError):
        m.get_or_create_sequence(99)  # tracking slots exhausted
    m.flush_sequence(2)
    assert m.get_sequence(2) is None and int(m.free_blocks[0]) == free0 - 6
    m.flush_sequence(1234)  # unknown uid: warning, no error
    # whole-sequence KV offload / reload keeps the data
    seq = m.get_sequence(0)
    cache = m.kv_cache.full_cache
    cache[:, seq.all_block_ids.long()] = torch.randn(2, 2, 2, 2, 16, 8)
    want = cache[:, seq.all_block_ids.long()].clone()
    host = m.offload_sequence(0)
    assert seq.cur_allocated_blocks == 0 and host.shape[0] == 2
    m.reload_sequence(0)
    torch.testing.assert_close(cache[:, seq.all_block_ids.long()], want)


def test_sequence_descriptor_accounting():
    s = DSSequenceDescriptor(0, max_blocks=4)
    s.extend_kv_cache(torch.tensor([5, 9], dtype=torch.int32))
    s.pre_forward(10)
    assert s.in_flight_tokens == 10 and s.seen_tokens == 0
    s.post_forward()
    assert s.seen_tokens == 10 and s.in_flight_tokens == 0
    with pytest.raises(ValueError):
        s.extend_kv_cache(torch.tensor([1, 2, 3], dtype=torch.int32))
    assert s.free_kv_cache([9]).tolist() == [9] and s.cur_allocated_blocks == 1


def test_continuous_batching_with_chunked_prefill_matches_sequential_oracle():
    """A serving loop in the style of the reference's MII scheduler: 20 requests, at most 6 sequences / 96 tokens per
    forward, prompts fed in chunks of <= 40 tokens next to running decodes (``query`` / ``can_schedule`` / ``put`` /
    ``flush``). Every request's greedy continuation equals the one it gets alone on a fresh engine, and every KV block is
    back in the pool at the end."""
    import random
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch
    from hcache_deepspeed_b200.inference.v2.scheduling_utils import SchedulingResult
    from hcache_deepspeed_b200.models.arch import named_arch
    spec = named_arch("tiny-llama-test")
    def mk(mem, T=96):
        cfg = RaggedInferenceEngineConfig(state_manager=dict(max_tracked_sequences=12, max_ragged_batch_size=T, max_ragged_sequence_count=6, max_context=256,
                                          memory_config=dict(mode="allocate", size=mem)), hcache={"enabled": False})
        return build_engine_from_arch(spec, cfg, seed=3, dtype=torch.float32, device=torch.device("cpu"))
    eng = mk(6 << 20)
    total_blocks = int(eng.free_blocks.min())
    rng = random.Random(0)
    g = torch.Generator().manual_seed(0)
    N = 20
    prompts = {u: torch.randint(0, spec.vocab, (rng.randint(3, 120),), generator=g) for u in range(N)}
    target_new = {u: rng.randint(1, 12) for u in range(N)}
    # oracle: one sequence at a time on a fresh engine
    oracle = {}
    ref = mk(8 << 20, 256)
    for u in range(N):
        toks = []
        lg, _ = ref.put([u], [prompts[u]])
        for _ in range(target_new[u]):
            t = int(lg[0].argmax()); toks.append(t)
            lg, _ = ref.put([u], [torch.tensor([t])])
        ref.flush(u); oracle[u] = toks
    # continuous batching with chunked prefill (split-fuse): token budget 96, <= 6 sequences per step
    pending = list(range(N)); active = {}   # uid -> dict(pos, out, last)
    done = {}
    steps = 0
    while pending or active:
        steps += 1
        uids, toks = [], []
        budget = 96
        # decodes first
        for u, st in list(active.items()):
            if st["pos"] >= len(prompts[u]) and len(uids) < 6 and budget > 0:
                uids.append(u); toks.append(torch.tensor([st["last"]])); budget -= 1
        # prefill chunks
        for u in list(active) + pending[:]:
            if len(uids) >= 6 or budget <= 0: break
            if u in uids: continue
            st = active.get(u)
            pos = st["pos"] if st else 0
            if pos >= len(prompts[u]): continue
            n = min(budget, len(prompts[u]) - pos, 40)
            max_tok, max_blk = eng.query(u, n, int(eng.free_blocks.min()))
            n = min(n, max_tok)
            if n <= 0: continue
            if eng.can_schedule(uids + [u], [int(t.numel()) for t in toks] + [n]) != SchedulingResult.Success:
                continue
            if st is None:
                pending.remove(u); active[u] = st = {"pos": 0, "out": [], "last": None}
            uids.append(u); toks.append(prompts[u][pos:pos + n]); budget -= n
        if not uids:
            raise RuntimeError(f"nothing schedulable: free={eng.free_blocks} active={len(active)}")
        lg, _ = eng.put(uids, toks)
        for i, (u, t) in enumerate(zip(uids, toks)):
            st = active[u]
            if st["pos"] < len(prompts[u]):
                st["pos"] += int(t.numel())
                if st["pos"] < len(prompts[u]): continue       # more prompt chunks to go: logits unused
            nxt = int(lg[i].argmax()); st["out"].append(nxt); st["last"] = nxt
            if len(st["out"]) == target_new[u]:
                done[u] = st["out"]; eng.flush(u); del active[u]
    bad = [u for u in range(N) if done[u] != oracle[u]]
    assert not bad, bad
    assert int(eng.free_blocks.min()) == total_blocks and steps < 60


def test_reference_accessor_surface_of_wrapper_cache_and_manager():
    """Accessors of the reference's ragged classes (ragged_wrapper.py:230-276, kv_cache.py:144-201, ragged_manager.py:170)."""
    import torch
    from hcache_deepspeed_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_arch
    from hcache_deepspeed_b200.models.arch import named_arch
    cfg = RaggedInferenceEngineConfig(state_manager=dict(max_tracked_sequences=4, max_ragged_batch_size=64, max_ragged_sequence_count=2,
                                                         max_context=512, memory_config=dict(mode="allocate", size=8 << 20)))
    eng = build_engine_from_arch(named_arch("tiny-llama-test"), cfg, dtype=torch.float32, device=torch.device("cpu"))
    eng.put([1, 2], [torch.arange(5), torch.arange(3)])
    b = eng._batch
    assert b.inflight_seq_descriptors().tolist() == [[0, 5, 0, 0], [5, 3, 0, 0]]
    assert b.batch_metadata_buffer().tolist() == [8, 2]
    assert b.kv_ptrs().shape[0] == 2 and b.masks() is None
    sm = eng._state_manager
    kv = sm.kv_cache
    before = kv.free_blocks[0]
    ids = kv.reserve(2)
    assert kv.free_blocks[0] == before - 2 and int(sm.free_blocks[0]) == before - 2
    kv.free(ids)
    assert kv.free_blocks[0] == before
    assert sm.get_sequence(1).max_context >= 1

    # restore_kv with a caller-supplied projection fills the sequence's pages, spanning several blocks
    bs = sm.kv_block_size
    layers, heads, hd = kv.num_caches, kv.get_cache(0).shape[2], kv.get_cache(0).shape[4]
    n_tok, hidden = bs + 3, 16
    g = torch.Generator().manual_seed(0)
    latent = [torch.randn(n_tok, hidden, generator=g) for _ in range(layers)]
    w = [torch.randn(hidden, 2 * heads * hd, generator=g) for _ in range(layers)]
    got = sm.restore_kv(9, latent, lambda x, wl, b=None: x @ wl, w)
    seq = sm.get_sequence(9)
    assert got == n_tok and seq.seen_tokens == n_tok and seq.cur_allocated_blocks == 2
    blocks = seq.all_block_ids.tolist()
    for l in range(layers):
        want = (latent[l] @ w[l]).reshape(n_tok, 2, heads, hd)
        cache = kv.get_cache(l)
        for t in (0, bs - 1, bs, n_tok - 1):
            torch.testing.assert_close(cache[blocks[t // bs], :, :, t % bs], want[t])
    sm.flush_sequence(9)
    assert kv.free_blocks[0] == before
