"""TP sharding rules: every rank's slice is disjoint and the union covers the parameter (the checks of the
reference's tests/unit/inference/v2/model_implementations/sharding/test_{qkv,attn_out,mlp}_sharding.py)."""
import pytest
import torch

from hcache_deepspeed_b200.parallel import sharding as sh


def test_endpoints_cover_and_balance():
    for dim, n, gran in [(4096, 8, 32), (11008, 8, 32), (352, 3, 32), (50272, 4, 1), (7, 3, 1)]:
        ends = [sh.get_shard_endpoints(dim, r, n, gran) for r in range(n)]
        assert ends[0][0] == 0 and ends[-1][1] == dim
        for (a, b), (c, d) in zip(ends, ends[1:]):
            assert b == c and a % gran == 0
        sizes = [b - a for a, b in ends]
        assert max(sizes) - min(sizes) <= gran


@pytest.mark.parametrize("n_q,n_kv,tp", [(32, 32, 8), (40, 40, 3), (64, 8, 8), (64, 8, 4), (8, 2, 8), (16, 1, 4), (71, 1, 1)])
def test_qkv_sharding_covers_every_head(n_q, n_kv, tp):
    d = 4
    # each row's value encodes (kind, head): q heads 0.., k heads 1000.., v heads 2000..
    rows = ([h for h in range(n_q) for _ in range(d)] + [1000 + h for h in range(n_kv) for _ in range(d)] +
            [2000 + h for h in range(n_kv) for _ in range(d)])
    param = torch.tensor(rows, dtype=torch.float32)[:, None].repeat(1, 3)
    seen_q, k_count = [], {}
    for r in range(tp):
        shard = sh.shard_qkv_param(param, r, tp, d, n_q, n_kv)
        lq, lkv = sh.get_local_heads(r, tp, n_q, n_kv)
        assert shard.shape[0] == (lq + 2 * lkv) * d
        vals = shard[::d, 0].long().tolist()
        q, k, v = vals[:lq], vals[lq:lq + lkv], vals[lq + lkv:]
        assert all(x < 1000 for x in q) and all(1000 <= x < 2000 for x in k) and all(x >= 2000 for x in v)
        assert [x - 1000 for x in k] == [x - 2000 for x in v]
        ratio = n_q // n_kv
        for qh in q:  # a query head always travels with its own KV head
            assert (qh // ratio) + 1000 in k
        seen_q += q
        for x in k:
            k_count[x] = k_count.get(x, 0) + 1
    assert sorted(seen_q) == list(range(n_q))
    assert sorted(k_count) == [1000 + h for h in range(n_kv)]
    if n_kv < tp and n_kv != n_q:
        assert all(c == tp // n_kv for c in k_count.values())  # replicated KV heads


def test_qkv_sharding_rejects_bad_gqa():
    with pytest.raises(ValueError):
        sh.local_head_indices(0, 3, 8, 2)   # 3 ranks cannot split 2 kv heads evenly
    with pytest.raises(ValueError):
        sh.local_head_indices(0, 16, 8, 8)  # fewer heads than shards


@pytest.mark.parametrize("n_q,n_kv,tp", [(32, 32, 4), (64, 8, 8), (8, 2, 8)])
def test_attn_out_matches_qkv_heads(n_q, n_kv, tp):
    d, H = 2, 6
    w = torch.arange(n_q * d, dtype=torch.float32)[None, :].repeat(H, 1)  # column c holds value c
    cols = []
    for r in range(tp):
        s = sh.shard_attn_out_param(w, r, tp, d, n_q, n_kv)
        q_heads, _ = sh.local_head_indices(r, tp, n_q, n_kv)
        assert s.shape == (H, len(q_heads) * d)
        assert s[0].long().tolist() == [h * d + i for h in q_heads for i in range(d)]
        cols += s[0].long().tolist()
    assert sorted(cols) == list(range(n_q * d))
    bias = torch.ones(H)
    assert sh.shard_attn_out_param(bias, 0, tp, d, n_q, n_kv) is bias
    assert sh.shard_attn_out_param(bias, 1, tp, d, n_q, n_kv) is None


@pytest.mark.parametrize("inter,tp", [(11008, 8), (352, 3), (256, 2)])
def test_mlp_sharding(inter, tp):
    H = 4
    up = torch.arange(2 * inter, dtype=torch.float32)[:, None].repeat(1, H)  # gated: [gate; up] halves
    down = torch.arange(inter, dtype=torch.float32)[None, :].repeat(H, 1)
    g_rows, u_rows, d_cols = [], [], []
    for r in range(tp):
        s1 = sh.shard_mlp_1_param(up, r, tp, gated=True)
        n = s1.shape[0] // 2
        g_rows += s1[:n, 0].long().tolist()
        u_rows += (s1[n:, 0].long() - inter).tolist()
        s2 = sh.shard_mlp_2_param(down, r, tp)
        d_cols += s2[0].long().tolist()
        assert s2.shape[1] == n == sh.sharded_intermediate_dim(inter, tp, r)
    assert g_rows == u_rows == d_cols == list(range(inter))
    b = torch.ones(H)
    assert sh.shard_mlp_2_param(b, 0, tp) is b and sh.shard_mlp_2_param(b, 1, tp) is None


def test_unembed_and_embedding():
    w = torch.arange(50, dtype=torch.float32)[:, None].repeat(1, 2)
    rows = []
    for r in range(3):
        s = sh.shard_unembed_param(w, r, 3)
        assert s.shape[0] == sh.sharded_unembed_dim(50, r, 3)
        rows += s[:, 0].long().tolist()
    assert rows == list(range(50))
    assert sh.shard_embedding_param(w, 1, 3) is w
