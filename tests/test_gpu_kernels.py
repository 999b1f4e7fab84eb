"""Numerics of every sm_100a kernel against the fp32 PyTorch restatement of the same op
(hcache_deepspeed_b200.ops.reference). Mirrors the reference's per-kernel tests
(tests/unit/inference/v2/kernels/{core_ops,ragged_ops}/*, tolerances inference_test_utils.py:14-22)."""
import math

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

from hcache_deepspeed_b200.ops import reference as ref  # noqa: E402

DTYPES = [torch.bfloat16, torch.float16]


def tol(dtype):
    # bf16 has 8 bits of mantissa: one rounding of the output is 2^-8 relative
    return dict(rtol=1.6e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=4e-3, atol=4e-3)


def K():
    from hcache_deepspeed_b200.ops import kernels
    return kernels


def rand(*shape, dtype=torch.bfloat16, scale=1.0):
    return (torch.randn(*shape, device="cuda", dtype=torch.float32) * scale).to(dtype)


# ------------------------------------------------------------------------------------------------
# GEMM
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("M,N,K_,bn", [
    (128, 256, 128, 0), (300, 520, 192, 0), (16, 4096, 4096, 0), (1, 1024, 512, 32), (77, 384, 320, 64),
    (2048, 1024, 1024, 256), (512, 768, 2048, 128), (9, 11008 // 8 * 2, 4096, 0),
])
def test_gemm_store(dtype, M, N, K_, bn):
    torch.manual_seed(0)
    a, w = rand(M, K_, dtype=dtype), rand(N, K_, dtype=dtype, scale=K_ ** -0.5)
    out = K().gemm(a, w, bn=bn)
    torch.cuda.synchronize()
    want = ref.gemm(a, w, out_dtype=torch.float32)
    torch.testing.assert_close(out.float(), want, **tol(dtype))


@pytest.mark.parametrize("M,N,K_,bn,splits,gated", [(16, 4096, 4096, 128, 4, False), (1, 12288, 4096, 0, 0, False),
                                                    (64, 2 * 1376, 2048, 64, 3, True), (100, 520, 1024, 128, 2, False),
                                                    (8, 22016, 4096, 0, 0, True)])
def test_gemm_store_splitk(M, N, K_, bn, splits, gated):
    torch.manual_seed(12)
    a, w, b = rand(M, K_), rand(N, K_, scale=K_ ** -0.5), rand(N)
    ws = torch.zeros(K().gemm_splitk_workspace_bytes(M, N), dtype=torch.uint8, device="cuda")
    want = ref.gemm(a, w, bias=b, act=3 if gated else 0, gated=gated, out_dtype=torch.float32)
    for _ in range(2):  # second call: scratch must have been left zeroed
        out = K().gemm(a, w, bias=b, act=3 if gated else 0, gated=gated, bn=bn, splits=splits, workspace=ws)
        torch.testing.assert_close(out.float(), want, **tol(torch.bfloat16))
    assert int(ws.view(torch.int32).abs().max()) == 0


@pytest.mark.parametrize("M,N,K_", [(8, 4096, 4096), (64, 1000, 520), (300, 2 * 688, 1024), (1, 96, 64)])
def test_gemm_tiled_weights(M, N, K_):
    """pre-tiled weight layout: same results as the row-major operand for store / gated / accum epilogues"""
    torch.manual_seed(13)
    k = K()
    a, w, b = rand(M, K_), rand(N, K_, scale=K_ ** -0.5), rand(N)
    wt = k.TiledWeight.from_dense(w)
    torch.testing.assert_close(wt.to_dense(), w, rtol=0, atol=0)
    want = ref.gemm(a, w, bias=b, out_dtype=torch.float32)
    torch.testing.assert_close(k.gemm(a, wt, bias=b).float(), want, **tol(torch.bfloat16))
    if N % 2 == 0:
        wantg = ref.gemm(a, w, act=3, gated=True, out_dtype=torch.float32)
        torch.testing.assert_close(k.gemm(a, wt, act=3, gated=True).float(), wantg, **tol(torch.bfloat16))
    acc = torch.zeros(M, N, device="cuda")
    k.gemm_accum(a, wt, acc, bias=b)
    torch.testing.assert_close(acc, want, rtol=2e-3, atol=2e-3)
    if N >= 64:
        sub = wt.rows_from(32)
        torch.testing.assert_close(k.gemm(a, sub).float(), ref.gemm(a, w[32:], out_dtype=torch.float32), **tol(torch.bfloat16))


@pytest.mark.parametrize("act", [1, 2, 3, 4])
def test_gemm_bias_act(act):
    torch.manual_seed(1)
    M, N, K_ = 200, 640, 256
    a, w, b = rand(M, K_), rand(N, K_, scale=K_ ** -0.5), rand(N)
    out = K().gemm(a, w, bias=b, act=act)
    want = ref.gemm(a, w, bias=b, act=act, out_dtype=torch.float32)
    torch.testing.assert_close(out.float(), want, **tol(torch.bfloat16))


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("M", [1, 16, 17, 33, 64])
@pytest.mark.parametrize("mode", ["store", "gated", "accum"])
@pytest.mark.parametrize("use_ws", [False, True])
def test_gemm_skinny(dtype, M, mode, use_ws):
    """Decode-size M takes the feature-major weight-streaming kernel (gemm_skinny.cuh): ragged N and K, every token
    tile (16/32/64; bn=-1 forces the kernel for every M <= 64), with and without split-K scratch; scratch must come back zeroed."""
    torch.manual_seed(M)
    N, K_ = (690 if mode == "gated" else 1000), 2120
    a, w, b = rand(M, K_, dtype=dtype), rand(N, K_, dtype=dtype, scale=K_ ** -0.5), rand(N, dtype=dtype)
    ws = torch.zeros(K().gemm_splitk_workspace_bytes(M, N), dtype=torch.uint8, device="cuda") if use_ws else None
    for rep in range(2):
        if mode == "accum":
            acc = torch.zeros(M, N, device="cuda", dtype=torch.float32)
            K().gemm_accum(a, w, acc, bias=b, bn=-1)
            want = ref.gemm(a, w, bias=b, out_dtype=torch.float32)
            torch.testing.assert_close(acc, want, rtol=2e-3, atol=2e-3)
        else:
            gated = mode == "gated"
            out = K().gemm(a, w, bias=b, act=3 if gated else 2, gated=gated, workspace=ws, bn=-1)
            want = ref.gemm(a, w, bias=b, act=3 if gated else 2, gated=gated, out_dtype=torch.float32)
            torch.testing.assert_close(out.float(), want, **tol(dtype))
    if ws is not None:
        assert int(ws.view(torch.int32).abs().max()) == 0


@pytest.mark.parametrize("M", [5, 128, 700])
def test_gemm_gated(M):
    torch.manual_seed(2)
    N, K_ = 2 * 688, 512
    a, w = rand(M, K_), rand(N, K_, scale=K_ ** -0.5)
    out = K().gemm(a, w, act=3, gated=True)
    assert out.shape == (M, N // 2)
    want = ref.gemm(a, w, act=3, gated=True, out_dtype=torch.float32)
    torch.testing.assert_close(out.float(), want, **tol(torch.bfloat16))


@pytest.mark.parametrize("M,N,K_,splits", [(8, 4096, 4096, 0), (64, 512, 11008, 4), (300, 1024, 1024, 2), (16, 256, 64, 0)])
def test_gemm_accum_splitk(M, N, K_, splits):
    torch.manual_seed(3)
    a, w, b = rand(M, K_), rand(N, K_, scale=K_ ** -0.5), rand(N)
    acc = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    K().gemm_accum(a, w, acc, bias=b, splits=splits)
    want = ref.gemm(a, w, bias=b, out_dtype=torch.float32)
    torch.testing.assert_close(acc, want, rtol=2e-3, atol=2e-3)
    K().gemm_accum(a, w, acc, bias=b, splits=splits)  # accumulates on top
    torch.testing.assert_close(acc, 2 * want, rtol=2e-3, atol=4e-3)


# ------------------------------------------------------------------------------------------------
# norms
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("cols", [768, 4096, 5120, 8192])
@pytest.mark.parametrize("layer_norm", [False, True])
def test_norm_plain(dtype, cols, layer_norm):
    torch.manual_seed(4)
    x, g, b = rand(37, cols, dtype=dtype), rand(cols, dtype=dtype), rand(cols, dtype=dtype)
    out = torch.empty_like(x)
    K().norm(x, g, out, beta=b if layer_norm else None, eps=1e-5, layer_norm=layer_norm)
    want = ref.norm(x, g, b if layer_norm else None, 1e-5, layer_norm)
    torch.testing.assert_close(out.float(), want.float(), **tol(dtype))


@pytest.mark.parametrize("delta_f32", [False, True])
@pytest.mark.parametrize("layer_norm", [False, True])
def test_norm_residual(delta_f32, layer_norm):
    torch.manual_seed(5)
    rows, cols = 19, 4096
    x, g, b = rand(rows, cols), rand(cols), rand(cols)
    delta = torch.randn(rows, cols, device="cuda") if delta_f32 else rand(rows, cols)
    delta_in = delta.clone()
    res = torch.empty_like(x)
    out = torch.empty_like(x)
    K().norm(x, g, out, beta=b if layer_norm else None, delta=delta, zero_delta=True, res_out=res, eps=1e-6,
             layer_norm=layer_norm)
    want_res = (x.float() + delta_in.float()).to(torch.bfloat16)
    want = ref.norm(want_res, g, b if layer_norm else None, 1e-6, layer_norm)
    torch.testing.assert_close(res.float(), want_res.float(), rtol=0, atol=0)
    torch.testing.assert_close(out.float(), want.float(), **tol(torch.bfloat16))
    if delta_f32:
        assert float(delta.abs().max()) == 0.0  # consumed and zeroed


def test_norm_gather_rows():
    torch.manual_seed(6)
    x, g = rand(50, 2048), rand(2048)
    idx = torch.tensor([49, 0, 7, 7, 31], device="cuda", dtype=torch.int32)
    out = torch.empty(5, 2048, device="cuda", dtype=torch.bfloat16)
    K().norm(x, g, out, row_idx=idx, eps=1e-5)
    want = ref.norm(x[idx.long()], g, None, 1e-5, False)
    torch.testing.assert_close(out.float(), want.float(), **tol(torch.bfloat16))


# ------------------------------------------------------------------------------------------------
# ragged helpers
# ------------------------------------------------------------------------------------------------
def make_cache(n_blocks, n_kv, bs, d, dtype=torch.bfloat16):
    return torch.zeros(n_blocks, 2, n_kv, bs, d, device="cuda", dtype=dtype)


def make_slots(T, n_blocks, bs, seed=0):
    g = torch.Generator().manual_seed(seed)
    perm = torch.randperm(n_blocks * bs, generator=g)[:T]
    return perm.to(torch.int32).cuda()


@pytest.mark.parametrize("max_pos", [5000, 200000])     # long contexts: the in-kernel range reduction must hold up
@pytest.mark.parametrize("d,rd,n_q,n_kv", [(128, 128, 4, 4), (128, 128, 8, 2), (64, 64, 12, 12), (80, 32, 4, 4), (96, 96, 2, 1)])
def test_rope_kv_append(d, rd, n_q, n_kv, max_pos):
    torch.manual_seed(7)
    T, bs, n_blocks = 45, 64, 6
    qkv = rand(T, (n_q + 2 * n_kv) * d)
    pos = torch.randint(max_pos // 2, max_pos, (T,), device="cuda", dtype=torch.int32)
    slot = make_slots(T, n_blocks, bs)
    slot[3] = -1
    inv = ref.default_inv_freq(rd, device="cuda")
    cache = make_cache(n_blocks, n_kv, bs, d)
    q_out = torch.empty(T, n_q * d, device="cuda", dtype=torch.bfloat16)
    K().rope_kv_append(qkv, q_out, cache, pos, slot, inv, n_q, n_kv, d, rd, bs)
    q, k, v = qkv.split([n_q * d, n_kv * d, n_kv * d], dim=1)
    want_q = ref.rope_rotate_half(q.view(T, n_q, d), pos, inv, rd)
    want_k = ref.rope_rotate_half(k.reshape(T, n_kv, d), pos, inv, rd)
    want_cache = make_cache(n_blocks, n_kv, bs, d)
    ref.kv_append(want_cache, want_k, v.reshape(T, n_kv, d), slot)
    torch.testing.assert_close(q_out.view(T, n_q, d).float(), want_q.float(), **tol(torch.bfloat16))
    torch.testing.assert_close(cache.float(), want_cache.float(), **tol(torch.bfloat16))


def test_embed_gather_argmax_cvt():
    torch.manual_seed(8)
    V, H, T = 1000, 768, 33
    W, P = rand(V, H), rand(64, H)
    ids = torch.randint(0, V, (T,), device="cuda", dtype=torch.int32)
    pos = torch.randint(0, 60, (T,), device="cuda", dtype=torch.int32)
    out = torch.empty(T, H, device="cuda", dtype=torch.bfloat16)
    K().embed(ids, W, out, pos_weight=P, tok_pos=pos, pos_offset=2)
    want = ref.embed(ids, W, P, pos, 2)
    torch.testing.assert_close(out.float(), want.float(), rtol=1e-2, atol=1e-2)
    K().embed(ids, W, out)
    torch.testing.assert_close(out, W[ids.long()], rtol=0, atol=0)

    idx = torch.tensor([5, 1, 32], device="cuda", dtype=torch.int32)
    g = torch.empty(3, H, device="cuda", dtype=torch.bfloat16)
    K().gather_rows(out, idx, g)
    torch.testing.assert_close(g, out[idx.long()], rtol=0, atol=0)

    logits = rand(7, 32000)
    got = K().argmax(logits)
    assert got.tolist() == logits.float().argmax(-1).tolist()
    lf = torch.randn(3, 50257, device="cuda")
    assert K().argmax(lf).tolist() == lf.argmax(-1).tolist()

    src = torch.randn(1000, device="cuda")
    dst = torch.empty(1000, device="cuda", dtype=torch.bfloat16)
    K().cvt_f32(src, dst, zero_src=True)
    assert float(src.abs().max()) == 0.0


def test_kv_blocks_copy_roundtrip():
    torch.manual_seed(9)
    L, nb, kvh, bs, d = 3, 10, 2, 64, 128
    cache = rand(L, nb, 2, kvh, bs, d)
    ids = torch.tensor([7, 2, 9], device="cuda", dtype=torch.int32)
    staging = torch.empty(3, L, 2, kvh, bs, d, device="cuda", dtype=torch.bfloat16)
    K().kv_blocks_copy(cache, staging, ids, to_staging=True)
    want = cache[:, ids.long()].permute(1, 0, 2, 3, 4, 5)
    torch.testing.assert_close(staging, want.contiguous(), rtol=0, atol=0)
    cache2 = torch.zeros_like(cache)
    K().kv_blocks_copy(cache2, staging, ids, to_staging=False)
    torch.testing.assert_close(cache2[:, ids.long()], cache[:, ids.long()], rtol=0, atol=0)


# ------------------------------------------------------------------------------------------------
# fused projections (restore kernel and forward QKV)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("T,H,n_kv,d,rope", [(200, 512, 4, 128, True), (1000, 1024, 8, 128, True), (70, 768, 12, 64, False),
                                             (33, 256, 2, 64, True), (1100, 512, 32, 128, True),   # last two: CTA-pair kernel
                                             (700, 256, 40, 128, False)])
def test_restore_kv_proj(dtype, T, H, n_kv, d, rope):
    torch.manual_seed(10)
    bs = 64 if d == 128 else 128
    n_blocks = (T + bs - 1) // bs + 3
    lat = rand(T, H, dtype=dtype)
    w_kv = rand(2 * n_kv * d, H, dtype=dtype, scale=H ** -0.5)
    bias = None if rope else rand(2 * n_kv * d, dtype=dtype)
    pos = torch.arange(T, device="cuda", dtype=torch.int32) + 17
    slot = make_slots(T, n_blocks, bs, seed=1)
    inv = ref.default_inv_freq(d, device="cuda") if rope else None
    cache = make_cache(n_blocks, n_kv, bs, d, dtype)
    K().restore_kv_proj(lat, w_kv, cache, pos, slot, inv, n_kv, d, bs, bias_kv=bias)
    kv = ref.gemm(lat, w_kv, bias=bias, out_dtype=torch.float32)
    k, v = kv[:, :n_kv * d].view(T, n_kv, d), kv[:, n_kv * d:].view(T, n_kv, d)
    if rope:
        k = ref.rope_rotate_half(k, pos, inv, d)
    want = make_cache(n_blocks, n_kv, bs, d, torch.float32)
    ref.kv_append(want, k, v, slot)
    torch.testing.assert_close(cache.float(), want, **tol(dtype))


@pytest.mark.parametrize("T,n_q,n_kv", [(150, 4, 2), (1100, 32, 8)])   # the second shape takes the CTA-pair kernel
def test_qkv_rope_proj(T, n_q, n_kv):
    torch.manual_seed(11)
    H, d, bs = 512, 128, 64
    n_blocks = (T + bs - 1) // bs + 2
    x = rand(T, H)
    w = rand((n_q + 2 * n_kv) * d, H, scale=H ** -0.5)
    pos = torch.arange(T, device="cuda", dtype=torch.int32)
    slot = make_slots(T, n_blocks, bs, seed=2)
    inv = ref.default_inv_freq(d, device="cuda")
    cache = make_cache(n_blocks, n_kv, bs, d)
    q_out = torch.empty(T, n_q * d, device="cuda", dtype=torch.bfloat16)
    K().qkv_rope_proj(x, w, q_out, cache, pos, slot, inv, n_q, n_kv, d, bs)
    qkv = ref.gemm(x, w, out_dtype=torch.float32)
    q, k, v = qkv.split([n_q * d, n_kv * d, n_kv * d], dim=1)
    want_q = ref.rope_rotate_half(q.view(T, n_q, d), pos, inv, d)
    want_k = ref.rope_rotate_half(k.reshape(T, n_kv, d), pos, inv, d)
    want = make_cache(n_blocks, n_kv, bs, d, torch.float32)
    ref.kv_append(want, want_k, v.reshape(T, n_kv, d), slot)
    torch.testing.assert_close(q_out.view(T, n_q, d).float(), want_q, **tol(torch.bfloat16))
    torch.testing.assert_close(cache.float(), want, **tol(torch.bfloat16))


# ------------------------------------------------------------------------------------------------
# paged attention
# ------------------------------------------------------------------------------------------------
def build_paged_problem(seq_specs, n_q_heads, n_kv, d, bs, dtype=torch.bfloat16, seed=0):
    """seq_specs: list of (ctx_before, n_new). Returns q, cache, block_tables, items(list)."""
    g = torch.Generator().manual_seed(seed)
    total_blocks = sum((c + n + bs - 1) // bs for c, n in seq_specs) + 4
    perm = torch.randperm(total_blocks, generator=g).tolist()
    max_blocks = max((c + n + bs - 1) // bs for c, n in seq_specs)
    cache = (torch.randn(total_blocks, 2, n_kv, bs, d, generator=g) * 0.5).to(dtype).cuda()
    tables = torch.zeros(len(seq_specs), max_blocks, dtype=torch.int32)
    items, row = [], 0
    for s, (ctx, n_new) in enumerate(seq_specs):
        nb = (ctx + n_new + bs - 1) // bs
        tables[s, :nb] = torch.tensor([perm.pop() for _ in range(nb)], dtype=torch.int32)
        items.append((s, row, n_new, ctx))
        row += n_new
    q = (torch.randn(row, n_q_heads, d, generator=g)).to(dtype).cuda()
    return q, cache, tables.cuda(), items


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("n_q_heads,n_kv,d,bs,n_splits", [(8, 8, 128, 64, 1), (8, 8, 128, 64, 4), (16, 2, 128, 64, 3),
                                                         (12, 12, 64, 128, 2), (4, 4, 128, 32, 5)])
def test_attention_splitkv_decode(dtype, n_q_heads, n_kv, d, bs, n_splits):
    specs = [(0, 1), (5, 1), (127, 1), (128, 1), (1000, 1), (4095, 1), (333, 1)]
    q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype)
    items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
    G = n_q_heads // n_kv
    out = torch.zeros_like(q)
    ws_bytes = K().splitkv_workspace_bytes(len(items), n_q_heads, n_kv, d, n_splits, G)
    ws = torch.zeros(max(ws_bytes, 16), dtype=torch.uint8, device="cuda")
    scale = 1.0 / math.sqrt(d)
    for _ in range(2):  # second call checks the scratch counters were left clean
        out.zero_()
        K().paged_attention_splitkv(q, out, cache, tables, items_t, len(items), G, n_q_heads, n_kv, d, bs, n_splits,
                                    scale, ws)
        want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale)
        torch.testing.assert_close(out.float(), want.float(), **tol(dtype))


def test_attention_splitkv_multirow():
    # short continuations: several query tokens per item (MHA: rows = tokens)
    n_q_heads = n_kv = 4
    d, bs = 128, 64
    specs = [(300, 8), (0, 5), (64, 8), (2000, 3)]
    q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs)
    items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
    out = torch.zeros_like(q)
    ws = torch.zeros(K().splitkv_workspace_bytes(len(items), n_q_heads, n_kv, d, 2, 8), dtype=torch.uint8, device="cuda")
    scale = 1.0 / math.sqrt(d)
    K().paged_attention_splitkv(q, out, cache, tables, items_t, len(items), 8, n_q_heads, n_kv, d, bs, 2, scale, ws)
    want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale)
    torch.testing.assert_close(out.float(), want.float(), **tol(torch.bfloat16))


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("n_q_heads,n_kv,d,bs", [(4, 4, 128, 64), (8, 2, 128, 64), (6, 6, 64, 128), (4, 4, 128, 16)])
def test_attention_prefill(dtype, n_q_heads, n_kv, d, bs):
    specs = [(0, 128), (0, 17), (100, 64), (1000, 128), (4000, 96), (63, 1), (128, 128)]
    q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=3)
    items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
    out = torch.zeros_like(q)
    scale = 1.0 / math.sqrt(d)
    K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, scale)
    want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale)
    torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
    # split-KV variant (partials merged through scratch); twice to check the tickets were reset
    for n_splits in (3, 8):
        ws = torch.zeros(K().prefill_workspace_bytes(len(items), n_q_heads, d, n_splits), dtype=torch.uint8, device="cuda")
        for _ in range(2):
            out.zero_()
            K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, scale,
                                        n_splits=n_splits, workspace=ws)
            torch.testing.assert_close(out.float(), want.float(), **tol(dtype))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("n_q_heads,n_kv,bs", [(8, 1, 64), (32, 4, 64), (16, 4, 32), (64, 8, 64), (8, 2, 128)])
def test_attention_packed_gqa_decode(dtype, n_q_heads, n_kv, bs):
    """Head-packed grouped-query decode on the tensor cores (one CTA per (item, KV head), MMA rows = (token, head of the
    group)): decode items of 1 token — and as many tokens as 128 / G allows — against the fp32 paged-attention oracle,
    unsplit and split-KV (tickets must come back zeroed: every launch runs twice)."""
    d = 128
    G = n_q_heads // n_kv
    specs = [(0, 1), (100, 1), (1000, 1), (4000, 1), (63, 1), (127, 1), (128, 1), (5, min(3, 128 // G)), (700, min(128 // G, 6))]
    q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=5)
    items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
    scale = 1.0 / math.sqrt(d)
    want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale)
    out = torch.zeros_like(q)
    K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, scale, packed=True)
    torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
    for n_splits in (2, 5):
        ws = torch.zeros(K().prefill_workspace_bytes(len(items), n_kv, d, n_splits), dtype=torch.uint8, device="cuda")
        for _ in range(2):
            out.zero_()
            K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, scale,
                                        n_splits=n_splits, workspace=ws, packed=True)
            torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
        assert int(ws[:len(items) * n_kv * 32].view(torch.int32).abs().max()) == 0


@pytest.mark.parametrize("dtype", DTYPES)
def test_attention_alibi_all_kernels(dtype):
    """ALiBi (``hds_attention_next_alibi``): score += slope[head] * (key - query position) in the split-KV decode kernel
    (1 and several rows per item, split and unsplit), the tensor-core prefill kernel (unsplit / split-KV) and the
    head-packed GQA kernel, against the fp32 oracle; the setting is consumed by ONE launch (the next one is plain)."""
    scale_of = lambda d: 1.0 / math.sqrt(d)
    # split-KV decode, MHA (Bloom-like: 6 heads = interleaved slope sequence) and GQA rows
    for n_q_heads, n_kv, d, bs, rows in ((6, 6, 64, 128, 1), (8, 8, 128, 64, 1), (16, 4, 128, 64, 1), (4, 4, 128, 64, 8)):
        specs = [(0, rows), (5, rows), (127, 1), (1000, rows), (4095, 1), (333, 1)]
        q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=11)
        items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
        slopes = ref.alibi_slopes(n_q_heads).cuda()
        G = n_q_heads // n_kv
        R = rows * G
        want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale_of(d), alibi=slopes)
        plain = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale_of(d))
        assert (want.float() - plain.float()).abs().max() > 0.05          # the bias matters in this problem
        for n_splits in (1, 3):
            ws = torch.zeros(max(K().splitkv_workspace_bytes(len(items), n_q_heads, n_kv, d, n_splits, R), 16), dtype=torch.uint8,
                             device="cuda")
            out = torch.zeros_like(q)
            K().paged_attention_splitkv(q, out, cache, tables, items_t, len(items), R, n_q_heads, n_kv, d, bs, n_splits,
                                        scale_of(d), ws, alibi=slopes)
            torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
            out.zero_()
            K().paged_attention_splitkv(q, out, cache, tables, items_t, len(items), R, n_q_heads, n_kv, d, bs, n_splits,
                                        scale_of(d), ws)
            torch.testing.assert_close(out.float(), plain.float(), **tol(dtype))
    # tensor-core prefill
    # (32 heads: the steepest slope moves the row maximum by > 126 log2 units from one key tile to the next, so the rescale
    # factor of the accumulator underflows to exactly 0 — stale contributions must still be wiped)
    for n_q_heads, n_kv, d, bs in ((4, 4, 128, 64), (6, 6, 64, 128), (8, 2, 128, 16), (32, 8, 128, 64)):
        specs = [(0, 128), (0, 17), (100, 64), (1000, 128), (4000, 96), (63, 1)]
        q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=12)
        items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
        slopes = ref.alibi_slopes(n_q_heads).cuda()
        want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale_of(d), alibi=slopes)
        out = torch.zeros_like(q)
        K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, scale_of(d), alibi=slopes)
        torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
        ws = torch.zeros(K().prefill_workspace_bytes(len(items), n_q_heads, d, 3), dtype=torch.uint8, device="cuda")
        out.zero_()
        K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, scale_of(d), n_splits=3,
                                    workspace=ws, alibi=slopes)
        torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
    # head-packed grouped-query decode: every row of the tile has its own head's slope
    n_q_heads, n_kv, d, bs = 32, 4, 128, 64
    specs = [(0, 1), (100, 1), (1000, 1), (4000, 1), (5, 3), (700, 6)]
    q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=13)
    items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
    slopes = ref.alibi_slopes(n_q_heads).cuda()
    want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, scale_of(d), alibi=slopes)
    out = torch.zeros_like(q)
    K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, scale_of(d), packed=True,
                                alibi=slopes)
    torch.testing.assert_close(out.float(), want.float(), **tol(dtype))


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("window", [1, 100, 128, 1000])
def test_attention_sliding_window_all_kernels(dtype, window):
    """Sliding-window attention (``hds_attention_next_window``): keys (p - window, p] only, key tiles before the window
    skipped, in the split-KV decode kernel (split and unsplit, one and several rows), the tensor-core prefill kernel and the
    head-packed GQA kernel; combined with ALiBi once; consumed by ONE launch."""
    sc = lambda d: 1.0 / math.sqrt(d)
    for n_q_heads, n_kv, d, bs, rows in ((8, 8, 128, 64, 1), (12, 12, 64, 128, 1), (16, 4, 128, 64, 1), (4, 4, 128, 64, 8)):
        specs = [(0, rows), (5, rows), (127, 1), (1000, rows), (4095, 1), (333, 1)]
        q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=21)
        items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
        R = rows * (n_q_heads // n_kv)
        want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, sc(d), window=window)
        plain = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, sc(d))
        for n_splits in (1, 3):
            ws = torch.zeros(max(K().splitkv_workspace_bytes(len(items), n_q_heads, n_kv, d, n_splits, R), 16), dtype=torch.uint8,
                             device="cuda")
            out = torch.zeros_like(q)
            for _ in range(2):
                out.zero_()
                K().paged_attention_splitkv(q, out, cache, tables, items_t, len(items), R, n_q_heads, n_kv, d, bs, n_splits, sc(d),
                                            ws, window=window)
                torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
            out.zero_()
            K().paged_attention_splitkv(q, out, cache, tables, items_t, len(items), R, n_q_heads, n_kv, d, bs, n_splits, sc(d), ws)
            torch.testing.assert_close(out.float(), plain.float(), **tol(dtype))
    for n_q_heads, n_kv, d, bs in ((4, 4, 128, 64), (6, 6, 64, 128), (8, 2, 128, 16)):
        specs = [(0, 128), (0, 17), (100, 64), (1000, 128), (4000, 96), (63, 1), (300, 128)]
        q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=22)
        items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
        want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, sc(d), window=window)
        out = torch.zeros_like(q)
        K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, sc(d), window=window)
        torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
        ws = torch.zeros(K().prefill_workspace_bytes(len(items), n_q_heads, d, 3), dtype=torch.uint8, device="cuda")
        for _ in range(2):
            out.zero_()
            K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, sc(d), n_splits=3,
                                        workspace=ws, window=window)
            torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
        slopes = ref.alibi_slopes(n_q_heads).cuda()
        want_a = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, sc(d), window=window, alibi=slopes)
        out.zero_()
        K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, sc(d), window=window,
                                    alibi=slopes)
        torch.testing.assert_close(out.float(), want_a.float(), **tol(dtype))
    n_q_heads, n_kv, d, bs = 32, 4, 128, 64
    specs = [(0, 1), (100, 1), (1000, 1), (4000, 1), (5, 3), (700, 6)]
    q, cache, tables, items = build_paged_problem(specs, n_q_heads, n_kv, d, bs, dtype, seed=23)
    items_t = torch.tensor(items, dtype=torch.int32, device="cuda")
    want = ref.paged_attention(q, cache, tables, items_t.cpu(), n_q_heads, n_kv, sc(d), window=window)
    out = torch.zeros_like(q)
    K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, sc(d), packed=True, window=window)
    torch.testing.assert_close(out.float(), want.float(), **tol(dtype))
    ws = torch.zeros(K().prefill_workspace_bytes(len(items), n_kv, d, 4), dtype=torch.uint8, device="cuda")
    out.zero_()
    K().paged_attention_prefill(q, out, cache, tables, items_t, len(items), n_q_heads, n_kv, d, bs, sc(d), packed=True, n_splits=4,
                                workspace=ws, window=window)
    torch.testing.assert_close(out.float(), want.float(), **tol(dtype))


def test_attention_prefill_64_key_tiles_subprocess():
    """The opt-in 64-key-tile / two-CTAs-per-SM shape of the tensor-core attention kernel (HDS_PREFILL_TILE=64, read once per
    process): the prefill attention tests of this file are re-run in a child process with the variable set."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, HDS_PREFILL_TILE="64")
    res = subprocess.run([sys.executable, "-m", "pytest", __file__, "-q", "-m", "gpu", "-k", "test_attention_prefill and not subprocess",
                          "--timeout", "120", "-p", "no:cacheprovider"], env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-1000:]
    assert " passed" in res.stdout


# ------------------------------------------------------------------------------------------------
# mixture of experts
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,E,k,norm", [(1, 8, 2, True), (37, 8, 2, False), (300, 60, 4, True), (5, 3, 1, False)])
def test_moe_pipeline(T, E, k, norm):
    """top-k gating -> scatter -> grouped gated GEMM -> grouped GEMM -> weighted gather vs a per-token fp32 loop."""
    torch.manual_seed(T + E)
    H, I = 256, 384
    kern = K()
    x = rand(T, H)
    w_router = rand(E, H, scale=H ** -0.5)
    w_up = rand(E * 2 * I, H, scale=H ** -0.5)     # stacked, gate/up interleaved per expert
    w_down = rand(E * H, I, scale=I ** -0.5)
    logits = kern.gemm(x, w_router)
    scores = torch.empty(T, k, dtype=torch.float32, device="cuda")
    ids = torch.empty(T, k, dtype=torch.int32, device="cuda")
    offs = torch.empty(T, k, dtype=torch.int32, device="cuda")
    slots = torch.empty(T, k, dtype=torch.int32, device="cuda")
    counts = torch.zeros(E, dtype=torch.int32, device="cuda")
    pad_off = torch.zeros(E + 1, dtype=torch.int32, device="cuda")
    rows = kern.moe_padded_rows(T, E, k)
    xs = torch.zeros(rows, H, dtype=x.dtype, device="cuda")
    hs = torch.zeros(rows, I, dtype=x.dtype, device="cuda")
    ys = torch.zeros(rows, H, dtype=x.dtype, device="cuda")
    out = torch.zeros(T, H, dtype=torch.float32, device="cuda")
    for rep in range(2):  # the counters must come back zeroed
        kern.moe_topk_gating(logits, k, norm, scores, ids, offs, counts)
        kern.moe_scatter(x, xs, ids, offs, counts, pad_off, slots, k)
        kern.moe_gemm(xs, w_up, pad_off, E, hs, act=3, gated=True)
        kern.moe_gemm(hs, w_down, pad_off, E, ys)
        out.zero_()
        kern.moe_gather(ys, slots, scores, out)
    torch.cuda.synchronize()
    assert int(counts.abs().max()) == 0 and int(pad_off[-1]) % 128 == 0
    # oracle
    probs = torch.softmax(logits.float(), -1)
    topv, topi = torch.topk(probs, k, dim=-1)
    assert torch.equal(topi.int().sort(-1).values, ids.sort(-1).values)
    if norm:
        topv = topv / topv.sum(-1, keepdim=True)
    torch.testing.assert_close(scores.sort(-1).values, topv.sort(-1).values, rtol=1e-4, atol=1e-5)
    want = torch.zeros(T, H, device="cuda")
    for t in range(T):
        for j in range(k):
            e = int(ids[t, j])
            h = ref.gemm(x[t:t + 1], w_up[e * 2 * I:(e + 1) * 2 * I], act=3, gated=True, out_dtype=torch.float32).to(x.dtype)
            y = ref.gemm(h, w_down[e * H:(e + 1) * H], out_dtype=torch.float32).to(x.dtype)
            want[t] += float(scores[t, j]) * y[0].float()
    torch.testing.assert_close(out, want, rtol=3e-2, atol=3e-2)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("M,N,K_", [(130, 512, 64), (256, 256, 128), (300, 1000, 2096), (1024, 4096, 1024), (4100, 1376, 512)])
@pytest.mark.parametrize("bn", [0, 128, 256])
@pytest.mark.parametrize("mode", ["store", "gated"])
def test_gemm_2sm(dtype, M, N, K_, bn, mode):
    """CTA-pair kernel (tcgen05 cta_group::2, 256 x bn tiles): ragged M / N / K, both tile widths, both epilogues."""
    torch.manual_seed(M)
    a, w, b = rand(M, K_, dtype=dtype), rand(N, K_, dtype=dtype, scale=K_ ** -0.5), rand(N, dtype=dtype)
    gated = mode == "gated"
    out = K().gemm_2sm(a, w, bias=b, act=3 if gated else 2, gated=gated, bn=bn)
    want = ref.gemm(a, w, bias=b, act=3 if gated else 2, gated=gated, out_dtype=torch.float32)
    torch.testing.assert_close(out.float(), want, **tol(dtype))


def test_v2_kernel_wrapper_objects_cuda_vs_host():
    """inference.v2.kernels wrapper classes: the CUDA path (native kernels) against their own host formulation."""
    from hcache_deepspeed_b200.inference.v2.kernels import (BlasLibLinear, CUDAFPPreLN, CUDARMSNorm, MoEGather, MoEGEMM,
                                                            MoEScatter, RaggedLogitsGather, RaggedTopKGating)
    torch.manual_seed(5)
    T, H, I, E, k = 70, 256, 384, 8, 2
    x = rand(T, H)
    gate_w = rand(E, H, scale=H ** -0.5)
    w1, w2 = rand(E, 2 * I, H, scale=H ** -0.5), rand(E, H, I, scale=I ** -0.5)

    lg = torch.empty(T, E, dtype=x.dtype, device="cuda")
    BlasLibLinear(x.dtype)(lg, x, gate_w)
    torch.testing.assert_close(lg.float(), x.float() @ gate_w.float().t(), rtol=2e-2, atol=2e-2)
    # one set of fp32 logits for both devices, ties broken, so both route identically
    shared_logits = lg.float() + torch.arange(E, device="cuda") * 1e-4

    def run(dev):
        xd, a, b = x.to(dev), w1.to(dev), w2.to(dev)
        logits = shared_logits.to(dev)
        counts = torch.zeros(E, dtype=torch.int32, device=dev)
        scores = torch.empty(T, k, dtype=torch.float32, device=dev)
        ids, offs, slots = (torch.empty(T, k, dtype=torch.int32, device=dev) for _ in range(3))
        RaggedTopKGating(xd.dtype)(counts, scores, ids, offs, logits)
        rows = int(((counts + 127) // 128 * 128).sum())
        xs = torch.zeros(rows, H, dtype=xd.dtype, device=dev)
        cum = torch.zeros(E + 1, dtype=torch.int32, device=dev)
        MoEScatter(xd.dtype)(xs, cum, slots, xd, counts, ids, offs)
        hs = torch.zeros(rows, I, dtype=xd.dtype, device=dev)
        ys = torch.zeros(rows, H, dtype=xd.dtype, device=dev)
        MoEGEMM(xd.dtype, "silu", gated=True)(hs, xs, a, cum)
        MoEGEMM(xd.dtype)(ys, hs, b, cum)
        out = torch.zeros(T, H, dtype=torch.float32, device=dev)
        MoEGather(xd.dtype)(out, ys, scores, slots)
        gamma, beta = torch.ones(H, dtype=xd.dtype, device=dev), torch.zeros(H, dtype=xd.dtype, device=dev)
        n1 = torch.empty_like(xd)
        CUDARMSNorm(H, xd.dtype, 1e-6)(n1, xd, gamma)
        zr, zh = torch.empty_like(xd), torch.empty_like(xd)
        CUDAFPPreLN(H, xd.dtype, 1e-5)(zr, zh, xd, n1, gamma, beta)
        last = torch.empty(3, H, dtype=xd.dtype, device=dev)
        RaggedLogitsGather(H, xd.dtype)(last, zh, torch.tensor([0, 9, T - 1], dtype=torch.int32, device=dev))
        return out.cpu(), n1.float().cpu(), zh.float().cpu(), last.float().cpu()

    got, want = run("cuda"), run("cpu")
    torch.testing.assert_close(got[0], want[0], rtol=4e-2, atol=4e-2)
    for g_, w_ in zip(got[1:], want[1:]):
        torch.testing.assert_close(g_, w_, rtol=2e-2, atol=2e-2)
