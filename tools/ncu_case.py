"""Runs ONE kernel case a handful of times so that ncu can capture it (see B200_PROFILING.md).
usage: python tools/ncu_case.py <prefill|prefill_full|decode_gemm|restore|attn_decode|gemm>"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hcache_deepspeed_b200.ops import kernels as K  # noqa: E402
from hcache_deepspeed_b200.ops import reference as ref  # noqa: E402

dev = "cuda"
case = sys.argv[1]
iters = 4


def rand(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


if case == "prefill":
    d, bs, nq, nkv, ctx, n_new = 128, 64, 32, 32, 4096, 128
    nb = (ctx + n_new + bs - 1) // bs
    cache = (torch.randn(nb, 2, nkv, bs, d, device=dev) * 0.3).to(torch.bfloat16)
    tables = torch.arange(nb, device=dev, dtype=torch.int32).view(1, nb)
    items = torch.tensor([(0, 0, n_new, ctx)], dtype=torch.int32, device=dev)
    q = rand(n_new, nq, d)
    out = torch.empty_like(q)
    fn = lambda: K.paged_attention_prefill(q, out, cache, tables, items, 1, nq, nkv, d, bs, 1 / math.sqrt(d))
elif case == "prefill_full":
    d, bs, nq, nkv, n_new = 128, 64, 32, 32, 4096
    nb = (n_new + bs - 1) // bs
    cache = (torch.randn(nb, 2, nkv, bs, d, device=dev) * 0.3).to(torch.bfloat16)
    tables = torch.arange(nb, device=dev, dtype=torch.int32).view(1, nb)
    items = torch.tensor([(0, r0, 128, r0) for r0 in range(0, n_new, 128)], dtype=torch.int32, device=dev)
    q = rand(n_new, nq, d)
    out = torch.empty_like(q)
    fn = lambda: K.paged_attention_prefill(q, out, cache, tables, items, items.shape[0], nq, nkv, d, bs, 1 / math.sqrt(d))
elif case == "decode_gemm":
    M, N, Kd = 16, 12288, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fn = lambda: K.gemm(a, w, out=out, bn=128)
elif case == "decode_gated":
    M, N, Kd = 8, 22016, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    out = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)
    ws = torch.zeros(K.gemm_splitk_workspace_bytes(M, N), dtype=torch.uint8, device=dev)
    fn = lambda: K.gemm(a, w, act=3, gated=True, out=out, workspace=ws)
elif case == "gemm":
    M, N, Kd = 4096, 12288, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fn = lambda: K.gemm(a, w, out=out, bn=256)
elif case == "gemm_2sm":
    M, N, Kd = 4096, 12288, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fn = lambda: K.gemm_2sm(a, w, out=out, bn=256)
elif case == "gemm_2sm_silu":
    M, N, Kd = 4160, 22016, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fn = lambda: K.gemm_2sm(a, w, out=out, act=3, bn=256)
elif case == "gemm_fp8_2sm_gated":
    M, N, Kd = 4160, 22016, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    aq, a_s = K.quantize_fp8_rows(a)
    wq, w_s = K.quantize_fp8_rows(w)
    out = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)
    fn = lambda: K.gemm_fp8(aq, a_s, wq, w_s, out, act=3, gated=True, bn=1256)
elif case == "decode_gated_fp8":
    M, N, Kd = 8, 22016, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    aq, a_s = K.quantize_fp8_rows(a)
    wq, w_s = K.quantize_fp8_rows(w)
    out = torch.empty(M, N // 2, device=dev, dtype=torch.bfloat16)
    ws = torch.zeros(K.gemm_fp8_skinny_workspace_bytes(M, N), dtype=torch.uint8, device=dev)
    fn = lambda: K.gemm_fp8(aq, a_s, wq, w_s, out, act=3, gated=True, workspace=ws)
elif case == "gemm_fp8":
    M, N, Kd = 4096, 12288, 4096
    a, w = rand(M, Kd), rand(N, Kd, scale=Kd ** -0.5)
    aq, a_s = K.quantize_fp8_rows(a)
    wq, w_s = K.quantize_fp8_rows(w)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    fn = lambda: K.gemm_fp8(aq, a_s, wq, w_s, out, bn=256)
elif case == "restore":
    T, H, n_kv, d, bs = 4096, 4096, 32, 128, 64
    lat = rand(T, H)
    w_kv = rand(2 * n_kv * d, H, scale=H ** -0.5)
    cache = torch.zeros(T // bs + 8, 2, n_kv, bs, d, device=dev, dtype=torch.bfloat16)
    pos = torch.arange(T, device=dev, dtype=torch.int32)
    slot = torch.arange(T, device=dev, dtype=torch.int32)
    inv = ref.default_inv_freq(d, device=dev)
    fn = lambda: K.restore_kv_proj(lat, w_kv, cache, pos, slot, inv, n_kv, d, bs)
elif case == "attn_decode":
    d, bs, nq, nkv, n_seqs, ctx = 128, 64, 32, 32, 16, 4096
    bps = (ctx + 1 + bs - 1) // bs
    cache = (torch.randn(n_seqs * bps, 2, nkv, bs, d, device=dev) * 0.3).to(torch.bfloat16)
    tables = torch.arange(n_seqs * bps, device=dev, dtype=torch.int32).view(n_seqs, bps)
    items = torch.tensor([(s, s, 1, ctx) for s in range(n_seqs)], dtype=torch.int32, device=dev)
    q = rand(n_seqs, nq, d)
    out = torch.empty_like(q)
    ws = torch.zeros(max(16, K.splitkv_workspace_bytes(n_seqs, nq, nkv, d, 2, 1)), dtype=torch.uint8, device=dev)
    fn = lambda: K.paged_attention_splitkv(q, out, cache, tables, items, n_seqs, 1, nq, nkv, d, bs, 2, 1 / math.sqrt(d), ws)
else:
    raise SystemExit(f"unknown case {case}")

for _ in range(iters):
    fn()
torch.cuda.synchronize()
print("ok", case)
